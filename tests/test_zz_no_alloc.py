"""tdod_ransac_plane_static (include/tdod.h): the RANSAC plane fit with its scratch in a fixed device array inside the
library -- no workspace argument, no allocation.  Called through the C ABI and compared with tdod_ransac_plane on two
calls in a row (the second call finds the first call's contents in the shared array), for both point flavours; sizes
beyond the stated maximum must be refused.  Collected last on purpose (the file name sorts after the other tests)."""
import pytest
import torch

from tdod import api, synth, lib as L

pytestmark = pytest.mark.gpu


def _static_call(pts, n_pts, seed, iters, thresh, frame_id_base, out):
    """one call of the new entry point; every buffer it touches was allocated by the caller beforehand"""
    F, _, P = pts.shape
    is64 = pts.dtype == torch.float64
    return L.lib().tdod_ransac_plane_static(L.ptr(pts, None, "pts"), 1 if is64 else 0, L.ptr(n_pts, torch.int32, "n_pts"), F, P,
                                            None, int(seed), int(frame_id_base), int(iters), float(thresh),
                                            L.ptr(out["eq"]), L.ptr(out["best_iter"]), L.ptr(out["count"]), L.ptr(out["normal"]),
                                            L.ptr(out["mask"]), L.stream_ptr())


def _outputs(pts):
    F, _, P = pts.shape
    d = pts.device
    return dict(eq=torch.full((F, 4), float("nan"), dtype=pts.dtype, device=d),
                normal=torch.full((F, 3), float("nan"), dtype=pts.dtype, device=d),
                best_iter=torch.full((F,), -7, dtype=torch.int32, device=d),
                count=torch.full((F,), -7, dtype=torch.int32, device=d),
                mask=torch.full((F, P), 9, dtype=torch.uint8, device=d))


def _same(a, b):
    """bit-exact: the two entry points run the same kernels on the same inputs (integers and floating point alike)"""
    for k in ("best_iter", "count", "mask"):
        assert torch.equal(a[k], b[k]), k
    for k in ("eq", "normal"):
        ia = a[k].view(torch.int64 if a[k].dtype == torch.float64 else torch.int32)
        ib = b[k].view(torch.int64 if b[k].dtype == torch.float64 else torch.int32)
        assert torch.equal(ia, ib), k


@pytest.mark.parametrize("flavour", [L.BACKPROJECT_NUMPY, L.BACKPROJECT_TORCH])
def test_static_scratch_ransac_equals_workspace_ransac_twice_in_a_row(flavour):
    d = torch.device("cuda:0")
    c = synth.CONFIGS["C2"]
    iters = 1000
    sets = []
    for first, F in ((0, 12), (40, 7)):                       # two different batches: the second call must not see the first
        batch, _, _ = synth.make_batch(range(first, first + F), c["H"], c["W"], c["n_obj"], d)
        pts, _, counts = api.backproject(batch.depth, batch.ground, batch.K, flavour=flavour, want_other=False)
        sets.append((pts, counts[:, 0].contiguous(), first))
    refs = [api.ransac_plane(p, n, seed=11, iters=iters, thresh=0.05, want_mask=True, frame_id_base=fb) for p, n, fb in sets]
    outs = [_outputs(p) for p, _, _ in sets]
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated(d)
    for (p, n, fb), o in zip(sets, outs):                     # two calls in a row, same stream, nothing in between
        assert _static_call(p, n, 11, iters, 0.05, fb, o) == 0
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated(d) == before          # (torch's allocator saw nothing; the library calls no cudaMalloc)
    for r, o in zip(refs, outs):
        assert int(o["count"].min()) > 100                    # a real plane was found
        _same(r, o)
    # and once more on the first batch: the array now holds the second call's hypotheses and counts
    o = _outputs(sets[0][0])
    assert _static_call(sets[0][0], sets[0][1], 11, iters, 0.05, sets[0][2], o) == 0
    torch.cuda.synchronize()
    _same(refs[0], o)


def test_static_scratch_ransac_refuses_sizes_beyond_its_maximum():
    d = torch.device("cuda:0")
    F, P = 4, 64
    pts = torch.randn((F, 3, P), dtype=torch.float32, device=d)
    n = torch.full((F,), P, dtype=torch.int32, device=d)
    o = _outputs(pts)
    lib = L.lib()
    rc = _static_call(pts, n, 1, 65536 // F + 1, 0.05, 0, o)  # F * iters = 65 540 > 65 536
    assert rc == -2                                           # TDOD_ERANGE
    assert b"static scratch" in lib.tdod_last_error_string()
    torch.cuda.synchronize()
    assert int(o["count"][0]) == -7                           # nothing was launched
    assert _static_call(pts, n, 1, 65536 // F, 0.05, 0, o) == 0      # exactly the maximum is accepted
    torch.cuda.synchronize()
    r = api.ransac_plane(pts, n, seed=1, iters=65536 // F, thresh=0.05, want_mask=True)
    _same(r, o)
    big = torch.zeros((1025, 3, 4), dtype=torch.float32, device=d)
    nb = torch.full((1025,), 4, dtype=torch.int32, device=d)
    assert _static_call(big, nb, 1, 8, 0.05, 0, _outputs(big)) == -2   # F beyond 1024
