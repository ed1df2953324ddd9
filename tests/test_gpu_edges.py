"""Edge cases through the C ABI: empty and ragged batches, minimum / large proposal counts, sizes the library must refuse
with a negative code (never a crash or a silent fallback)."""
import numpy as np
import pytest
import torch

import tdod_oracle as O
from tdod import api, pipeline, synth, lib as L

pytestmark = pytest.mark.gpu


def dev():
    return torch.device("cuda:0")


def test_empty_inputs_are_no_ops():
    d = dev()
    z = lambda *s, dt=torch.float32: torch.zeros(s, dtype=dt, device=d)
    assert api.score_iou(z(4), z(0, 4)).shape == (0,)
    assert api.cube_corners(z(0, 15)).shape == (0, 8, 3)
    r, h = api.mask_rect(torch.zeros((0, 64, 64), dtype=torch.uint8, device=d))
    assert r.shape == (0, 4, 2) and h.shape == (0,)
    o = api.propose_score_select(z(1, 48, 64), torch.eye(3, device=d)[None], z(0, 4), z(0, 3), z(0, 3),
                                 torch.zeros(0, dtype=torch.int32, device=d), z(0, 4, 2), None, 16, seed=1)
    assert o.best_idx.shape == (0,)
    pts, _, cnt = api.backproject(z(0, 48, 64), None, z(0, 3, 3))
    assert pts.shape[0] == 0 and cnt.shape == (0, 2)
    torch.cuda.synchronize()


def test_ragged_batch_with_object_free_frames():
    """frames without objects in the middle of a batch: winners equal the per-frame runs (global Philox ids)"""
    d = dev()
    counts = [3, 0, 1, 0, 0, 7, 2]
    frames = [synth.make_frame(60 + k, 240, 320, c) for k, c in enumerate(counts)]
    frames = [f for f in frames]
    keep = [f for f in frames if f.n_obj() > 0]
    # from_frames needs masks on every frame that has objects; object-free frames carry masks=None
    batch = pipeline.FrameBatch.from_frames(frames, d)
    assert batch.n_obj == sum(counts) and batch.n_frames == len(counts)
    s = pipeline.select(batch, 300, seed=3)
    torch.cuda.synchronize()
    off = np.cumsum([0] + counts)
    for k, f in enumerate(frames):
        if counts[k] == 0:
            continue
        b1 = pipeline.FrameBatch.from_frames([f], d)
        s1 = pipeline.select(b1, 300, seed=3, obj_id_base=int(off[k]), frame_id_base=k)
        torch.cuda.synchronize()
        assert torch.equal(s1.best_idx, s.best_idx[off[k]:off[k + 1]])
        assert torch.equal(s1.normal[0], s.normal[k])
    assert len(keep) == 4


@pytest.mark.parametrize("n", [2, 3, 33, 2049, 16000])
def test_proposal_count_extremes_against_the_oracle_sampler(n):
    """n = 2 is the smallest the reference's linspace supports; 2049 crosses from one CTA to a cluster; 16 000 uses all
    eight CTAs.  Host-drawn randomness -> dimensions and rotations bit-exact, centres to the last-ulp trig tolerance."""
    d = dev()
    f = synth.make_frame(77, 240, 320, 2)
    g = torch.Generator().manual_seed(n)
    nv = torch.tensor([0.02, 0.9995, 0.02], dtype=torch.float64)
    nv = nv / nv.norm()
    o = O.propose(f.boxes, f.depth, f.prior_mean, f.prior_std, f.K, n, nv, g)
    got = api.propose_score_select(f.depth[None].to(d), f.K[None].to(d), f.boxes.to(d), f.prior_mean.to(d), f.prior_std.to(d),
                                   torch.zeros(2, dtype=torch.int32, device=d), None, None, n,
                                   yaw_tab=api.yaw_table(nv[None].to(d)), noise=o["noise"].permute(1, 0, 2).contiguous().to(d),
                                   yaw_idx=o["yaw_idx"].to(torch.int16).to(d), propose_only=True)
    torch.cuda.synchronize()
    a, b = got.all_cubes.cpu().numpy(), o["cubes"].numpy()
    assert np.array_equal(a[..., 3:], b[..., 3:])
    ok = np.isclose(a[..., :3], b[..., :3], rtol=1e-5, atol=1e-6) | (np.isnan(a[..., :3]) & np.isnan(b[..., :3]))
    assert ok.all()


def test_sizes_outside_the_supported_range_are_refused():
    d = dev()
    f = synth.make_frame(5, 64, 96, 1)
    args = (f.depth[None].to(d), f.K[None].to(d), f.boxes.to(d), f.prior_mean.to(d), f.prior_std.to(d),
            torch.zeros(1, dtype=torch.int32, device=d), torch.zeros((1, 4, 2), device=d), None)
    with pytest.raises(ValueError, match="n_prop"):
        api.propose_score_select(*args, 1, seed=0)                       # linspace needs n >= 2
    with pytest.raises(ValueError, match="shared memory"):
        api.propose_score_select(*args, 100000, seed=0)                  # 7 arrays x n/8 floats do not fit
    with pytest.raises(ValueError):
        api.mask_rect(torch.zeros((1, 8, 8), dtype=torch.uint8, device=d))   # W < 16
    with pytest.raises(ValueError, match="CUDA"):
        api.score_iou(torch.zeros(4), torch.zeros((3, 4)))               # CPU tensors: no fallback
    with pytest.raises(ValueError):
        api.propose_score_select(*args, 16, seed=0, cluster_size=3)
