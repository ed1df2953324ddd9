"""CPU: the C-ABI library builds, loads, and exports exactly what include/tdod.h declares; the host shim refuses
non-CUDA inputs instead of falling back; nothing in the product imports the oracle."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "tdod.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(tdod_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from tdod import lib
    h = lib.lib()
    names = _declared()
    assert len(names) >= 17
    for n in names:
        assert hasattr(h, n), n
    assert sorted(lib.SIGNATURES) == names, "ctypes table and header disagree"
    assert h.tdod_version() >= 100
    assert h.tdod_strerror(0) == b"ok"


def test_struct_mirror_matches_header():
    from tdod import lib
    src = open(os.path.join(ROOT, "include", "tdod.h")).read()
    body = src[src.index("typedef struct tdod_pss_args {") + len("typedef struct tdod_pss_args {"):src.index("} tdod_pss_args;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    fields = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        for part in decl.split(","):
            fields.append(re.findall(r"([A-Za-z_][A-Za-z0-9_]*)\s*$", part.strip())[0])
    assert fields == [f[0] for f in lib.PssArgs._fields_]


def test_argument_checks_return_negative_codes_without_a_gpu():
    from tdod import lib
    h = lib.lib()
    a = lib.PssArgs()
    a.n_obj, a.n_prop = 1, 1
    rc = h.tdod_propose_score_select(ctypes.byref(a), None)
    assert rc == -1 and b"n_prop" in h.tdod_last_error_string()
    a.n_prop = 100
    rc = h.tdod_propose_score_select(ctypes.byref(a), None)
    assert rc == -1 and b"NULL" in h.tdod_last_error_string()
    assert h.tdod_mask_rect(None, 1, 10, 10, None, None, None, 0, None) == -1
    assert h.tdod_backproject(None, None, None, 1, 4, 4, 5, 0, None, None, None, None) == -1
    assert h.tdod_ransac_workspace_bytes(2, 1000, 1) > h.tdod_ransac_workspace_bytes(2, 1000, 0) > 0


def test_no_cpu_fallback():
    from tdod import api
    import ProposalNetwork.utils.spaces as spaces
    import ProposalNetwork.utils.conversions as conv
    cubes = spaces.Cubes(torch.zeros(1, 4, 15))
    with pytest.raises(ValueError):
        cubes.get_all_corners()
    with pytest.raises(ValueError):
        conv.cubes_to_box(cubes, torch.eye(3), (64, 48))
    with pytest.raises(NameError):
        cubes.get_bube_corners(torch.eye(3))
    with pytest.raises(ValueError):
        api.score_iou(torch.zeros(4), torch.zeros(3, 4))


def test_missing_library_fails_loudly(monkeypatch):
    from tdod import lib
    monkeypatch.setattr(lib, "_lib", None)
    monkeypatch.setattr(lib, "LIB_PATH", "/nonexistent/libtdod.so")
    with pytest.raises(RuntimeError, match="no CPU path"):
        lib.lib()


def test_product_never_imports_the_oracle():
    """the product path may not import the oracle, the reference harness, or the CPU libraries the reference's
    own implementation leans on (cv2, scipy)"""
    pkg = os.path.join(ROOT, "3dod_b200")
    banned = re.compile(r"^\s*(import|from)\s+(tdod_oracle|weak_loss_oracle|cv_spec|ref_import|oracle|cv2|scipy)\b", re.M)
    n = 0
    for top in (pkg, os.path.join(ROOT, "tools")):          # the measurement tools run the product only, too
        for dp, _, fns in os.walk(top):
            for fn in fns:
                if fn.endswith(".py"):
                    n += 1
                    assert not banned.search(open(os.path.join(dp, fn)).read()), os.path.join(dp, fn)
    assert n >= 10


def test_cubes_container_semantics():
    import ProposalNetwork.utils.spaces as spaces
    t = torch.arange(2 * 3 * 15, dtype=torch.float32).reshape(2, 3, 15)
    c = spaces.Cubes(t)
    assert c.num_instances == 2 and len(c) == 2 and tuple(c.shape) == (2, 3, 15)
    assert tuple(c[1].tensor.shape) == (1, 3, 15) and torch.equal(c[1].tensor[0], t[1])
    assert tuple(c[1, 2].tensor.shape) == (1, 1, 15) and torch.equal(c[1, 2].tensor[0, 0], t[1, 2])
    assert tuple(c.dimensions.shape) == (2, 3, 3) and tuple(c.rotations.shape) == (2, 3, 3, 3)
    assert tuple(spaces.Cubes.cat([c, c]).tensor.shape) == (4, 3, 15)
    assert [tuple(x.tensor.shape) for x in c.split(2, dim=1)] == [(2, 2, 15), (2, 1, 15)]
    assert tuple(spaces.Cubes(torch.zeros(15)).tensor.shape) == (1, 1, 15)
    with pytest.raises(AssertionError):
        spaces.Cubes(t, scores=torch.zeros(3))


def test_host_mask_packer_matches_the_bitmap_layout():
    """tdod_host_pack_masks is a HOST function (the transfer codec of the end-to-end path): runs without a GPU; its
    output must be the layout tdod_mask_rect_bitmap reads, i.e. what pipeline.pack_masks builds with torch ops"""
    import numpy as np
    import torch
    from tdod import lib as L, pipeline
    rs = np.random.RandomState(2)
    for (N, H, W) in ((3, 17, 33), (2, 40, 97), (4, 9, 64), (1, 53, 730), (2, 5, 16)):
        m = torch.from_numpy(((rs.rand(N, H, W) < 0.4) * rs.randint(1, 255, size=(N, H, W))).astype(np.uint8))
        m[0, :, W - 1] = 7
        want = pipeline.pack_masks(m)
        for nt in (1, 2, 5):
            got = torch.full(want.shape, -1, dtype=torch.int32)
            assert L.lib().tdod_host_pack_masks(m.data_ptr(), N, H, W, got.data_ptr(), nt) == 0
            assert torch.equal(got, want), (N, H, W, nt)
        b = torch.from_numpy(rs.rand(N, H, W) < 0.5)
        got = torch.empty(want.shape, dtype=torch.int32)
        assert L.lib().tdod_host_pack_masks(b.data_ptr(), N, H, W, got.data_ptr(), 3) == 0
        assert torch.equal(got, pipeline.pack_masks(b))
    assert L.lib().tdod_host_pack_masks(None, 1, 4, 4, None, 1) < 0
