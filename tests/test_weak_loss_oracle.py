"""Oracle (oracle/weak_loss_oracle.py) against the fixture generated from the UNMODIFIED reference
(tests/golden/weak_loss.npz, `python tests/golden/make_golden.py weak_loss`): SURVEY 8(f) row 4."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "oracle"), os.path.join(ROOT, "3dod_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import weak_loss_oracle as WO  # noqa: E402
from tdod import synth  # noqa: E402

G = np.load(os.path.join(ROOT, "tests", "golden", "weak_loss.npz"), allow_pickle=True)


def frame():
    fid, H, W, n = [int(v) for v in G["frame"]]
    return synth.make_frame(fid, H, W, n, with_masks=True), H, W


def test_projection_chain_and_gradients():
    f, H, W = frame()
    cubes = torch.tensor(G["cubes"], requires_grad=True)
    Ks = torch.tensor(G["Ks"])
    N = cubes.shape[0]
    ims = [(W, H)] * N
    corners = WO.bube_corners(cubes, Ks, ims)
    boxes = WO.cubes_to_box(cubes, Ks, ims)
    np.testing.assert_allclose(corners.detach().numpy(), G["corners"], rtol=1e-6, atol=1e-5)
    np.testing.assert_allclose(boxes.detach().numpy(), G["boxes"], rtol=1e-6, atol=1e-5)
    g = torch.autograd.grad((boxes * torch.tensor(G["wb"])).sum(), cubes, retain_graph=True)[0]
    np.testing.assert_allclose(g.numpy(), G["grad_box"], rtol=1e-4, atol=1e-4)
    giou = WO.giou_loss(f.boxes, boxes)
    np.testing.assert_allclose(giou.detach().numpy(), G["giou"], rtol=1e-6, atol=1e-6)
    g = torch.autograd.grad((giou * torch.tensor(G["wl"])).sum(), cubes)[0]
    np.testing.assert_allclose(g.numpy(), G["grad_giou"], rtol=1e-4, atol=1e-6)


def test_z_loss_sweep():
    f, H, W = frame()
    cubes = torch.tensor(G["cubes"])
    N = cubes.shape[0]
    z = WO.z_loss(f.boxes, cubes, torch.tensor(G["Ks"]), [(W, H)] * N, torch.tensor(G["boxes"]))
    np.testing.assert_array_equal(z.numpy(), G["z_loss"])


def test_hull_and_segment_loss():
    f, H, W = frame()
    N = G["cubes"].shape[0]
    segc = torch.tensor(G["seg_corners"], requires_grad=True)
    for i in range(N):
        hull, order, plus = WO.jarvis_march(segc[i].detach())
        assert len(hull) == int(G["hull_len"][i])
        np.testing.assert_array_equal(hull.numpy(), G["hull"][i, :len(hull)])
    for kind in ("focal", "bce", "dice"):
        sl = WO.segment_loss(f.masks, segc, list(range(N)), loss=kind)
        np.testing.assert_allclose(sl.detach().numpy(), G["seg_" + kind], rtol=2e-6, atol=1e-7)
        w = torch.tensor(G["wl"]) if sl.dim() else torch.tensor(1.0)
        g = torch.autograd.grad((sl * w).sum(), segc)[0]
        ref = G["grad_seg_%s_corners" % kind]
        np.testing.assert_allclose(g.numpy(), ref, rtol=1e-4, atol=1e-6 * max(1.0, float(np.abs(ref).max())))


def test_segment_loss_gradient_reaches_the_cubes():
    f, H, W = frame()
    cubes = torch.tensor(G["cubes"], requires_grad=True)
    N = cubes.shape[0]
    ims = [(W, H)] * N
    segc = WO.segment_corners(WO.bube_corners(cubes, torch.tensor(G["Ks"]), ims), ims)
    np.testing.assert_allclose(segc.detach().numpy(), G["seg_corners"], rtol=1e-6, atol=1e-5)
    sl = WO.segment_loss(f.masks, segc, list(range(N)))
    g = torch.autograd.grad((sl * torch.tensor(G["wl"])).sum(), cubes)[0]
    ref = G["grad_seg_focal_cubes"]
    np.testing.assert_allclose(g.numpy(), ref, rtol=1e-3, atol=1e-5 * max(1.0, float(np.abs(ref).max())))


def _depth_case():
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    from depth_case import depth_loss_case
    return depth_loss_case(), np.load(os.path.join(ROOT, "tests", "golden", "depth_losses.npz"), allow_pickle=True)


def test_depth_map_losses_match_the_reference():
    """pseudo_gt_z_box_loss / pseudo_gt_z_point_loss / depth_range_loss (roi_heads.py:1196-1231, :1256-1305) against
    tests/golden/depth_losses.npz (`python tests/golden/make_golden.py depth_losses`)"""
    c, D = _depth_case()
    keep = torch.tensor(D["keep"])
    n1 = c["counts"][0]
    ok = [b.clone() for b in c["boxes"]]
    ok[1] = ok[1][keep[n1:n1 + c["counts"][1]]]
    z = WO.pseudo_gt_z_box_loss(list(c["depth"]), ok, c["pred_z"][keep])
    assert np.array_equal(z.numpy(), D["z_box"])
    # with the box whose slice is empty (the reference raises there) the restatement gives NaN for it alone
    zf = WO.pseudo_gt_z_box(list(c["depth"]), c["boxes"])
    assert int(torch.isnan(zf).sum()) == 1
    zp = WO.pseudo_gt_z_point_loss(list(c["depth"]), c["pred_xy"], c["pred_z"], c["counts"])
    assert np.array_equal(zp.numpy(), D["z_point"])
    cubes = c["cubes"].clone().requires_grad_(True)
    dr = WO.depth_range_loss(c["masks"], c["mask_idx"].tolist(), list(c["depth"]), cubes, c["gt_boxes"], c["counts"])
    np.testing.assert_allclose(dr.detach().numpy(), D["depth_range"], rtol=0, atol=2e-6)
    g = torch.autograd.grad(dr.sum(), cubes)[0]
    np.testing.assert_allclose(g.numpy(), D["grad_depth_range"], rtol=1e-6, atol=1e-6)
