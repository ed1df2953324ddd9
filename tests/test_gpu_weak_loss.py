"""GPU parity of the weak-loss geometry: tdod.weak_loss through the C ABI against
tests/golden/weak_loss.npz (values and autograd gradients of the UNMODIFIED reference) and against the oracle
(oracle/weak_loss_oracle.py, plain torch autograd on the CPU) on larger seeded inputs.

Bars: hull vertices, argmin / argmax choices and z_loss values bit-exact; fp32 values within 1e-5 relative; gradients
within 1e-4 relative to the largest component of the tensor (the reference's own fp32 autograd carries that much
rounding noise: it sums H x W products per edge in fp32), and within 1e-6 of the oracle evaluated in fp64."""
import os

import numpy as np
import pytest
import torch

import tdod_oracle as O
import weak_loss_oracle as WO
from tdod import synth, weak_loss as WL

pytestmark = pytest.mark.gpu


def dev():
    return torch.device("cuda:0")


def close(got, ref, rtol, atol_rel=0.0, atol=0.0, what=""):
    got = np.asarray(got, np.float64); ref = np.asarray(ref, np.float64)
    assert got.shape == ref.shape, what
    tol = atol + atol_rel * (np.abs(ref).max() if ref.size else 0.0) + rtol * np.abs(ref)
    bad = ~(np.abs(got - ref) <= tol)
    assert not bad.any(), "%s: %d / %d outside tolerance, worst abs %.3g (ref scale %.3g)" % (
        what, int(bad.sum()), bad.size, float(np.abs(got - ref).max()), float(np.abs(ref).max()))


@pytest.fixture(scope="module")
def G(golden_dir):
    return np.load(os.path.join(golden_dir, "weak_loss.npz"), allow_pickle=True)


def fixture_frame(G):
    fid, H, W, n = [int(v) for v in G["frame"]]
    return synth.make_frame(fid, H, W, n, with_masks=True), H, W


def test_projection_chain_matches_reference_values_and_gradients(G):
    f, H, W = fixture_frame(G)
    d = dev()
    cubes = torch.tensor(G["cubes"], device=d, requires_grad=True)
    N = cubes.shape[0]
    out = WL.project_boxes(cubes, torch.tensor(G["Ks"], device=d), [(W, H)] * N, gt_boxes=f.boxes.to(d))
    close(out["corners"].detach().cpu(), G["corners"], 1e-5, atol=1e-4, what="corners")
    close(out["seg_corners"].detach().cpu(), G["seg_corners"], 1e-5, atol=1e-4, what="seg corners")
    close(out["boxes"].detach().cpu(), G["boxes"], 1e-5, atol=1e-4, what="boxes")
    close(out["loss_iou"].detach().cpu(), G["giou"], 1e-5, atol=1e-6, what="giou")
    g = torch.autograd.grad((out["boxes"] * torch.tensor(G["wb"], device=d)).sum(), cubes, retain_graph=True)[0]
    close(g.cpu(), G["grad_box"], 1e-4, atol_rel=1e-5, what="d box / d cubes")
    g = torch.autograd.grad((out["loss_iou"] * torch.tensor(G["wl"], device=d)).sum(), cubes)[0]
    close(g.cpu(), G["grad_giou"], 1e-4, atol_rel=1e-5, what="d giou / d cubes")


def test_z_loss_matches_reference(G):
    f, H, W = fixture_frame(G)
    d = dev()
    N = G["cubes"].shape[0]
    z = WL.z_loss(f.boxes.to(d), torch.tensor(G["cubes"], device=d), torch.tensor(G["Ks"], device=d), [(W, H)] * N,
                  torch.tensor(G["boxes"], device=d))
    close(z.cpu(), G["z_loss"], 0, atol=2e-7, what="z_loss")


def test_hull_matches_reference_bit_exact(G):
    d = dev()
    h, meta = WL.hull(torch.tensor(G["seg_corners"], device=d))
    h, meta = h.cpu().numpy(), meta.cpu().numpy()
    np.testing.assert_array_equal(meta[:, 0], G["hull_len"])
    for i in range(h.shape[0]):
        m = int(meta[i, 0])
        np.testing.assert_array_equal(h[i, :m], G["hull"][i, :m])


@pytest.mark.parametrize("kind", ["focal", "bce", "dice"])
def test_segment_loss_matches_reference(G, kind):
    f, H, W = fixture_frame(G)
    d = dev()
    N = G["cubes"].shape[0]
    segc = torch.tensor(G["seg_corners"], device=d, requires_grad=True)
    sl = WL.segment_loss(f.masks.to(d), segc, list(range(N)), loss=kind)
    close(sl.detach().cpu(), G["seg_" + kind], 1e-5, atol=1e-7, what="segment_loss " + kind)
    w = torch.tensor(G["wl"], device=d) if sl.dim() else torch.tensor(1.0, device=d)
    g = torch.autograd.grad((sl * w).sum(), segc)[0]
    close(g.cpu(), G["grad_seg_%s_corners" % kind], 1e-4, atol_rel=1e-4, what="d segment_loss / d corners " + kind)


def test_full_chain_gradient_reaches_the_cubes(G):
    f, H, W = fixture_frame(G)
    d = dev()
    cubes = torch.tensor(G["cubes"], device=d, requires_grad=True)
    N = cubes.shape[0]
    out = WL.project_boxes(cubes, torch.tensor(G["Ks"], device=d), [(W, H)] * N)
    sl = WL.segment_loss([m[None] for m in f.masks.to(d)], out["seg_corners"], list(range(N)))     # the reference's list form
    g = torch.autograd.grad((sl * torch.tensor(G["wl"], device=d)).sum(), cubes)[0]
    close(g.cpu(), G["grad_seg_focal_cubes"], 1e-3, atol_rel=1e-4, what="d segment_loss / d cubes")


def _random_case(n, H, W, seed):
    f = synth.make_frame(seed, H, W, n, with_masks=True)
    N = f.boxes.shape[0]
    g = torch.Generator().manual_seed(seed)
    cubes = O.propose(f.boxes, f.depth, f.prior_mean, f.prior_std, f.K, 3, normal=torch.tensor([0.0, 1.0, 0.05]),
                      generator=g)["cubes"][:, 0].float().clone()
    cubes[::7, 2] *= 0.3                       # some cubes close to / across the camera plane: clamped corners
    cubes[3::11, 0] += 2.5
    Ks = f.K[None].repeat(N, 1, 1).clone()
    Ks[:, 1, 2] += torch.arange(N) * 0.25
    return f, cubes, Ks


def test_projection_chain_against_oracle_autograd_on_more_cubes():
    d = dev()
    H, W = 530, 730
    tot = 0
    for seed in (3, 4, 5):
        f, cubes, Ks = _random_case(40, H, W, seed)
        N = cubes.shape[0]
        tot += N
        ims = [(W, H)] * N
        g = torch.Generator().manual_seed(seed)
        wb, wl, wc = torch.rand((N, 4), generator=g), torch.rand(N, generator=g), torch.rand((N, 8, 2), generator=g)
        ref_c = cubes.double().requires_grad_(True)
        rc = WO.bube_corners(ref_c, Ks.double(), ims)
        rs = WO.segment_corners(rc, ims)
        rb = WO.cubes_to_box(ref_c, Ks.double(), ims)
        rg = WO.giou_loss(f.boxes.double(), rb)
        ref_grad = torch.autograd.grad((rb * wb).sum() + (rg * wl).sum() + (rc * wc).sum() + (rs * wc).sum() * 0.5, ref_c)[0]
        c = cubes.to(d).requires_grad_(True)
        out = WL.project_boxes(c, Ks.to(d), ims, gt_boxes=f.boxes.to(d))
        close(out["boxes"].detach().cpu(), rb.detach(), 1e-5, atol=2e-4, what="boxes")
        close(out["loss_iou"].detach().cpu(), rg.detach(), 1e-5, atol=2e-6, what="giou")
        got = torch.autograd.grad((out["boxes"] * wb.to(d)).sum() + (out["loss_iou"] * wl.to(d)).sum() +
                                  (out["corners"] * wc.to(d)).sum() + (out["seg_corners"] * wc.to(d)).sum() * 0.5, c)[0]
        # fp32 kernel against fp64 autograd, per-cube scale.  Cubes with a corner near the camera plane are
        # ill-conditioned in fp32 (pz -> 0): there the bar is the error of the reference's own fp32 autograd
        r32 = cubes.clone().requires_grad_(True)
        c32 = WO.bube_corners(r32, Ks, ims)
        b32 = WO.cubes_to_box(r32, Ks, ims)
        g32 = torch.autograd.grad((b32 * wb).sum() + (WO.giou_loss(f.boxes, b32) * wl).sum() + (c32 * wc).sum() +
                                  (WO.segment_corners(c32, ims) * wc).sum() * 0.5, r32)[0].double().numpy()
        ref = ref_grad.numpy(); gg = got.cpu().numpy().astype(np.float64)
        scale = np.abs(ref).max(1) + 1e-6
        err = np.abs(gg - ref).max(1) / scale
        err32 = np.abs(g32 - ref).max(1) / scale
        assert (err <= 2e-4 + 4 * err32).all(), (float(err.max()), float(err32.max()))
        assert np.median(err) < 1e-5
        z = WL.z_loss(f.boxes.to(d), cubes.to(d), Ks.to(d), ims, out["boxes"].detach())
        zr = WO.z_loss(f.boxes, cubes, Ks, ims, out["boxes"].detach().cpu())
        # an argmin over 50 areas computed in fp32 on both sides: a near-tie may pick the neighbouring 0.1 m step
        diff = np.abs(z.cpu().numpy() - zr.numpy())
        assert (diff < 1e-6).mean() > 0.95 and diff.max() <= 0.05 + 1e-6, (float((diff < 1e-6).mean()), float(diff.max()))
    assert tot > 60


@pytest.mark.parametrize("kind", ["focal", "dice"])
def test_segment_loss_against_fp64_oracle_at_c2_resolution(kind):
    d = dev()
    H, W = 530, 730
    f, cubes, Ks = _random_case(6, H, W, 9)
    N = cubes.shape[0]
    ims = [(W, H)] * N
    segc = WO.segment_corners(WO.bube_corners(cubes, Ks, ims), ims)
    # the hull and the soft mask are fp32 objects (separately rounded products) and the oracle reproduces them in fp32;
    # its pixel sums are fp32 too (torch autograd), the kernel's are fp64: hence the 1e-3 bar on the gradients
    ref_in = segc.clone().requires_grad_(True)
    ref = WO.segment_loss(f.masks, ref_in, list(range(N)), loss=kind)
    w = torch.rand(N, generator=torch.Generator().manual_seed(1)) if ref.dim() else torch.tensor(1.0)
    ref_g = torch.autograd.grad((ref * w).sum(), ref_in)[0]
    x = segc.to(d).requires_grad_(True)
    got = WL.segment_loss(f.masks.to(d), x, list(range(N)), loss=kind)
    close(got.detach().cpu(), ref.detach(), 2e-5, atol=1e-7, what=kind)
    g = torch.autograd.grad((got * w.to(d)).sum(), x)[0]
    close(g.cpu(), ref_g, 1e-3, atol_rel=1e-3, what="grad " + kind)


def test_hull_edge_cases_duplicates_and_collinear_points():
    d = dev()
    sq = [[0., 0.], [10., 0.], [10., 10.], [0., 10.]]
    cases = [
        sq + sq,                                                                   # every point twice: shifted copies
        [[5., 5.]] * 8,                                                            # all identical
        [[0., 0.], [5., 0.], [10., 0.], [10., 5.], [10., 10.], [5., 10.], [0., 10.], [0., 5.]],   # collinear mid points
        [[0., 0.], [0., 0.], [0., 0.], [729., 0.], [729., 529.], [0., 529.], [300., 200.], [300., 200.]],
        [[3., 1.], [3., 9.], [3., 5.], [8., 5.], [3., 1.], [6., 2.], [7., 8.], [4., 4.]],
    ]
    pts = torch.tensor(cases)
    h, meta = WL.hull(pts.to(d))
    h, meta = h.cpu(), meta.cpu()
    for i in range(len(cases)):
        rh, order, plus = WO.jarvis_march(pts[i])
        m = int(meta[i, 0])
        assert m == len(rh), (i, m, len(rh))
        np.testing.assert_array_equal(h[i, :m].numpy(), rh.numpy())
        np.testing.assert_array_equal(meta[i, 1:1 + m].numpy(), np.array(order))
    # integer vertices on the image border: edge functions hit 0 and 1 exactly -> halved gradients (torch.max / min ties)
    H, W = 40, 60
    masks = (torch.rand((len(cases), H, W), generator=torch.Generator().manual_seed(2)) > 0.5)
    small = (pts * torch.tensor([W - 1, H - 1]) / torch.tensor([729., 529.])).round().clamp(min=0)
    small = torch.minimum(small, torch.tensor([W - 1., H - 1.]))
    ref_in = small.clone().requires_grad_(True)
    ref = WO.segment_loss(masks, ref_in, list(range(len(cases))))
    ref_g = torch.autograd.grad(ref.sum(), ref_in)[0]
    x = small.to(d).requires_grad_(True)
    got = WL.segment_loss(masks.to(d), x, list(range(len(cases))))
    close(got.detach().cpu(), ref.detach(), 1e-5, atol=1e-7, what="border focal")
    g = torch.autograd.grad(got.sum(), x)[0]
    close(g.cpu(), ref_g, 1e-4, atol_rel=1e-4, what="border focal grad")


def test_normal_vector_from_maps_is_the_fp32_plane_path():
    d = dev()
    fr = [synth.make_frame(i, 240, 320, 4, device=d, with_masks=False) for i in range(3)]
    depth = torch.stack([f.depth for f in fr]); ground = torch.stack([f.ground for f in fr]); K = torch.stack([f.K for f in fr])
    nv = WL.normal_vector_from_maps(ground, depth, K, seed=3)
    assert nv.shape == (3, 3) and nv.dtype == torch.float32
    from tdod import pipeline, lib as L
    ref = pipeline.ground_normals(depth, ground, K, seed=3, flavour=L.BACKPROJECT_TORCH)[0]
    assert torch.equal(nv, ref)                       # the fp32 K-A path (parity of that path: test_gpu_parity.py)
    for i in range(3):
        n = nv[i].cpu().double()
        assert abs(float(n.norm()) - 1.0) < 1e-5 and float(n[1]) > 0
    with pytest.raises(ValueError):
        WL.project_boxes(torch.zeros((2, 15)), torch.eye(3), [(10, 10)] * 2)          # CPU tensors are refused


def test_dense_convex_hull_and_fill_polygon_mirror():
    """ProposalNetwork.utils.utils.convex_hull / fill_polygon / jarvis_march (utils.py:424-507) as dense images"""
    from ProposalNetwork.utils import utils as U
    d = dev()
    H, W = 90, 120
    g = torch.Generator().manual_seed(4)
    pts = torch.rand((5, 8, 2), generator=g) * torch.tensor([W - 1., H - 1.])
    pts[1, 3] = pts[1, 0]                                  # a duplicate
    pts[2] = pts[2].round()                                # integer vertices: exact 0 / 1 edge values
    wgt = torch.rand((5, H, W), generator=g)
    mask = torch.zeros((H, W), dtype=torch.bool)
    ref_in = pts.clone().requires_grad_(True)
    ref = torch.stack([WO.convex_hull((H, W), ref_in[i]) for i in range(5)])
    ref_g = torch.autograd.grad((ref * wgt).sum(), ref_in)[0]
    x = pts.to(d).requires_grad_(True)
    got = U.convex_hull(mask.to(d), x)
    assert got.shape == (5, H, W)
    np.testing.assert_array_equal(got.detach().cpu().numpy(), ref.detach().numpy())       # same fp32 products, same order
    gg = torch.autograd.grad((got * wgt.to(d)).sum(), x)[0]
    close(gg.cpu(), ref_g, 1e-4, atol_rel=1e-4, what="dense hull grad")
    one = U.convex_hull(mask.to(d), pts[0].to(d))
    np.testing.assert_array_equal(one.cpu().numpy(), ref[0].detach().numpy())
    hull0 = U.jarvis_march(pts[0].to(d))
    np.testing.assert_array_equal(hull0.cpu().numpy(), WO.jarvis_march(pts[0])[0].numpy())
    tri = torch.tensor([[10., 10.], [100., 20.], [50., 80.]])
    ref_t = tri.clone().requires_grad_(True)
    rf = WO.fill_polygon((H, W), ref_t)
    rg = torch.autograd.grad((rf * wgt[0]).sum(), ref_t)[0]
    xt = tri.to(d).requires_grad_(True)
    gf = U.fill_polygon(mask.to(d), xt)
    np.testing.assert_array_equal(gf.detach().cpu().numpy(), rf.detach().numpy())
    close(torch.autograd.grad((gf * wgt[0].to(d)).sum(), xt)[0].cpu(), rg, 1e-4, atol_rel=1e-4, what="fill_polygon grad")


# ---- the losses that read the depth map (roi_heads.py:1196-1231, :1256-1305) ------------------------------------------
def _depth_case(golden_dir):
    import sys
    if golden_dir not in sys.path:
        sys.path.insert(0, golden_dir)
    from depth_case import depth_loss_case
    return depth_loss_case(), np.load(os.path.join(golden_dir, "depth_losses.npz"), allow_pickle=True)


def test_depth_map_losses_match_reference(golden_dir):
    c, D = _depth_case(golden_dir)
    d = dev()
    depth = c["depth"].to(d)
    # all boxes, the one with an empty slice included: NaN for it, the reference's values for the others
    z = WL.pseudo_gt_z_box_loss(depth, [b.to(d) for b in c["boxes"]], c["pred_z"].to(d)).cpu()
    zo = WO.pseudo_gt_z_box_loss(list(c["depth"]), c["boxes"], c["pred_z"])
    assert int(torch.isnan(z).sum()) == 1
    assert np.array_equal(z.numpy(), zo.numpy(), equal_nan=True)
    # the reference's own run (without that box): bit-equal
    keep = torch.tensor(D["keep"])
    n1 = c["counts"][0]
    ok = [b.clone() for b in c["boxes"]]
    ok[1] = ok[1][keep[n1:n1 + c["counts"][1]]]
    pz = c["pred_z"][keep].to(d).requires_grad_(True)
    z = WL.pseudo_gt_z_box_loss(list(depth), [b.to(d) for b in ok], pz)
    assert np.array_equal(z.detach().cpu().numpy(), D["z_box"])
    z.sum().backward()
    assert torch.all(pz.grad.abs() == 1)
    zp = WL.pseudo_gt_z_point_loss(depth, c["pred_xy"].to(d), c["pred_z"].to(d), c["counts"])
    assert np.array_equal(zp.cpu().numpy(), D["z_point"])
    cubes = c["cubes"].to(d).requires_grad_(True)
    dr = WL.depth_range_loss(c["masks"].to(d), c["mask_idx"].tolist(), depth, cubes, c["gt_boxes"].to(d), c["counts"])
    close(dr.detach().cpu().numpy(), D["depth_range"], rtol=0, atol=2e-6, what="depth_range_loss")
    dr.sum().backward()
    close(cubes.grad.cpu().numpy(), D["grad_depth_range"], rtol=1e-6, atol=1e-6, what="depth_range_loss gradient")


def test_depth_select_matches_torch_on_large_crops():
    """medians bit-equal to torch.median and quantile ranges within an ulp of torch.quantile on the device, at the frame
    size of the bench (730 x 530), duplicate-heavy depths and negative values included"""
    d = dev()
    g = torch.Generator().manual_seed(3)
    F_, H, W = 3, 530, 730
    depth = torch.rand((F_, H, W), generator=g) * 8 - 1
    depth[1] = torch.round(depth[1] * 4) / 4                              # 36 distinct values: long runs of ties
    depth = depth.to(d)
    n = 40
    x1 = torch.rand(n, generator=g) * W * 0.8; y1 = torch.rand(n, generator=g) * H * 0.8
    boxes = torch.stack((x1, y1, x1 + torch.rand(n, generator=g) * W * 0.6 + 2, y1 + torch.rand(n, generator=g) * H * 0.6 + 2), 1)
    boxes[0] = torch.tensor([-5.0, -5.0, W + 5.0, H + 5.0])                # the whole frame
    boxes[1] = torch.tensor([7.2, 9.9, 8.1, 11.0])                         # one pixel
    frame = torch.arange(n) % F_
    got, flag = WL.depth_select(depth, frame, boxes.to(d), 0)
    assert bool(flag.all())
    for i in range(n):
        b = boxes[i]
        xs = [int(b[0].clamp(0, W)), int(b[2].clamp(0, W))]; ys = [int(b[1].clamp(0, H)), int(b[3].clamp(0, H))]
        ref = torch.median(depth[frame[i], ys[0]:ys[1], xs[0]:xs[1]])
        assert got[i].item() == ref.item(), i
    yy, xx = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    masks = torch.stack([((xx - 100 - 13 * i) ** 2 + (yy - 90 - 9 * i) ** 2 < (20 + 11 * i) ** 2) for i in range(n)]).to(d)
    masks[5] = False
    got, flag = WL.depth_select(depth, frame, boxes.to(d), 2, masks=masks, mask_idx=torch.arange(n))
    assert flag.tolist() == [0 if i == 5 else 1 for i in range(n)]
    for i in range(n):
        v = depth[frame[i]][masks[i]]
        if i == 5:
            b = boxes[i].long()
            v = depth[frame[i], b[1]:b[3], b[0]:b[2]].reshape(-1)
        ref = torch.quantile(v, 0.9) - torch.quantile(v, 0.1)
        assert abs(got[i].item() - ref.item()) <= 1e-6 * max(1.0, abs(ref.item())), (i, got[i].item(), ref.item())
