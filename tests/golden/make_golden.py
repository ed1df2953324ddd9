"""Generates tests/golden/*.npz by running the UNMODIFIED reference (imported from /root/reference via
oracle/ref_import.py) on synthetic frames from 3dod_b200/tdod/synth.py.  Runs only in the build
container (the GPU box has no /root/reference); the resulting fixtures are committed.

Inputs are regenerated from (frame_id, H, W, n_obj) by synth.make_frame -- each fixture stores a SHA-256
of the regenerated inputs so a drift in the generator is detected -- plus everything the reference
computed: accepted standard normals, cubes, projected corners, boxes, every score, the argmax.

    python tests/golden/make_golden.py
"""
import hashlib
import os
import pickle
import random
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "3dod_b200"))

import ref_import  # noqa: E402
import tdod_oracle as O  # noqa: E402
from tdod import synth  # noqa: E402

R = ref_import.load()


def digest(*arrays):
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def frame_digest(f):
    parts = [f.depth.numpy(), f.ground.numpy(), f.boxes.numpy(), f.prior_mean.numpy(),
             f.prior_std.numpy(), f.K.numpy()]
    if f.masks is not None:
        parts.append(f.masks.numpy())
    return digest(*parts)


def ref_normal(f, seed):
    """roi_heads.py:345-428 flavour: fp64 NumPy back-projection, Plane_np (== pyransac3d), rectify."""
    pg, pn = O.backproject_np(f.depth.numpy(), f.K.numpy(), f.ground.numpy())
    random.seed(seed)
    eq, inl = R.plane.Plane_np().fit(pg, thresh=0.05, maxIteration=1000)
    random.seed(seed)
    ids = O.draw_sample_ids(len(pg), 1000, random)
    nv = O.rectify_normal(np.array(eq[:-1]))
    return pg, pn, ids, np.array(eq), np.asarray(inl), nv


def ref_frame(f, n, seed, normal, W, H, mode_b_objs=()):
    """Runs the reference's PRED-branch selection (roi_heads.py:493-505) for one frame."""
    out = {}
    boxes = R.Boxes(f.boxes)
    torch.manual_seed(seed)
    cubes, _, _ = R.proposals.propose(boxes, f.depth, (f.prior_mean, f.prior_std), (W, H), f.K,
                                      number_of_proposals=n,
                                      ground_normal=None if normal is None else torch.from_numpy(normal))
    # accepted standard normals: re-run the oracle's sampler on the same stream (bit-equal cubes is
    # asserted, so the noise is the one the reference consumed)
    g = torch.Generator().manual_seed(seed)
    o = O.propose(f.boxes, f.depth, f.prior_mean, f.prior_std, f.K, n,
                  None if normal is None else torch.from_numpy(normal), g)
    assert torch.equal(o["cubes"], cubes.tensor), "oracle sampler diverged from the reference"
    out["noise"] = o["noise"].numpy()
    if o["yaw_idx"] is not None:
        out["yaw_idx"] = o["yaw_idx"].numpy().astype(np.int16)
    out["tries"] = np.array(o["tries"])
    out["cubes"] = cubes.tensor.numpy()
    pb = R.conversions.cubes_to_box(cubes, f.K, (W, H))
    out["boxes"] = np.stack([b.tensor.numpy() for b in pb])
    N = len(boxes)
    iou = np.zeros((N, n), np.float32); dim = np.zeros((N, n), np.float32); cor = np.zeros((N, n), np.float32)
    comb = np.zeros((N, n), np.float32); best = np.zeros(N, np.int64)
    rects = np.full((N, 4, 2), np.nan, np.float32)
    c2d_all = cubes.get_bube_corners(f.K, (W, H))
    out["corners2d_obj0"] = c2d_all[0].numpy()
    for i in range(N):
        c2d = cubes[i].get_bube_corners(f.K, (W, H))
        s_iou = R.scorefunction.score_iou(boxes[i], pb[i])
        s_dim = R.scorefunction.score_dimensions((f.prior_mean[i], f.prior_std[i]), cubes[i].dimensions[0],
                                                 boxes[i], pb[i])
        s_cor = R.scorefunction.score_corners(f.masks[i], c2d)
        c = np.asarray(s_iou) * np.asarray(s_dim) * np.asarray(s_cor)
        iou[i], dim[i], cor[i], comb[i] = s_iou.numpy(), s_dim.numpy(), s_cor.numpy(), c
        best[i] = int(np.argmax(c))
        r = O.mask_rect(f.masks[i].numpy().astype(np.uint8))
        if r is not None:
            rects[i] = r
    out.update(iou=iou, dim=dim, cor=cor, combined=comb, best=best, rects=rects)
    for i in mode_b_objs:
        c2d = cubes[i].get_bube_corners(f.K, (W, H))
        out["seg_%d" % i] = R.scorefunction.score_segmentation(f.masks[i], c2d).numpy()
        out["modseg_%d" % i] = R.scorefunction.score_mod_segmentation(f.masks[i], c2d).numpy()
    return out, cubes


def main():
    ver = O.versions()
    # ---- G1: BASELINE config 1: 640x480, 5 boxes x 1000 cubes, mode A, ground normal from RANSAC ----
    W, H, n = 640, 480, 1000
    f = synth.make_frame(0, H, W, 5)
    pg, pn, ids, eq, inl, nv = ref_normal(f, seed=0)
    out, cubes = ref_frame(f, n, seed=13, normal=nv, W=W, H=H)
    # point-cloud score of object 0 against the no-ground points (fp64), roi_heads.py:535
    pc = R.scorefunction.score_point_cloud(torch.from_numpy(pn), cubes[0]).numpy()
    np.savez_compressed(os.path.join(HERE, "c1_modeA.npz"), versions=ver, frame_id=0, H=H, W=W, n_obj=5, n=n,
                        seed=13, ransac_seed=0, input_sha=frame_digest(f), plane_eq=eq, n_inliers=len(inl),
                        inlier_sha=digest(inl.astype(np.int64)), sample_ids=ids.astype(np.int32),
                        normal=nv, n_plane_pts=len(pg), n_noground_pts=len(pn), pc_obj0=pc, **out)

    # ---- G2: mode B on a smaller proposal count (segmentation / mod-seg are slow in the reference) ----
    n2 = 256
    out2, cubes2 = ref_frame(f, n2, seed=21, normal=nv, W=W, H=H, mode_b_objs=(0, 1, 2, 3, 4))
    pcs = np.stack([R.scorefunction.score_point_cloud(torch.from_numpy(pn), cubes2[i]).numpy() for i in range(5)])
    np.savez_compressed(os.path.join(HERE, "c1_modeB.npz"), versions=ver, frame_id=0, H=H, W=W, n_obj=5, n=n2,
                        seed=21, input_sha=frame_digest(f), normal=nv, pc=pcs, **out2)

    # ---- G3: no ground normal -> random rotations (randn_orthobasis_torch), other frame / shape ----
    W3, H3, n3 = 730, 530, 128
    f3 = synth.make_frame(7, H3, W3, 3)
    out3, _ = ref_frame(f3, n3, seed=5, normal=None, W=W3, H=H3, mode_b_objs=(0,))
    np.savez_compressed(os.path.join(HERE, "c2_randrot.npz"), versions=ver, frame_id=7, H=H3, W=W3, n_obj=3, n=n3,
                        seed=5, input_sha=frame_digest(f3), **out3)

    # ---- G4: planes.  torch flavour (fp32 fit_parallel) on the synthetic frame; NumPy flavour and
    #      fit_parallel on a fixed subsample of the reference's caixa.ply fixture (test_plane.py:29-35) ----
    f4 = synth.make_frame(1, 480, 640, 5)
    pts32 = O.backproject_torch(f4.depth, f4.K, f4.ground)
    random.seed(3)
    eq32, inl32 = R.plane.Plane().fit_parallel(pts32, 0.05, maxIteration=1000)
    random.seed(3)
    ids32 = O.draw_sample_ids(len(pts32), 1000, random)
    raw = open(os.path.join(ref_import.REF_ROOT, "ProposalNetwork/utils/tests/caixa.ply"), "rb").read()
    hd = raw.index(b"end_header\n") + len(b"end_header\n")
    arr = np.frombuffer(raw[hd:], dtype=np.dtype([("p", "<f8", 3), ("c", "u1", 3)]))
    caixa = np.ascontiguousarray(arr["p"])
    random.seed(0)
    eq_full, inl_full = R.plane.Plane_np().fit(caixa, thresh=0.01, maxIteration=1000)
    sub = caixa[::6].copy()                                    # 7729 points, fp64
    random.seed(1)
    eq_sub, inl_sub = R.plane.Plane_np().fit(sub, thresh=0.01, maxIteration=1000)
    random.seed(1)
    ids_sub = O.draw_sample_ids(len(sub), 1000, random)
    sub32 = torch.from_numpy(sub).float()
    random.seed(1)
    eq_sub32, inl_sub32 = R.plane.Plane().fit_parallel(sub32, 0.01, maxIteration=1000)
    np.savez_compressed(os.path.join(HERE, "planes.npz"), versions=ver, frame_id=1, H=480, W=640, n_obj=5,
                        input_sha=frame_digest(f4), torch_n_pts=len(pts32), torch_ids=ids32.astype(np.int32),
                        torch_eq=eq32.numpy(), torch_inlier_sha=digest(inl32.numpy()), torch_n_inliers=len(inl32),
                        caixa_full_eq=np.array(eq_full), caixa_full_n_inliers=len(inl_full),
                        caixa_sub=sub, caixa_sub_ids=ids_sub.astype(np.int32), caixa_sub_eq=np.array(eq_sub),
                        caixa_sub_n_inliers=len(inl_sub), caixa_sub_inlier_sha=digest(np.asarray(inl_sub, np.int64)),
                        caixa_sub32_eq=eq_sub32.numpy(), caixa_sub32_n_inliers=len(inl_sub32))

    # ---- G5: the reference's real SAM-HQ masks (ProposalNetwork/mask0.pkl, mask1.pkl) -> rectangle ----
    real = {}
    for name in ("mask0", "mask1"):
        m = pickle.load(open(os.path.join(ref_import.REF_ROOT, "ProposalNetwork", name + ".pkl"), "rb"))
        m = np.asarray(m).astype(bool).reshape(m.shape[-2], m.shape[-1])
        real[name + "_shape"] = np.array(m.shape)
        real[name + "_bits"] = np.packbits(m)
        real[name + "_rect"] = O.mask_rect(m.astype(np.uint8))
        # score_corners on a handful of synthetic projected corners through the reference
        rs = np.random.RandomState(3)
        c2d = torch.tensor(rs.uniform(0, max(m.shape), (1, 64, 8, 2)).astype(np.float32))
        real[name + "_c2d"] = c2d[0].numpy()
        real[name + "_cor"] = R.scorefunction.score_corners(torch.from_numpy(m), c2d).numpy()
    np.savez_compressed(os.path.join(HERE, "real_masks.npz"), versions=ver, **real)

    # ---- G6: degenerate semantics: empty mask (fallback rectangle), zero-height boxes -> NaN-first argmax,
    #      cubes behind the camera (no z>0 guard), point-cloud known answer (scorefunction.py:199-201) ----
    Wd, Hd = 320, 240
    Kd = torch.tensor([[288.0, 0, 160.0], [0, 288.0, 120.0], [0, 0, 1]])
    rs = np.random.RandomState(11)
    nd = 64
    t = np.zeros((1, nd, 15), np.float32)
    t[0, :, 0:3] = rs.uniform([-1, -1, 1.5], [1, 1, 4], (nd, 3))
    t[0, :, 3:6] = rs.uniform(0.2, 1.0, (nd, 3))
    Rm = O.random_rotations(1, nd, torch.Generator().manual_seed(2))
    t[0, :, 6:] = Rm.reshape(1, nd, 9).numpy()
    t[0, 5, 2] = -2.0            # behind the camera
    t[0, 9, 2] = 0.05            # straddles the image plane
    t[0, 17, 0:3] = [30.0, 40.0, 1.0]   # projects far outside -> clamped to a zero-height box -> NaN score
    t[0, 40, 0:3] = [30.0, 40.0, 1.0]
    cub = R.spaces.Cubes(torch.from_numpy(t))
    gt = R.Boxes(torch.tensor([[60.0, 50.0, 200.0, 170.0]]))
    pbd = R.conversions.cubes_to_box(cub, Kd, (Wd, Hd))
    c2d = cub.get_bube_corners(Kd, (Wd, Hd))
    pm, ps = torch.tensor([0.6, 0.5, 0.7]), torch.tensor([0.2, 0.15, 0.25])
    s_iou = R.scorefunction.score_iou(gt, pbd[0])
    s_dim = R.scorefunction.score_dimensions((pm, ps), cub[0].dimensions[0], gt, pbd[0])
    empty = torch.zeros((Hd, Wd), dtype=torch.bool)
    s_cor = R.scorefunction.score_corners(empty, c2d)
    comb = np.asarray(s_iou) * np.asarray(s_dim) * np.asarray(s_cor)
    pts3 = torch.tensor([[0.1, 0.1, 0.1], [0.2, 0.2, 0.2], [-3, 0, 0]])
    unit = torch.tensor([[[0.5, 0.5, 0.5, 1, 1, 1, 1, 0, 0, 0, 1, 0, 0, 0, 1]]], dtype=torch.float32)
    pc_unit = R.scorefunction.score_point_cloud(pts3, R.spaces.Cubes(unit)).numpy()
    np.savez_compressed(os.path.join(HERE, "degenerate.npz"), versions=ver, W=Wd, H=Hd, K=Kd.numpy(), cubes=t,
                        ref_box=gt.tensor.numpy(), prior_mean=pm.numpy(), prior_std=ps.numpy(),
                        boxes=pbd[0].tensor.numpy(), corners2d=c2d[0].numpy(), iou=s_iou.numpy(), dim=s_dim.numpy(),
                        cor_emptymask=s_cor.numpy(), combined=comb, best=int(np.argmax(comb)),
                        pc_points=pts3.numpy(), pc_cube=unit.numpy(), pc_count=pc_unit)
    make_variants(ver)
    for fn in sorted(os.listdir(HERE)):
        if fn.endswith(".npz"):
            print(fn, os.path.getsize(os.path.join(HERE, fn)) // 1024, "KiB")


VARIANTS = ("random", "xy", "z", "dim", "aspect", "rotation")


def make_variants(ver):
    """the six ablation samplers (proposals.py:20-336) of the UNMODIFIED reference on one C1-shaped frame; the random
    numbers each consumed are recovered by re-running the oracle's restatement on the same seed (bit-equal cubes asserted)"""
    fid, H, W, nobj, n = 11, 480, 640, 5, 400
    f = synth.make_frame(fid, H, W, nobj)
    fn = dict(random=R.proposals.propose_random, xy=R.proposals.propose_xy_patch, z=R.proposals.propose_z,
              dim=R.proposals.propose_random_dim, aspect=R.proposals.propose_aspect_ratio,
              rotation=R.proposals.propose_random_rotation)
    out = dict(versions=ver, frame_id=fid, H=H, W=W, n_obj=nobj, n=n, input_sha=frame_digest(f))
    for k, name in enumerate(VARIANTS):
        seed = 500 + k
        torch.manual_seed(seed)
        cubes, _, _ = fn[name](R.Boxes(f.boxes), f.depth, (f.prior_mean, f.prior_std), (W, H), f.K, number_of_proposals=n)
        torch.manual_seed(seed)
        o = O.propose_variant(name, f.boxes, f.depth, f.prior_mean, f.prior_std, f.K, (W, H), n)
        assert torch.equal(o["cubes"], cubes.tensor), "oracle %s diverged from the reference" % name
        out[name + "_cubes"] = cubes.tensor.numpy()
        out[name + "_noise"] = o["noise"].numpy()
        out[name + "_rot"] = o["rot"].numpy()
        if o["aux"] is not None:
            out[name + "_aux"] = o["aux"].numpy()
    np.savez_compressed(os.path.join(HERE, "variants.npz"), **out)


def _ref_methods(names):
    """the UNMODIFIED source of methods of ROIHeads3DScore (cubercnn/modeling/roi_heads/roi_heads.py), exec'd from the
    file where it lies (the module itself needs detectron2 / pytorch3d / segment_anything, absent here)"""
    import ast
    import textwrap
    import torch.nn.functional as F
    from torchvision.ops import sigmoid_focal_loss
    path = os.path.join(ref_import.REF_ROOT, "cubercnn/modeling/roi_heads/roi_heads.py")
    src = open(path).read()
    tree = ast.parse(src)
    cls = [n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "ROIHeads3DScore"][0]
    ns = {"torch": torch, "F": F, "Boxes": R.Boxes, "Cubes": R.spaces.Cubes, "cubes_to_box": R.conversions.cubes_to_box,
          "convex_hull": R.utils.convex_hull, "sigmoid_focal_loss": sigmoid_focal_loss}
    for fn in cls.body:
        if isinstance(fn, ast.FunctionDef) and fn.name in names:
            exec(textwrap.dedent(ast.get_source_segment(src, fn)), ns)

    class Self:
        pass
    me = Self()
    for n in names:
        setattr(me, n, ns[n].__get__(me))
    return me


def make_weak_loss(ver):
    """cubes_to_box with autograd, GIoU loss, z_loss, segment_loss of the UNMODIFIED reference on
    the winners-like cubes of a small frame (120x160 so that the fixture stays small)."""
    from torchvision.ops import generalized_box_iou_loss
    me = _ref_methods(["l1_loss", "dice_loss", "segment_loss", "z_loss"])
    H, W, n = 120, 160, 14
    f = synth.make_frame(21, H, W, n, with_masks=True)
    N = f.boxes.shape[0]
    g = torch.Generator().manual_seed(5)
    cubes = O.propose(f.boxes, f.depth, f.prior_mean, f.prior_std, f.K, 4, normal=torch.tensor([0.05, 0.99, 0.1]),
                      generator=g)["cubes"]
    cubes = cubes[:, 1].float().clone()                                   # one proposal per object, [N,15]
    cubes[0, 2] = 0.2          # straddles the camera plane: projections blow up and clamp
    cubes[1, 0] += 3.0         # far to the right: clamped on x
    cubes[2, 3:6] *= 0.05      # tiny cube: corners nearly coincide
    Ks = f.K[None].repeat(N, 1, 1).clone()
    Ks[:, 0, 2] += torch.arange(N) * 0.5                                  # per-object intrinsics (roi_heads.py:1539)
    ims = [(W, H)] * N
    cubes.requires_grad_(True)
    wb = torch.rand((N, 4), generator=g)
    wl = torch.rand(N, generator=g)
    # roi_heads.py:1571-1583
    corners, segc, boxes = [], [], []
    for i in range(N):
        cb = R.spaces.Cubes(cubes[i][None, None])
        bc = cb.get_bube_corners(Ks[i], ims[i])
        corners.append(bc[0, 0])
        x = torch.clamp(bc[..., 0], 0, int(ims[i][0] - 1))
        y = torch.clamp(bc[..., 1], 0, int(ims[i][1] - 1))
        segc.append(torch.stack((x, y), dim=-1)[0, 0])
        boxes.append(R.conversions.cubes_to_box(cb, Ks[i], ims[i])[0].tensor[0])
    corners, segc, boxes = torch.stack(corners), torch.stack(segc), torch.stack(boxes)
    gt = R.Boxes(f.boxes)
    out = {"versions": ver, "frame": np.array([21, H, W, n]), "inputs_sha256": np.array(frame_digest(f)),
           "cubes": cubes.detach().numpy(), "Ks": Ks.numpy(), "wb": wb.numpy(), "wl": wl.numpy(),
           "corners": corners.detach().numpy(), "seg_corners": segc.detach().numpy(), "boxes": boxes.detach().numpy()}
    out["grad_box"] = torch.autograd.grad((boxes * wb).sum(), cubes, retain_graph=True)[0].numpy()
    giou = generalized_box_iou_loss(gt.tensor, boxes, reduction='none').view(N, -1).mean(dim=1)        # :1599
    out["giou"] = giou.detach().numpy()
    out["grad_giou"] = torch.autograd.grad((giou * wl).sum(), cubes, retain_graph=True)[0].numpy()
    z = me.z_loss(gt, R.spaces.Cubes(cubes[:, None]), Ks, ims, R.Boxes(boxes.detach()))
    out["z_loss"] = z.detach().numpy()
    masks = [m[None] for m in f.masks]
    idx = list(range(N))
    for kind in ("focal", "bce", "dice"):
        sl = me.segment_loss(masks, segc, idx, loss=kind)
        out["seg_" + kind] = sl.detach().numpy()
        wv = wl if sl.dim() else torch.tensor(1.0)
        gc, gq = torch.autograd.grad((sl * wv).sum(), (segc, cubes), retain_graph=True, allow_unused=True)
        out["grad_seg_%s_corners" % kind] = gc.numpy()
        out["grad_seg_%s_cubes" % kind] = gq.numpy()
    hulls = [R.utils.jarvis_march(segc[i].detach()) for i in range(N)]
    out["hull_len"] = np.array([len(h) for h in hulls])
    out["hull"] = np.stack([np.pad(h.numpy(), ((0, 8 - len(h)), (0, 0))) for h in hulls])
    np.savez_compressed(os.path.join(HERE, "weak_loss.npz"), **out)
    print("weak_loss.npz", {k: getattr(v, "shape", None) for k, v in out.items() if k.startswith(("seg_", "z_", "giou"))})
    print("hull_len", out["hull_len"], "z", out["z_loss"], "focal", out["seg_focal"], "dice", out["seg_dice"])


def make_depth_losses(ver):
    """SURVEY 8(f) row 4, the losses that read the depth map: pseudo_gt_z_box_loss, pseudo_gt_z_point_loss and
    depth_range_loss of the UNMODIFIED reference (roi_heads.py:1196-1231, :1256-1305)."""
    me = _ref_methods(["l1_loss", "pseudo_gt_z_box_loss", "pseudo_gt_z_point_loss", "depth_range_loss"])
    sys.path.insert(0, HERE)
    from depth_case import depth_loss_case
    c = depth_loss_case()
    ok = [b.clone() for b in c["boxes"]]
    ok[1] = torch.cat((ok[1][:3], ok[1][4:]))                          # the reference raises on the empty slice
    keep = torch.ones(sum(c["counts"]), dtype=torch.bool)
    keep[c["counts"][0] + 3] = False
    out = {"versions": ver, "keep": keep.numpy()}
    out["z_box"] = me.pseudo_gt_z_box_loss(list(c["depth"]), ok, c["pred_z"][keep]).numpy()
    out["z_point"] = me.pseudo_gt_z_point_loss(list(c["depth"]), c["pred_xy"], c["pred_z"], c["counts"]).numpy()
    cubes = c["cubes"].clone().requires_grad_(True)
    dr = me.depth_range_loss(c["masks"][:, None], c["mask_idx"].tolist(), list(c["depth"]),
                             R.spaces.Cubes(cubes[:, None]), R.Boxes(c["gt_boxes"]), c["counts"])
    out["depth_range"] = dr.detach().numpy()
    out["grad_depth_range"] = torch.autograd.grad(dr.sum(), cubes)[0].numpy()
    np.savez_compressed(os.path.join(HERE, "depth_losses.npz"), **out)
    print("depth_losses.npz", {k: (v.shape, v.dtype) for k, v in out.items() if k != "versions"})
    print("z_box", out["z_box"][:8], "depth_range", out["depth_range"][:8])


def make_pseudo_record(ver):
    """one image record + its annotation records built by the UNMODIFIED record-building statements of
    tools/eval_boxes.py (the `img = {...}` assignment :646 and the annotation loop :647-655), exec'd from the file where
    they lie, on synthetic `instances` / `input_` stand-ins.  `bboxes` (:637-640) comes from
    GeneralizedRCNN._postprocess -> detectron2's detector_postprocess (absent here: public behaviour restated in
    oracle/tdod_oracle.py::postprocess_boxes)."""
    import ast
    import json
    import textwrap
    import types
    path = os.path.join(ref_import.REF_ROOT, "tools/eval_boxes.py")
    src = open(path).read()
    tree = ast.parse(src)
    img_stmt = loop_stmt = None
    for node in ast.walk(tree):
        if isinstance(node, ast.Assign) and isinstance(node.targets[0], ast.Name) and node.targets[0].id == "img" \
                and isinstance(node.value, ast.Dict) and img_stmt is None and node.lineno > 600:
            img_stmt = node
        if isinstance(node, ast.For) and node.lineno > 600 and any(
                isinstance(b, ast.Assign) and getattr(b.targets[0], "id", "") == "ann" for b in node.body):
            loop_stmt = node
    assert img_stmt is not None and loop_stmt is not None
    code = textwrap.dedent(ast.get_source_segment(src, img_stmt)) + "\n" + textwrap.dedent(
        " " * loop_stmt.col_offset + ast.get_source_segment(src, loop_stmt))
    g = torch.Generator().manual_seed(5)
    n = 4
    net_h, net_w, H0, W0 = 512, 704, 530, 730                       # network resolution / original image size
    gt = torch.tensor([[10.5, 20.25, 300.0, 400.0], [-4.0, 30.0, 720.0, 520.0], [100.0, 100.0, 100.0, 180.0],
                       [50.0, 60.0, 200.0, 260.0]])
    gt_classes = torch.tensor([3, 17, 5, 9])
    cubes = torch.rand(n, 15, generator=g)
    corners = torch.rand(n, 8, 3, generator=g)
    inst = types.SimpleNamespace(pred_center_cam=cubes[:, :3], pred_dimensions=cubes[:, 3:6], pred_bbox3D=corners,
                                 pred_pose=cubes[:, 6:].reshape(n, 3, 3))
    input_ = {"image_id": 4711, "width": W0, "height": H0, "file_name": "datasets/SUNRGBD/kv2/x/image/0001.jpg",
              "K": [[657.0, 0.0, 365.0], [0.0, 657.0, 265.0], [0.0, 0.0, 1.0]],
              "instances": types.SimpleNamespace(gt_classes=gt_classes)}
    thing = ["class%02d" % k for k in range(50)]
    c2id = {k: 100 + k for k in range(50)}
    util = types.SimpleNamespace(MetadataCatalog=types.SimpleNamespace(get=lambda name: types.SimpleNamespace(thing_classes=thing)))
    bboxes = O.postprocess_boxes(gt, (net_h, net_w), H0, W0)        # all four kept here (the empty one is dropped there)
    keep = O.postprocess_keep(gt, (net_h, net_w), H0, W0)
    ns = {"input_": input_, "img_id": input_["image_id"], "instances": inst, "bboxes": bboxes, "util": util,
          "contiguous_to_id": c2id, "global_id": 1, "idd": 12, "annotations": []}
    exec(code, ns)
    doc = {"versions": str(ver), "net_hw": [net_h, net_w], "gt_boxes": gt.tolist(), "gt_classes": gt_classes.tolist(),
           "keep": keep.tolist(), "cubes": cubes.tolist(), "corners": corners.tolist(), "thing_classes": thing,
           "input": {k: v for k, v in input_.items() if k != "instances"}, "image": ns["img"],
           "annotations": ns["annotations"], "exec_lines": [img_stmt.lineno, loop_stmt.end_lineno]}
    json.dump(doc, open(os.path.join(HERE, "pseudo_record.json"), "w"), indent=1)
    print("pseudo_record.json", len(ns["annotations"]), "annotations; image", ns["img"])


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "pseudo_record":
        make_pseudo_record(np.array(str(O.versions())))
    elif len(sys.argv) > 1 and sys.argv[1] == "depth_losses":
        make_depth_losses(np.array(str(O.versions())))
    elif len(sys.argv) > 1 and sys.argv[1] == "variants":
        make_variants(np.array(str(O.versions())))
    elif len(sys.argv) > 1 and sys.argv[1] == "weak_loss":
        make_weak_loss(np.array(str(O.versions())))
    else:
        main()
