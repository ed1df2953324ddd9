"""One-off stress comparison (not a pytest): weak-loss geometry (tdod.weak_loss) against the torch-autograd oracle
(oracle/weak_loss_oracle.py) on many seeded cubes at the C2 frame shape: hull vertices bit-equal, loss values, gradients
with respect to the corners.    python tests/probes/weakloss_sweep.py [n_frames]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (os.path.join(ROOT, "3dod_b200"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import numpy as np
import torch

import tdod_oracle as O
import weak_loss_oracle as WO
from tdod import synth, weak_loss as WL

n_frames = int(sys.argv[1]) if len(sys.argv) > 1 else 12
d = torch.device("cuda:0")
H, W = 530, 730
tot = dict(cubes=0, hull_bad=0, box_bad=0, giou_bad=0, z_near_tie=0, z_bad=0, focal_bad=0, bce_bad=0, dice_bad=0, grad_bad=0,
           worst_grad=0.0, worst_val=0.0)
for k in range(n_frames):
    f = synth.make_frame(13000 + 3 * k, H, W, 12, with_masks=True)
    N = f.n_obj()
    g = torch.Generator().manual_seed(k)
    cubes = O.propose(f.boxes, f.depth, f.prior_mean, f.prior_std, f.K, 3, normal=torch.tensor([0.0, 1.0, 0.05]),
                      generator=g)["cubes"][:, k % 3].float().clone()
    cubes[::5, 2] *= 0.3
    cubes[2::7, 0] += 2.0
    Ks = f.K[None].repeat(N, 1, 1)
    ims = [(W, H)] * N
    idx = list(range(N))
    # oracle
    rc = cubes.clone().requires_grad_(True)
    rcor = WO.bube_corners(rc, Ks, ims)
    rseg = WO.segment_corners(rcor, ims)
    rbox = WO.cubes_to_box(rc, Ks, ims)
    rgiou = WO.giou_loss(f.boxes, rbox)
    rz = WO.z_loss(f.boxes, cubes, Ks, ims, rbox.detach())
    seg_in = rseg.detach().clone().requires_grad_(True)
    rl = {kind: WO.segment_loss(f.masks, seg_in, idx, loss=kind) for kind in ("focal", "bce", "dice")}
    w = torch.rand(N, generator=g)
    rgrad = torch.autograd.grad((rl["focal"] * w).sum(), seg_in)[0]
    # device
    c = cubes.to(d).requires_grad_(True)
    o = WL.project_boxes(c, Ks.to(d), ims, gt_boxes=f.boxes.to(d))
    hull, meta = WL.hull(rseg.detach().to(d))                  # the same fp32 corners on both sides
    for i in range(N):
        rh = WO.jarvis_march(rseg[i].detach())[0].numpy()
        m = int(meta[i, 0])
        tot["hull_bad"] += not (m == len(rh) and np.array_equal(hull[i, :m].cpu().numpy(), rh))
    close = lambda a, b, rt, at: bool(np.all(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64)) <= at + rt * np.abs(np.asarray(b, np.float64))))
    tot["box_bad"] += not close(o["boxes"].detach().cpu(), rbox.detach(), 1e-5, 2e-4)
    tot["giou_bad"] += not close(o["loss_iou"].detach().cpu(), rgiou.detach(), 1e-5, 2e-6)
    z = WL.z_loss(f.boxes.to(d), c.detach(), Ks.to(d), ims, o["boxes"].detach()).cpu().numpy()
    dz = np.abs(z - rz.numpy())
    tot["z_near_tie"] += int(((dz > 1e-6) & (dz <= 0.05 + 1e-6)).sum())     # argmin over fp32 areas: neighbouring 0.1 m step
    tot["z_bad"] += int((dz > 0.05 + 1e-6).sum())
    tot.setdefault("corner_bad", 0)
    tot["corner_bad"] += not close(o["seg_corners"].detach().cpu(), rseg.detach(), 1e-5, 2e-4)
    x = rseg.detach().to(d).requires_grad_(True)               # identical inputs: the band pixels are decided by the last bit
    for kind in ("focal", "bce", "dice"):
        got = WL.segment_loss(f.masks.to(d), x, idx, loss=kind)
        ok = close(got.detach().cpu(), rl[kind].detach(), 2e-5, 1e-7)
        tot[kind + "_bad"] += not ok
        tot["worst_val"] = max(tot["worst_val"], float((got.detach().cpu() - rl[kind].detach()).abs().max()))
    gg = torch.autograd.grad((WL.segment_loss(f.masks.to(d), x, idx) * w.to(d)).sum(), x)[0].cpu()
    scale = float(rgrad.abs().max()) + 1e-12
    e = float((gg - rgrad).abs().max()) / scale
    tot["worst_grad"] = max(tot["worst_grad"], e)
    tot["grad_bad"] += e > 1e-3
    tot["cubes"] += N
    print("frame", k, tot, flush=True)
print(tot)
