"""Stage times of the weak-loss geometry on the cubes of 8 C2-shaped frames: CUDA events around
project_boxes / segment_loss / z_loss / backward, and the two fill kernels alone.  python tools/weakloss_stages.py"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))  # tools/ -> repo root
sys.path.insert(0, os.path.join(ROOT, "3dod_b200"))
import torch
from tdod import pipeline, synth, weak_loss as WL, lib as L
dev = torch.device("cuda:0")
c = synth.CONFIGS["C2"]; H, W = c["H"], c["W"]
fr = [synth.make_frame(i, H, W, c["n_obj"], device=dev) for i in range(8)]
batch = pipeline.FrameBatch.from_frames(fr, dev)
dc = pipeline.select(batch, 8, seed=3, return_all=True).all.all_cubes[:, 0].contiguous()    # proposals of the product path
dK, db, dm = batch.K[batch.obj_frame.long()].contiguous(), batch.boxes, batch.masks
N = dc.shape[0]
im_t = torch.tensor([(W, H)] * N, dtype=torch.int32, device=dev); idx_t = torch.arange(N, dtype=torch.int32, device=dev)
def ev(): return torch.cuda.Event(enable_timing=True)
acc = {}
for it in range(13):
    e = [ev() for _ in range(6)]
    x = dc.clone().requires_grad_(True)
    e[0].record()
    o = WL.project_boxes(x, dK, im_t, gt_boxes=db)
    e[1].record()
    seg = WL.segment_loss(dm, o["seg_corners"], idx_t)
    e[2].record()
    z = WL.z_loss(db, x, dK, im_t, o["boxes"])
    e[3].record()
    loss = o["loss_iou"].mean() + seg.mean() + z.mean()
    e[4].record()
    loss.backward()
    e[5].record()
    torch.cuda.synchronize()
    if it >= 3:
        for k, (a, b) in {"project_fwd": (0, 1), "segment_fwd": (1, 2), "z_loss": (2, 3), "loss_sum": (3, 4), "backward": (4, 5), "total": (0, 5)}.items():
            acc.setdefault(k, []).append(e[a].elapsed_time(e[b]))
print(json.dumps({k: sum(v) / len(v) for k, v in acc.items()}))
# kernels alone
h, meta = WL.hull(o["seg_corners"])
mom = torch.empty((N, 12), dtype=torch.float64, device=dev)
ws = torch.empty((N, 24), dtype=torch.float64, device=dev); gout = torch.empty((N, 8, 2), device=dev)
gm = torch.rand((N, 12), dtype=torch.float64, device=dev)
mu8 = dm.view(torch.uint8)
def t(fn, n=20):
    for _ in range(3): fn()
    a, b = ev(), ev(); a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize(); return a.elapsed_time(b) / n
print("moments kernel ms", t(lambda: L.check(L.lib().tdod_hull_fill_moments(L.ptr(mu8), L.ptr(idx_t), L.ptr(h), L.ptr(meta), N, H, W, 0, L.ptr(mom), L.stream_ptr()))))
print("grad kernels ms", t(lambda: L.check(L.lib().tdod_hull_fill_grad(L.ptr(mu8), L.ptr(idx_t), L.ptr(h), L.ptr(meta), N, H, W, L.ptr(gm), L.ptr(ws), L.ptr(gout), L.stream_ptr()))))
# the depth-map losses: one launch for all boxes against the reference's per-box loop of torch.median / torch.quantile
depth, fidx = batch.depth, batch.obj_frame
print("depth_select median (mode 0) ms", t(lambda: WL.depth_select(depth, fidx, db, 0)))
print("depth_select q90-q10 under masks (mode 2) ms", t(lambda: WL.depth_select(depth, fidx, db, 2, masks=mu8, mask_idx=idx_t)))
def torch_loop():
    bl = db.long()
    return [torch.median(depth[fidx[i], bl[i, 1]:bl[i, 3], bl[i, 0]:bl[i, 2]]) for i in range(N)]
def torch_loop_q():
    out = []
    for i in range(N):
        v = depth[fidx[i]][dm[i]]
        out.append(torch.quantile(v, 0.9) - torch.quantile(v, 0.1))
    return out
print("torch per-box median loop ms", t(torch_loop, 5), " torch per-mask quantile loop ms", t(torch_loop_q, 5), " boxes", N)
