"""Where the time of tdod.mabo.mabo goes (C2-shaped frames): host wall per phase with a synchronize after each."""
import os, sys, time, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "3dod_b200"))
from tdod import api, mabo, pipeline, synth
d = torch.device("cuda:0")
F = int(sys.argv[1]) if len(sys.argv) > 1 else 8
b = pipeline.FrameBatch.from_frames(synth.make_frames("C2", range(F), device=d), d)
n = 1000
gt = pipeline.select(b, n, seed=99).best_cube.clone()
N = b.n_obj
H, W = int(b.depth.shape[1]), int(b.depth.shape[2])
def T(label, fn, reps=3):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): r = fn()
    torch.cuda.synchronize()
    print("%-34s %8.3f ms" % (label, (time.perf_counter() - t0) * 1e3 / reps)); return r
def gtboxes():
    g = torch.empty((N, 4), dtype=torch.float32, device=d)
    for f in torch.unique(b.obj_frame).tolist():
        m = b.obj_frame == f
        g[m] = api.project_cubes(gt[m].contiguous(), b.K[f], W, H, corners=False)[1]
    return g
print("frames", F, "objects", N)
T("whole mabo", lambda: mabo.mabo(b, gt, n, seed=5))
g = T("gt boxes (per-frame loop)", gtboxes)
T("select mode A (winners only)", lambda: pipeline.select(b, n, seed=5, iou_boxes=g))
sel = T("select mode_b=True", lambda: pipeline.select(b, n, seed=5, iou_boxes=g, mode_b=True))
o = sel.all
cg = api.cube_corners(gt.contiguous())
T("cube_corners(all)", lambda: api.cube_corners(o.all_cubes))
ca = api.cube_corners(o.all_cubes)
iou3d = T("box3d_iou", lambda: api.box3d_iou(cg, ca))
ch = T("score_channels", lambda: mabo.score_channels(o))
ch["random"] = torch.rand((N, n), dtype=torch.float64, device=d)
singles = T("stack singles", lambda: torch.stack([ch[k].to(torch.float64) for k in mabo.CURVES], 1).contiguous())
T("accumulate 8 curves", lambda: api.accumulate_scores(singles, iou3d))
prods = T("26 numpy products", lambda: torch.stack([mabo.numpy_product([ch[k] for k in names]).to(torch.float64) for names in mabo.COMBINATIONS], 1).contiguous())
T("accumulate 26 products (n_out=1)", lambda: api.accumulate_scores(prods, iou3d, n_out=1))
