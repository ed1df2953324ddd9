import os, sys, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (os.path.join(ROOT, "3dod_b200"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import tdod_oracle as O
from tdod import api, pipeline, synth, lib as L
d = torch.device("cuda:0")
g = np.load(os.path.join(ROOT, "tests/golden/c1_modeB.npz"))
f = synth.make_frame(int(g["frame_id"]), int(g["H"]), int(g["W"]), int(g["n_obj"]))
n, N, H, W = int(g["n"]), int(g["n_obj"]), int(g["H"]), int(g["W"])
gd = lambda t: t.to(d)
pts, other, cnt = api.backproject(gd(f.depth)[None], gd(f.ground)[None], gd(f.K)[None], 5, L.BACKPROJECT_NUMPY, want_other=True)
rects, has = api.mask_rect(gd(f.masks))
o = api.propose_score_select(gd(f.depth)[None], gd(f.K)[None], gd(f.boxes), gd(f.prior_mean), gd(f.prior_std),
                             torch.zeros(N, dtype=torch.int32, device=d), rects, has, n,
                             yaw_tab=api.yaw_table(torch.from_numpy(g["normal"]).to(d)[None]),
                             noise=torch.from_numpy(g["noise"]).permute(1, 0, 2).contiguous().to(d),
                             yaw_idx=torch.from_numpy(g["yaw_idx"]).to(d), masks=gd(f.masks), points=other,
                             n_points=cnt[:, 1].contiguous())
torch.cuda.synchronize()
for i in range(N):
    c2d, _ = api.project_cubes(o.all_cubes[i], gd(f.K), W, H, boxes=False)
    pf = api.score_segmentation(gd(f.masks[i]), c2d)
    print(i, "fused", o.all_scores[i, 4, :6].cpu().numpy(), "per-fn", pf[:6].cpu().numpy(), "ref", g["seg_%d" % i][:6])
    print("   fused==ref", float((o.all_scores[i, 4].cpu().numpy() == g["seg_%d" % i]).mean()), "per-fn==ref", float((pf.cpu().numpy() == g["seg_%d" % i]).mean()),
          "cubes close", float(np.abs(o.all_cubes[i].cpu().numpy() - g["cubes"][i]).max()))
    pcf = api.score_point_cloud(other[0, :, :int(cnt[0, 1])].t().contiguous(), o.all_cubes[i])
    print("   pc fused==ref", float((o.all_pc[i].cpu().numpy() == g["pc"][i]).mean()), "per-fn==ref", float((pcf.cpu().numpy() == g["pc"][i]).mean()))
