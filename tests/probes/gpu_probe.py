"""Diagnostic (not a pytest): prints max deviations of the CUDA path from the golden fixtures / oracle."""
import os, sys, hashlib
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "3dod_b200")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
from tdod import api, synth, lib as L
import tdod_oracle as O

dev = torch.device("cuda")
G = os.path.join(ROOT, "tests", "golden")

def rel(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.nanmax(np.abs(a - b) / np.maximum(np.abs(b), 1e-30)))

def run(name, S):
    g = np.load(os.path.join(G, name))
    H, W, n, N = int(g["H"]), int(g["W"]), int(g["n"]), int(g["n_obj"])
    f = synth.make_frame(int(g["frame_id"]), H, W, N)
    depth = f.depth[None].to(dev); K = f.K[None].to(dev)
    noise = torch.from_numpy(g["noise"]).permute(1, 0, 2).contiguous().to(dev)
    has_normal = "yaw_idx" in g.files
    yaw_tab = yaw_idx = rot = None
    cubes_g = torch.from_numpy(g["cubes"])
    if has_normal:
        normal = torch.from_numpy(g["normal"]).to(dev)[None]
        yaw_tab = api.yaw_table(normal)
        tab_o = O.yaw_table(torch.from_numpy(g["normal"])).to(torch.float32)
        print("  yaw table max abs diff vs oracle:", (yaw_tab[0].cpu() - tab_o).abs().max().item(),
              "n unequal", int((yaw_tab[0].cpu() != tab_o).sum()))
        yaw_tab = tab_o[None].contiguous().to(dev)   # exact table so R parity is about the kernel only
        yaw_idx = torch.from_numpy(g["yaw_idx"]).to(dev)
    else:
        rot = cubes_g[:, :, 6:].contiguous().to(dev)
    rects = torch.from_numpy(g["rects"]).to(dev)
    has = torch.ones(N, dtype=torch.uint8, device=dev)
    o = api.propose_score_select(depth, K, f.boxes.to(dev), f.prior_mean.to(dev), f.prior_std.to(dev),
                                 torch.zeros(N, dtype=torch.int32, device=dev), rects, has, n, yaw_tab=yaw_tab,
                                 noise=noise, yaw_idx=yaw_idx, rot=rot, return_all=True, cluster_size=S)
    torch.cuda.synchronize()
    c = o.all_cubes.cpu().numpy(); cg = g["cubes"]
    print("  [S=%d] whl exact: %s   R exact: %s   xyz max rel: %.3g  (n unequal xyz %d / %d)" % (
        S, np.array_equal(c[..., 3:6], cg[..., 3:6]), np.array_equal(c[..., 6:], cg[..., 6:]),
        rel(c[..., :3], cg[..., :3]), int((c[..., :3] != cg[..., :3]).sum()), c[..., :3].size))
    print("  boxes max abs:", float(np.abs(o.all_boxes.cpu().numpy() - g["boxes"]).max()))
    sc = o.all_scores.cpu().numpy()
    for k, nm in enumerate(["iou", "dim", "cor", "combined"]):
        d = np.abs(sc[:, k] - g[nm]); r = d / np.maximum(np.abs(g[nm]), 1e-30)
        print("  %-8s max abs %.3g   max rel (|ref|>1e-3) %.3g" % (nm, d.max(), r[np.abs(g[nm]) > 1e-3].max()))
    print("  best:", o.best_idx.cpu().numpy(), "golden:", g["best"], " score", o.best_score.cpu().numpy())

for nm in ("c1_modeA.npz", "c2_randrot.npz", "c1_modeB.npz"):
    for S in (1, 2, 4, 8):
        print(nm, "cluster", S); run(nm, S)

# planes
g = np.load(os.path.join(G, "planes.npz"))
f = synth.make_frame(1, 480, 640, 5)
pts, other, cnt = api.backproject(f.depth[None].to(dev), f.ground[None].to(dev), f.K[None].to(dev), 5, L.BACKPROJECT_TORCH)
torch.cuda.synchronize()
print("torch flavour n_pts", cnt.cpu().numpy(), "golden", int(g["torch_n_pts"]))
po = O.backproject_torch(f.depth, f.K, f.ground)
npl = int(cnt[0, 0])
print("  points equal oracle:", torch.equal(pts[0, :, :npl].t().cpu(), po))
ids = torch.from_numpy(g["torch_ids"])[None].contiguous().to(dev)
r = api.ransac_plane(pts, cnt[:, 0].contiguous(), ids, thresh=0.05, want_mask=True)
torch.cuda.synchronize()
print("  eq", (-r["eq"][0]).cpu().numpy(), "golden", g["torch_eq"], "count", int(r["count"][0]), int(g["torch_n_inliers"]),
      "best", int(r["best_iter"][0]))
sub = torch.from_numpy(g["caixa_sub"]).t().contiguous()[None].to(dev)
ids = torch.from_numpy(g["caixa_sub_ids"])[None].contiguous().to(dev)
n_pts = torch.tensor([sub.shape[2]], dtype=torch.int32, device=dev)
r = api.ransac_plane(sub, n_pts, ids, thresh=0.01, want_mask=True)
torch.cuda.synchronize()
print("caixa fp64 eq", r["eq"][0].cpu().numpy(), "golden", g["caixa_sub_eq"], "count", int(r["count"][0]), int(g["caixa_sub_n_inliers"]))
inl = torch.where(r["mask"][0] > 0)[0].cpu().numpy().astype(np.int64)
print("  inlier sha equal:", hashlib.sha256(inl.tobytes()).hexdigest() == str(g["caixa_sub_inlier_sha"]))
r = api.ransac_plane(sub.float(), n_pts, ids, thresh=0.01)
print("caixa fp32 eq", (-r["eq"][0]).cpu().numpy(), "golden", g["caixa_sub32_eq"], int(r["count"][0]), int(g["caixa_sub32_n_inliers"]))
# numpy flavour backprojection of frame 0 vs c1 golden
g1 = np.load(os.path.join(G, "c1_modeA.npz"))
f0 = synth.make_frame(0, 480, 640, 5)
pts, other, cnt = api.backproject(f0.depth[None].to(dev), f0.ground[None].to(dev), f0.K[None].to(dev), 5, L.BACKPROJECT_NUMPY)
pg, pn = O.backproject_np(f0.depth.numpy(), f0.K.numpy(), f0.ground.numpy())
print("numpy flavour counts", cnt.cpu().numpy(), len(pg), len(pn), "equal:",
      np.array_equal(pts[0, :, :len(pg)].t().cpu().numpy(), pg), np.array_equal(other[0, :, :len(pn)].t().cpu().numpy(), pn))
ids = torch.from_numpy(g1["sample_ids"])[None].contiguous().to(dev)
r = api.ransac_plane(pts, cnt[:, 0].contiguous(), ids, thresh=0.05)
print("  eq", r["eq"][0].cpu().numpy(), g1["plane_eq"], "normal", r["normal"][0].cpu().numpy(), g1["normal"], int(r["count"][0]), int(g1["n_inliers"]))
