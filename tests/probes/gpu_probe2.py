"""Diagnostic: mask_rect / segmentation / point-cloud kernels vs golden fixtures and cv2 on the box."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "3dod_b200")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
from tdod import api, synth
import tdod_oracle as O
from scipy import ndimage as ndi
dev = torch.device("cuda")
G = os.path.join(ROOT, "tests", "golden")

def rect_dev(a, b):
    # set distance between two 4-point sets
    return max(min(float(np.hypot(*(p - q))) for q in b) for p in a)

g = np.load(os.path.join(G, "c1_modeA.npz"))
f = synth.make_frame(0, 480, 640, 5)
rects, has = api.mask_rect(f.masks.to(dev)); torch.cuda.synchronize()
print("c1 has", has.cpu().numpy(), "max dev", max(rect_dev(rects[i].cpu().numpy(), g["rects"][i]) for i in range(5)),
      "bit-equal sets", [sorted(map(tuple, rects[i].cpu().numpy().tolist())) == sorted(map(tuple, g["rects"][i].tolist())) for i in range(5)])
gr = np.load(os.path.join(G, "real_masks.npz"))
for nm in ("mask0", "mask1"):
    shp = tuple(gr[nm + "_shape"]); m = np.unpackbits(gr[nm + "_bits"])[:shp[0] * shp[1]].reshape(shp).astype(bool)
    r, h = api.mask_rect(torch.from_numpy(m)[None].to(dev)); torch.cuda.synchronize()
    print(nm, shp, "has", int(h[0]), "dev", rect_dev(r[0].cpu().numpy(), gr[nm + "_rect"]))
# random masks vs cv2
rs = np.random.RandomState(0)
stats = {"n": 0, "bit": 0, "close": 0, "far": 0, "has_mismatch": 0, "codes": {}}
H, W = 200, 240
ms = []
for t in range(300):
    k = t % 5
    if k == 0: m = rs.rand(H, W) < rs.uniform(0.02, 0.7)
    elif k == 1: m = ndi.gaussian_filter(rs.randn(H, W), rs.uniform(1, 8)) > rs.uniform(-0.05, 0.1)
    elif k == 2:
        m = np.zeros((H, W), bool)
        for _ in range(rs.randint(1, 6)):
            x, y = rs.randint(0, W), rs.randint(0, H); s = rs.randint(1, 14); m[y:y + s, x:x + s] = True
    elif k == 3:
        m = ndi.gaussian_filter(rs.randn(H, W), 4) > 0; m &= rs.rand(H, W) < 0.97
    else:
        yy, xx = np.mgrid[0:H, 0:W]; cx, cy = rs.uniform(60, 180), rs.uniform(60, 140); a, b = rs.uniform(10, 60, 2)
        m = ((xx - cx) / a) ** 2 + ((yy - cy) / b) ** 2 <= 1
        m &= ~(((xx - cx) / (a * 0.5)) ** 2 + ((yy - cy) / (b * 0.5)) ** 2 <= 1) if t % 2 else m   # ring with an island sometimes
        if t % 4 == 0: m[int(cy) - 2:int(cy) + 3, int(cx) - 2:int(cx) + 3] = True
    ms.append(m)
mt = torch.from_numpy(np.stack(ms)).to(dev)
t0 = time.time(); rects, has = api.mask_rect(mt); torch.cuda.synchronize(); print("300 masks %.1f ms" % ((time.time() - t0) * 1e3))
rects = rects.cpu().numpy(); has = has.cpu().numpy()
for t, m in enumerate(ms):
    ref = O.mask_rect(m.astype(np.uint8))
    stats["n"] += 1
    stats["codes"][int(has[t])] = stats["codes"].get(int(has[t]), 0) + 1
    if has[t] == 3: continue
    if (ref is None) != (has[t] == 0): stats["has_mismatch"] += 1; continue
    if ref is None: continue
    d = rect_dev(rects[t], ref)
    if sorted(map(tuple, rects[t].tolist())) == sorted(map(tuple, ref.tolist())): stats["bit"] += 1
    elif d < 1e-3: stats["close"] += 1
    else:
        stats["far"] += 1
        if stats["far"] < 6: print("FAR", t, t % 5, d, rects[t].tolist(), ref.tolist())
print("random masks:", stats)

# segmentation + point cloud vs golden mode B
gb = np.load(os.path.join(G, "c1_modeB.npz"))
cubes = torch.from_numpy(gb["cubes"])
c2d, _ = api.project_cubes(cubes.to(dev), f.K.to(dev), 640, 480)
c2d_o = O.project_corners(O.cube_corners(cubes), f.K, 640, 480)
print("corners2d max abs vs oracle", (c2d.cpu() - c2d_o).abs().max().item())
for i in range(5):
    for p5, nm in ((False, "seg_%d"), (True, "modseg_%d")):
        # feed the oracle's corners so that the comparison is about the raster only
        s = api.score_segmentation(f.masks[i].to(dev), c2d_o[i].contiguous().to(dev), p5).cpu().numpy()
        ref = gb[nm % i]
        print(nm % i, "equal", np.array_equal(s, ref), "n diff", int((s != ref).sum()), "max abs", float(np.abs(s - ref).max()))
pg, pn = O.backproject_np(f.depth.numpy(), f.K.numpy(), f.ground.numpy())
pts = torch.from_numpy(pn).to(dev)
for i in range(5):
    pc = api.score_point_cloud(pts, cubes[i].contiguous().to(dev)).cpu().numpy()
    print("pc", i, "equal", np.array_equal(pc, gb["pc"][i]), "n diff", int((pc != gb["pc"][i]).sum()))
pc32 = api.score_point_cloud(pts.float(), cubes[0].contiguous().to(dev)).cpu().numpy()
print("pc fp32 points vs oracle:", np.array_equal(pc32, O.point_cloud_score(torch.from_numpy(pn).float(), cubes[0]).numpy()))
