"""One-off stress comparison (not a pytest): mode-B scores (score_segmentation / score_mod_segmentation: cv2.convexHull +
cv2.fillPoly + mask IoU; score_point_cloud: points in cube) against the oracle on seeded frames, bit-exact.
    python tests/probes/modeb_sweep.py [n_frames]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (os.path.join(ROOT, "3dod_b200"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import numpy as np
import torch

import tdod_oracle as O
from tdod import api, synth

n_frames = int(sys.argv[1]) if len(sys.argv) > 1 else 30
d = torch.device("cuda:0")
n = 300
tot = dict(frames=0, seg_props=0, seg_bad=0, modseg_bad=0, pc_props=0, pc_bad=0, seg_worst=0.0)
for k in range(n_frames):
    cfg = ("C1", "C2", "C4")[k % 3]
    c = synth.CONFIGS[cfg]
    H, W = c["H"], c["W"]
    fid = 7000 + 13 * k
    f = synth.make_frame(fid, H, W, 4)
    g = torch.Generator().manual_seed(fid)
    o = O.propose(f.boxes, f.depth, f.prior_mean, f.prior_std, f.K, n, normal=torch.tensor([0.02, 0.99, 0.1]), generator=g)
    cubes = o["cubes"].clone()
    cubes[:, ::9, 2] *= 0.25                         # some cubes near / across the camera plane: clamped, huge polygons
    cubes[:, 5::17, 0] += 1.5
    c2d = O.project_corners(O.cube_corners(cubes), f.K, W, H)
    _, pn = O.backproject_np(f.depth.numpy(), f.K.numpy(), f.ground.numpy())
    cloud = torch.from_numpy(pn).to(d)
    for i in range(f.n_obj()):
        for p5, key in ((False, "seg_bad"), (True, "modseg_bad")):
            got = api.score_segmentation(f.masks[i].to(d), c2d[i].contiguous().to(d), p5).cpu().numpy()
            ref = O.segmentation_score(f.masks[i], c2d[i], p5).numpy()
            same = (got == ref) | (np.isnan(got) & np.isnan(ref))
            tot[key] += int((~same).sum())
            if (~same).any():
                tot["seg_worst"] = max(tot["seg_worst"], float(np.nanmax(np.abs(got - ref))))
        tot["seg_props"] += n
        gp = api.score_point_cloud(cloud, cubes[i].contiguous().to(d)).cpu().numpy()
        rp = O.point_cloud_score(torch.from_numpy(pn), cubes[i]).numpy()
        tot["pc_bad"] += int((gp != rp).sum()); tot["pc_props"] += n
    tot["frames"] += 1
    print("frame %d (%s): running totals %s" % (k, cfg, tot), flush=True)
print(tot)
