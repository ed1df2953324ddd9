"""One-off stress comparison (not a pytest): K-A in the torch / fit_parallel flavour (fp32: roi_heads.py:1084-1136,
plane.py:79-134) against the oracle on many seeded frames with identical hypothesis streams: identical points, best
hypothesis, inlier count, and (with and without a ground map) rectified normal.   python tests/probes/plane_sweep.py [n]"""
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (os.path.join(ROOT, "3dod_b200"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import numpy as np
import torch

import tdod_oracle as O
from tdod import api, synth, lib as L

n_frames = int(sys.argv[1]) if len(sys.argv) > 1 else 40
d = torch.device("cuda:0")
tot = dict(frames=0, points_bad=0, best_bad=0, count_bad=0, normal_bad=0)
for k in range(n_frames):
    cfg = ("C1", "C2", "C4", "C5")[k % 4]
    c = synth.CONFIGS[cfg]
    H, W = c["H"], c["W"]
    f = synth.make_frame(11000 + 7 * k, H, W, 3, with_masks=False)
    use_ground = (k % 5) != 4                               # every fifth frame has no ground map (roi_heads.py:1103-1110)
    ground = f.ground if use_ground else None
    pts, _, cnt = api.backproject(f.depth[None].to(d), None if ground is None else ground[None].to(d), f.K[None].to(d), 5,
                                  L.BACKPROJECT_TORCH, want_other=False)
    po = O.backproject_torch(f.depth, f.K, ground)
    tot["points_bad"] += not (int(cnt[0, 0]) == len(po) and torch.equal(pts[0, :, :len(po)].t().cpu(), po))
    ids = O.draw_sample_ids(len(po), 1000, random.Random(k))
    eq, inl, best, _ = O.ransac_parallel(po, ids, 0.05)
    r = api.ransac_plane(pts, cnt[:, 0].contiguous(), torch.from_numpy(ids.astype(np.int32))[None].contiguous().to(d), thresh=0.05)
    tot["best_bad"] += int(r["best_iter"][0]) != int(best)
    tot["count_bad"] += int(r["count"][0]) != len(inl)
    nv = O.rectify_normal(-np.asarray(eq[:3], dtype=np.float64))     # fit_parallel returns the negated equation (plane.py:134)
    tot["normal_bad"] += not np.allclose(r["normal"][0].cpu().numpy(), nv, rtol=0, atol=1e-6)
    tot["frames"] += 1
print(tot)
