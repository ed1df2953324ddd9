"""One-off stress comparison (not a pytest): the whole path (K-A plane -> yaw table, K-B rectangle, fused K-C) against the
oracle chain on many seeded frames in parity-RNG mode, with the gates of tests/test_gpu_parity.py::test_selection_parity_*.
    python tests/probes/select_sweep.py [n_frames]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (os.path.join(ROOT, "3dod_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np
import torch

import test_gpu_parity as T
from tdod import pipeline, synth

n_frames = int(sys.argv[1]) if len(sys.argv) > 1 else 24
d = torch.device("cuda:0")
stats = dict(frames=0, objects=0, proposals=0, plane_bad=0, score_bad=0, worst_rel=0.0, argmax_checked=0, argmax_bad=0, nan_objects=0,
             nan_bad=0)
t0 = time.time()
for k in range(n_frames):
    cfg = ("C1", "C2", "C4")[k % 3]
    c = synth.CONFIGS[cfg]
    H, W, n = c["H"], c["W"], 1000
    fid = 5000 + 37 * k
    f = synth.make_frame(fid, H, W, c["n_obj"])
    r = T._oracle_chain(f, n, W, H, seed=fid)
    batch = pipeline.FrameBatch.from_frames([f], d)
    sel = pipeline.select(batch, n, noise=r["noise"].permute(1, 0, 2).contiguous().to(d), yaw_idx=r["yaw_idx"].to(torch.int16).to(d),
                          sample_ids=torch.from_numpy(r["ids"].astype(np.int32))[None].contiguous().to(d), return_all=True)
    torch.cuda.synchronize()
    ok_plane = int(sel.inlier_count[0]) == r["n_inl"] and np.allclose(sel.plane_eq[0].cpu().numpy(), r["eq"], rtol=1e-13, atol=0) \
        and np.allclose(sel.normal[0].cpu().numpy(), r["normal"], rtol=0, atol=1e-12)
    stats["plane_bad"] += not ok_plane
    comb = sel.all.all_scores[:, 3].cpu().numpy().astype(np.float64)
    ref = r["comb"].astype(np.float64)
    nanpat = np.array_equal(np.isnan(comb), np.isnan(ref))
    err = np.abs(comb - ref) - (1e-6 + 2e-5 * np.abs(ref))
    bad = int((err > 0).sum()) if nanpat else comb.size
    stats["score_bad"] += bad
    with np.errstate(invalid="ignore", divide="ignore"):
        rel = np.nanmax(np.abs(comb - ref) / np.maximum(np.abs(ref), 1e-3))
    stats["worst_rel"] = max(stats["worst_rel"], float(rel))
    got = sel.best_idx.cpu().tolist()
    for i in range(f.n_obj()):
        rr = r["comb"][i]
        if np.isnan(rr).any():
            stats["nan_objects"] += 1
            stats["nan_bad"] += got[i] != int(np.argmax(np.isnan(rr)))
            continue
        top2 = np.sort(rr)[-2:]
        if top2[1] - top2[0] > 2e-5 * abs(top2[1]):
            stats["argmax_checked"] += 1
            stats["argmax_bad"] += got[i] != r["best"][i]
    stats["frames"] += 1; stats["objects"] += f.n_obj(); stats["proposals"] += f.n_obj() * n
    print("frame %d (%s, id %d): %d objects, plane %s, scores outside tolerance %d" % (k, cfg, fid, f.n_obj(), "ok" if ok_plane else "BAD", bad), flush=True)
stats["seconds"] = round(time.time() - t0, 1)
print(stats)
