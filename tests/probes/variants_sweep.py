"""One-off stress comparison (not a pytest): the six ablation samplers (proposals.py:20-336) through the fused kernel's centre
/ dimension modes against the oracle's restatement (itself bit-equal to the unmodified reference, tests/golden/variants.npz)
on many seeded frames, fed the random numbers the oracle consumed.    python tests/probes/variants_sweep.py [n_frames]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (os.path.join(ROOT, "3dod_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np
import torch

import tdod_oracle as O
import test_gpu_parity as T
from tdod import api, synth, lib as L

n_frames = int(sys.argv[1]) if len(sys.argv) > 1 else 15
d = torch.device("cuda:0")
n = 300
tot = {}
for k in range(n_frames):
    cfg = ("C1", "C2", "C4")[k % 3]
    c = synth.CONFIGS[cfg]
    H, W = c["H"], c["W"]
    fid = 9000 + 11 * k
    f = synth.make_frame(fid, H, W, 5)
    N = f.n_obj()
    for name, (cm, dm) in T.VARIANT_MODES.items():
        g = torch.Generator().manual_seed(fid)
        r = O.propose_variant(name, f.boxes, f.depth, f.prior_mean, f.prior_std, f.K, (W, H), n, generator=g)
        noise = r["noise"].permute(1, 0, 2).contiguous().to(d)
        rot = r["rot"].reshape(N, n, 9).contiguous().to(d)
        aux = r["aux"].to(d) if r.get("aux") is not None else None
        o = api.propose_score_select(f.depth[None].to(d), f.K[None].to(d), f.boxes.to(d), f.prior_mean.to(d), f.prior_std.to(d),
                                     torch.zeros(N, dtype=torch.int32, device=d), None, None, n, noise=noise, rot=rot,
                                     propose_only=True, centre_mode=cm, dims_mode=dm, aux=aux, im_shape=(W, H))
        torch.cuda.synchronize()
        got, ref = o.all_cubes.cpu().numpy(), r["cubes"].numpy()
        s = tot.setdefault(name, dict(cubes=0, dims_rot_bad=0, centre_bad=0))
        s["cubes"] += N * n
        s["dims_rot_bad"] += int((got[..., 3:] != ref[..., 3:]).any(-1).sum())
        if cm == L.CENTRE_DEPTH:
            s["centre_bad"] += int((~np.isclose(got[..., :3], ref[..., :3], rtol=1e-5, atol=1e-6)).any(-1).sum())
        else:
            s["centre_bad"] += int((got[..., :3] != ref[..., :3]).any(-1).sum())
    print("frame", k, cfg, flush=True)
print(tot)
