"""fused against two-kernel K-B on simple block masks over a grid of small shapes (debugging aid)"""
import os, sys, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "3dod_b200"))
from tdod import api
d = torch.device("cuda:0")
for H in (40, 127, 128, 129, 130, 257):
    for W in (32, 33, 47, 48, 63, 64, 65, 96):
        m = np.zeros((3, H, W), bool)
        m[0, (3 * H) // 4:, W // 2:] = True
        m[1, H // 3: H // 2, 2: W - 3] = True
        m[2, H - 1, :] = True
        t = torch.from_numpy(m).to(d)
        r1, h1 = api.mask_rect(t)
        r0, h0 = api.mask_rect(t, two_kernel=True)
        same = torch.equal(h0, h1) and torch.equal(r0.view(torch.int32), r1.view(torch.int32))
        if not same:
            bad = [i for i in range(3) if not torch.equal(r0[i], r1[i]) or h0[i] != h1[i]]
            print("H %d W %d DIFFER masks %s  fused %s  two-kernel %s" % (H, W, bad, r1[bad[0]].flatten().tolist(), r0[bad[0]].flatten().tolist()))
print("done")
