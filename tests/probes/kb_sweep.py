"""One-off stress comparison (not a pytest): K-B (mask -> largest external contour -> minimum-area rectangle) against
cv2 through the oracle on a few thousand random masks of several families and shapes.  Prints one summary line per
shape; the output of the last run is kept in profiles/.    python tests/probes/kb_sweep.py [n_per_shape]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "3dod_b200")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import torch
from scipy import ndimage as ndi

import tdod_oracle as O
from tdod import api

n_per = int(sys.argv[1]) if len(sys.argv) > 1 else 600
dev = torch.device("cuda:0")


def rect_dev(a, b):
    return max(min(float(np.hypot(*(p - q))) for q in b) for p in a)


def family(k, rs, H, W):
    yy, xx = np.mgrid[0:H, 0:W]
    if k == 0:                                                       # smooth blobs, several components
        return ndi.gaussian_filter(rs.randn(H, W), rs.uniform(1.5, 10)) > rs.uniform(-0.05, 0.12)
    if k == 1:                                                       # a few squares, often touching the border
        m = np.zeros((H, W), bool)
        for _ in range(rs.randint(1, 7)):
            x, y, s = rs.randint(-5, W), rs.randint(-5, H), rs.randint(1, max(2, min(H, W) // 3))
            m[max(y, 0):y + s, max(x, 0):x + s] = True
        return m
    if k == 2:                                                       # blob with pin holes
        return (ndi.gaussian_filter(rs.randn(H, W), 5) > 0) & (rs.rand(H, W) < 0.98)
    if k == 3:                                                       # ellipse / ring (+ rival square)
        cx, cy = rs.uniform(0.25, 0.75) * W, rs.uniform(0.25, 0.75) * H
        a, b = rs.uniform(0.05, 0.3) * W, rs.uniform(0.05, 0.3) * H
        m = ((xx - cx) / a) ** 2 + ((yy - cy) / b) ** 2 <= 1
        if rs.rand() < 0.5:
            f = rs.uniform(0.3, 0.9)
            m &= ~(((xx - cx) / (a * f)) ** 2 + ((yy - cy) / (b * f)) ** 2 <= 1)
            if rs.rand() < 0.5:
                s = int(np.sqrt(rs.uniform(0.3, 0.9) * np.pi * a * b))
                m[0:s, 0:s] = True
        return m
    if k == 4:                                                       # thin lines and needles
        m = np.zeros((H, W), bool)
        for _ in range(rs.randint(1, 4)):
            if rs.rand() < 0.5:
                m[rs.randint(0, H), rs.randint(0, W // 2):rs.randint(W // 2, W)] = True
            else:
                m[rs.randint(0, H // 2):rs.randint(H // 2, H), rs.randint(0, W)] = True
        return m
    if k == 5:                                                       # rotated rectangle
        cx, cy, th = rs.uniform(0.3, 0.7) * W, rs.uniform(0.3, 0.7) * H, rs.uniform(0, np.pi)
        u = (xx - cx) * np.cos(th) + (yy - cy) * np.sin(th)
        v = -(xx - cx) * np.sin(th) + (yy - cy) * np.cos(th)
        return (np.abs(u) < rs.uniform(3, 0.3 * W)) & (np.abs(v) < rs.uniform(1, 0.2 * H))
    return rs.rand(H, W) < rs.choice([0.01, 0.08, 0.5])                # speckle


tot = bad = 0
for (H, W) in ((200, 240), (530, 730), (480, 640), (97, 1031), (120, 400), (90, 900), (70, 1024)):
    rs = np.random.RandomState(H + W)
    ms = [family(t % 7, rs, H, W) for t in range(n_per)]
    rects, has = api.mask_rect(torch.from_numpy(np.stack(ms)).to(dev))
    rects, has = rects.cpu().numpy(), has.cpu().numpy()
    nb = exact = 0
    for t, m in enumerate(ms):
        ref = O.mask_rect(m.astype(np.uint8))
        ok = (ref is None) == (has[t] == 0) and has[t] in (0, 1)
        if ok and ref is not None:
            ok = rect_dev(rects[t], ref) < 1e-3
            exact += sorted(map(tuple, rects[t].tolist())) == sorted(map(tuple, ref.tolist()))
        if not ok:
            nb += 1
            print("MISMATCH", (H, W), t, "family", t % 7, "has", has[t], rects[t].tolist(), None if ref is None else ref.tolist())
    print("shape %dx%d: %d masks, %d mismatches, %d bit-equal rectangles" % (H, W, len(ms), nb, exact), flush=True)
    tot += len(ms); bad += nb
print("total %d masks, %d mismatches" % (tot, bad))
