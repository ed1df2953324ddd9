"""Round-2 probe 3: the fused K-B path (tdod_mask_rect) against the two-kernel path (pack + rectangle passes) on the C2
batch, its timing, and against cv2 on the kb_sweep families."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (os.path.join(ROOT, "3dod_b200"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
from tdod import api, pipeline, synth  # noqa: E402

d = torch.device("cuda:0")
F = int(sys.argv[1]) if len(sys.argv) > 1 else 256
frames = synth.make_frames("C2", range(F), device=d) if "device" in synth.make_frames.__code__.co_varnames else synth.make_frames("C2", range(F))
b = pipeline.FrameBatch.from_frames(frames, d)
del frames
masks = b.masks
print("masks", tuple(masks.shape))
r_old, h_old = api.mask_rect(masks, two_kernel=True)
r_new, h_new = api.mask_rect(masks)
torch.cuda.synchronize()
same_h = torch.equal(h_old, h_new)
has1 = h_old == 1
same_r = torch.equal(r_old[has1].view(torch.int32), r_new[has1].view(torch.int32))
print("state bytes equal:", same_h, " rectangles bit-equal:", same_r, " codes:", torch.bincount(h_new.long(), minlength=5).tolist())
if not (same_h and same_r):
    bad = (~((r_old == r_new).flatten(1).all(1))) & has1
    idx = torch.nonzero(bad).flatten()[:10].tolist()
    print("first differing masks", idx, [(h_old[i].item(), h_new[i].item()) for i in idx])
    for i in idx[:3]:
        print(r_old[i].tolist(), r_new[i].tolist())
for name, kw in (("two-kernel", dict(two_kernel=True)), ("fused", dict())):
    for _ in range(3):
        api.mask_rect(masks, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        api.mask_rect(masks, **kw)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print("%s: %.4f ms per %d masks (%.1f GB/s of mask bytes)" % (name, ms, masks.shape[0], masks.numel() / ms / 1e6))
