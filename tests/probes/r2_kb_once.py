import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "3dod_b200"))
from tdod import api, pipeline, synth
d = torch.device("cuda:0")
F = int(sys.argv[1]) if len(sys.argv) > 1 else 256
b = pipeline.FrameBatch.from_frames(synth.make_frames("C2", range(F)), d)
for _ in range(3):
    api.mask_rect(b.masks)
torch.cuda.synchronize()
print("ok", b.masks.shape)
