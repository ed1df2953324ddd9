"""C1 (one 640x480 frame, 5 objects x 1000 proposals): eager / graph latency of one select; under ncu: the per-kernel list."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "3dod_b200"))
from tdod import pipeline, synth
d = torch.device("cuda:0")
name = sys.argv[1] if len(sys.argv) > 1 else "C1"
c = synth.CONFIGS[name]
fr = synth.make_frames(name, range(1), device=d)
b = pipeline.FrameBatch.from_frames(fr, d)
n = c.get("n_prop", 1000)
for _ in range(5): pipeline.select(b, n, seed=7)
torch.cuda.synchronize()
def timed(fn, k=50):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(k): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / k
print(name, "objects", b.n_obj, "n", n)
print("eager overlap %.4f ms   eager one stream %.4f ms" % (timed(lambda: pipeline.select(b, n, seed=7)), timed(lambda: pipeline.select(b, n, seed=7, overlap=False))))
for ov in (True, False):
    gs = pipeline.GraphedSelect(b, n, seed=7, overlap=ov)
    print("graph overlap=%s %.4f ms" % (ov, timed(gs.replay)))
import numpy as np
marks = [dict() for _ in range(10)]
for i in range(10):
    pipeline.select(b, n, seed=7, overlap=False, marks=marks[i])
torch.cuda.synchronize()
st = [float(np.mean([m[a].elapsed_time(m[c_]) for m in marks])) for a, c_ in (("ka0", "ka1"), ("kb0", "kb1"), ("kc0", "kc1"))]
print("serial stages (eager, one stream): K-A %.3f  K-B %.3f  K-C %.3f ms" % tuple(st))
