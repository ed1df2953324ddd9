"""Phase clocks of the fused K-B front kernel (library built with -DTDOD_KB_PROF; TDOD_LIB points at it)."""
import ctypes, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "3dod_b200"))
from tdod import api, lib, pipeline, synth
d = torch.device("cuda:0")
b = pipeline.FrameBatch.from_frames(synth.make_frames("C2", range(256)), d)
L = lib.lib()
buf = (ctypes.c_ulonglong * 16)()
for _ in range(3):
    api.mask_rect(b.masks)
L.tdod_debug_kb_prof(buf, 1)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    api.mask_rect(b.masks)
e1.record(); torch.cuda.synchronize()
L.tdod_debug_kb_prof(buf, 0)
names = ["pack: packing", "pack: wait stream free", "pack: wait bytes", "-", "-", "-", "-", "-",
         "-", "-", "-", "ex: labels/winner/rows (3)", "ex: wait pack (4)", "ex: extraction (5)", "ex: labelling (6)", "ex: winner (7)"]
print("ms per call", e0.elapsed_time(e1) / 10)
for i in range(16):
    print("%2d %-30s %9.1f us per CTA per call" % (i, names[i], buf[i] / 10 / 296 / 1965.0))
