"""End-to-end C2 step from pinned host memory: raw bool masks over PCIe against masks packed on the host cores per chunk."""
import os, sys, time, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "3dod_b200"))
from tdod import pipeline, synth
d = torch.device("cuda:0")
c = synth.CONFIGS["C2"]
batch, off, _ = synth.make_batch(range(256), c["H"], c["W"], c["n_obj"], d)
ref = pipeline.select(batch, 1000, seed=7).best_idx.cpu()
hb = batch.to("cpu").pin()
N = hb.n_obj
del batch
print("cpus", os.cpu_count())
def run(label, **kw):
    pipe = pipeline.Pipeline(c["H"], c["W"], 1000, device=d, seed=7, **kw)
    out = torch.empty((N, 21), dtype=torch.float32).pin_memory()
    for _ in range(2): pipe.run_host(hb, off, out=out)
    t0 = time.perf_counter()
    for _ in range(5): pipe.run_host(hb, off, out=out)
    ms = (time.perf_counter() - t0) * 1e3 / 5
    print("%-40s %7.2f ms per step  %.3e proposals/s  same winners %s" % (label, ms, N * 1000 / (ms * 1e-3), bool(torch.equal(out[:, 20].int(), ref))))
run("bool masks over PCIe, chunk 32", chunk_frames=32)
for rep in range(2):
    for ch in (64, 32, 16):
        run("host_pack, chunk %d, 16 threads" % ch, chunk_frames=ch, host_pack=True)
