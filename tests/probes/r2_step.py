"""Step time of the C2 hot path (pipeline.select, K-A beside K-B) and the stage times on one stream; TDOD_LIB picks the build."""
import os, sys, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "3dod_b200"))
from tdod import pipeline, synth
d = torch.device("cuda:0")
c = synth.CONFIGS["C2"]
batch, off, _ = synth.make_batch(range(256), c["H"], c["W"], c["n_obj"], d)
def run(k, overlap, marks=None):
    for i in range(k):
        o = pipeline.select(batch, 1000, seed=7, overlap=overlap, marks=None if marks is None else marks[i])
    return o
run(3, True); torch.cuda.synchronize()
res = []
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); o = run(20, True); e1.record(); torch.cuda.synchronize()
    res.append(e0.elapsed_time(e1) / 20)
marks = [dict() for _ in range(10)]
run(10, False, marks); torch.cuda.synchronize()
st = [float(np.mean([m[a].elapsed_time(m[b]) for m in marks])) for a, b in (("ka0", "ka1"), ("kb0", "kb1"), ("kc0", "kc1"))]
print("step ms (3 x 20): %s   serial stages K-A %.3f K-B %.3f K-C %.3f   best idx checksum %d"
      % (" ".join("%.4f" % r for r in res), st[0], st[1], st[2], int(o.best_idx.sum().item())))
# timeline of overlapped steps: stage marks of both streams against the step's first mark
marks = [dict() for _ in range(6)]
run(6, True, marks); torch.cuda.synchronize()
for m in marks[2:5]:
    o0 = m["kb0"]
    print("timeline (ms from kb0): " + "  ".join("%s %.3f" % (k, o0.elapsed_time(m[k])) for k in ("ka0", "ka1", "kb0", "kb1", "kc0", "kc1")))
# one step captured in a CUDA graph (K-A forked onto the auxiliary stream inside the capture), replayed
for ov in (True, False):
    try:
        st = torch.cuda.Stream(d)
        st.wait_stream(torch.cuda.current_stream(d))
        with torch.cuda.stream(st):
            for _ in range(3):
                pipeline.select(batch, 1000, seed=7, overlap=ov)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            og = pipeline.select(batch, 1000, seed=7, overlap=ov)
        for _ in range(3): g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): g.replay()
        e1.record(); torch.cuda.synchronize()
        print("graph overlap=%s: %.4f ms per step, same winners %s" % (ov, e0.elapsed_time(e1) / 20, bool(torch.equal(og.best_idx, o.best_idx))))
    except Exception as ex:
        print("graph overlap=%s failed: %r" % (ov, ex))
# NG graphs of the step replayed round-robin on NG streams: step i+1's K-A / K-B overlap step i's K-C
for NG in (2, 3, 4):
    try:
        g2 = [pipeline.GraphedSelect(batch, 1000, seed=7) for _ in range(NG)]
        ss = [torch.cuda.Stream(d) for _ in range(NG)]
        def pingpong(k):
            cur = torch.cuda.current_stream(d)
            for s_ in ss: s_.wait_stream(cur)
            for i in range(k):
                with torch.cuda.stream(ss[i % NG]):
                    g2[i % NG].replay()
            for s_ in ss: cur.wait_stream(s_)
        pingpong(2 * NG); torch.cuda.synchronize()
        res2 = []
        for rep in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); pingpong(24); e1.record(); torch.cuda.synchronize()
            res2.append(e0.elapsed_time(e1) / 24)
        print("%d graphs, round-robin streams: %s ms per step, same winners %s"
              % (NG, " ".join("%.4f" % r for r in res2), all(bool(torch.equal(g.out.best_idx, o.best_idx)) for g in g2)))
        del g2
    except Exception as ex:
        print("round-robin %d failed: %r" % (NG, ex))
