"""Do the K-B front kernel and K-A's inlier counting share the SMs?  mask_rect alone, ransac_plane alone, both on two streams."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "3dod_b200"))
from tdod import api, synth
d = torch.device("cuda:0")
c = synth.CONFIGS["C2"]
batch, off, _ = synth.make_batch(range(256), c["H"], c["W"], c["n_obj"], d)
pts, _, counts = api.backproject(batch.depth, batch.ground, batch.K, want_other=False)
n_pts = counts[:, 0].contiguous()
side = torch.cuda.Stream(d)
main = torch.cuda.current_stream(d)
def kb(): return api.mask_rect(batch.masks)
def ka(): return api.ransac_plane(pts, n_pts, seed=3)
def timed(fn, k=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(k): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / k
def both(first_side=True):
    side.wait_stream(main)
    if first_side:
        with torch.cuda.stream(side): r = ka()
        kb()
    else:
        kb()
        with torch.cuda.stream(side): r = ka()
    main.wait_stream(side)
    return r
print("K-B alone %.4f  K-A(ransac) alone %.4f  both (K-A first) %.4f  both (K-B first) %.4f ms"
      % (timed(kb), timed(ka), timed(lambda: both(True)), timed(lambda: both(False))))
# timeline of one co-run (K-B first): events on both streams against a common origin
for rep in range(2):
    torch.cuda.synchronize()
    ev = {k: torch.cuda.Event(enable_timing=True) for k in ("o", "kb1", "ka0", "ka1")}
    side.wait_stream(main)
    ev["o"].record(main)
    kb(); ev["kb1"].record(main)
    with torch.cuda.stream(side):
        ev["ka0"].record(side); ka(); ev["ka1"].record(side)
    main.wait_stream(side)
    torch.cuda.synchronize()
    print("timeline: K-B ends %.3f   K-A starts %.3f ends %.3f" % tuple(ev["o"].elapsed_time(ev[k]) for k in ("kb1", "ka0", "ka1")))
