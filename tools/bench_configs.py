"""Secondary measurements: the BASELINE.json configs that are not the bench line (bench.py measures configs[1]).

    python tools/bench_configs.py [C1] [C3] [C5] [--steps K] [--frames F]

One JSON line per config, device-resident inputs, CUDA events on the launching stream:
  C1  one 640x480 frame, 5 objects x 1000 proposals (the reference's CPU-runnable case)
  C3  one 1024x768 frame, 50 objects x 10 000 proposals (8-CTA cluster path of the fused kernel)
  C5  back-projection + RANSAC alone at 1920x1080 (K-A only; value in frames/s); both flavours
  WL  weak-loss geometry, forward + backward for the cubes of 8 C2-shaped frames
Algorithmic bytes follow SURVEY.md 8(d)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "3dod_b200"))

import numpy as np  # noqa: E402
import torch  # noqa: E402

from tdod import api, pipeline, synth, lib as L  # noqa: E402


def peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    return float(json.load(open(p))["hbm_gbs"]) if os.path.exists(p) else 6650.0


def flush_l2(buf):
    buf.zero_()      # 512 MB > 126 MB L2


def timed(fn, steps, flush):
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    for a, b in ev:
        flush_l2(flush)
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    return float(np.mean([a.elapsed_time(b) for a, b in ev]))


def run_select(name, steps, dev, flush, frames=None):
    c = synth.CONFIGS[name]
    F = frames or c["frames"]
    fr = [synth.make_frame(i, c["H"], c["W"], c["n_obj"], device=dev) for i in range(F)]
    batch = pipeline.FrameBatch.from_frames(fr, dev)
    n = c["n_prop"]
    N = batch.n_obj
    marks = {}
    ms = timed(lambda: pipeline.select(batch, n, seed=7, overlap=False, marks=marks), steps, flush)
    torch.cuda.synchronize()
    st = {"K-A": marks["ka0"].elapsed_time(marks["ka1"]), "K-B": marks["kb0"].elapsed_time(marks["kb1"]),
          "K-C": marks["kc0"].elapsed_time(marks["kc1"])}
    H, W = c["H"], c["W"]
    P = ((H + 4) // 5) * ((W + 4) // 5)
    ab = {"K-A": F * (5 * P + 64), "K-B": N * (H * W + 33), "K-C": N * (4 * n + 160)}
    tot = sum(ab.values())
    pk = peak()
    return {"config": name, "frames": F, "objects": N, "n_prop": n, "ms_per_step": ms,
            "value": N * n / (ms * 1e-3), "unit": "proposals/s", "stages_ms_last_step": st,
            "algorithmic_bytes": ab, "hbm_frac_step": tot / (ms * 1e-3) / 1e9 / pk,
            "hbm_frac_stage": {k: ab[k] / (st[k] * 1e-3) / 1e9 / pk for k in ab},
            "l2": "512 MB buffer rewritten between timed iterations"}


def run_mode_b(steps, dev, flush):
    """C1 shape, mode B scores (MABO branch, roi_heads.py:512-582): segmentation / mod-segmentation IoU and points-in-cube
    for 5 objects x 1000 proposals through the per-function entry points the reference's call sites use"""
    c = synth.CONFIGS["C1"]
    H, W, n = c["H"], c["W"], c["n_prop"]
    f = synth.make_frame(0, H, W, c["n_obj"], device=dev)
    batch = pipeline.FrameBatch.from_frames([f], dev)
    sel = pipeline.select(batch, n, seed=7, return_all=True)
    pts, other, cnt = api.backproject(batch.depth, batch.ground, batch.K, 5, L.BACKPROJECT_NUMPY, want_other=True)
    torch.cuda.synchronize()
    P_other = int(cnt[0, 1])
    cloud = other[0, :, :P_other].t().contiguous()
    cubes = sel.all.all_cubes
    N = cubes.shape[0]
    c2d = [api.project_cubes(cubes[i], batch.K[0], W, H, boxes=False)[0] for i in range(N)]

    def seg():
        for i in range(N):
            api.score_segmentation(batch.masks[i], c2d[i])

    def modseg():
        for i in range(N):
            api.score_segmentation(batch.masks[i], c2d[i], power5=True)

    def pc():
        for i in range(N):
            api.score_point_cloud(cloud, cubes[i])
    out = {"config": "C1 mode B", "objects": N, "n_prop": n, "points": P_other}
    for nm, fn in (("score_segmentation", seg), ("score_mod_segmentation", modseg), ("score_point_cloud", pc)):
        ms = timed(fn, steps, flush)
        out[nm + "_ms"] = ms
        out[nm + "_proposals_per_s"] = N * n / (ms * 1e-3)
    out["pc_compares_per_s"] = 6.0 * P_other * N * n / (out["score_point_cloud_ms"] * 1e-3)
    return out


def run_weak_loss(steps, dev, flush, n_img=8):
    """SURVEY 8(f) row 4 at the C2 frame shape: forward + backward of the weak-loss geometry for the cubes of n_img frames
    (project_boxes + GIoU + z_loss + focal segment_loss).  The cubes are proposals of the product path itself; the
    oracle stays test infrastructure (its CPU time for the same losses: `cpu_oracle` in profiles/r01_e_bench_configs.jsonl,
    recorded once from a test session)."""
    from tdod import weak_loss as WL
    c = synth.CONFIGS["C2"]
    H, W = c["H"], c["W"]
    fr = [synth.make_frame(i, H, W, c["n_obj"], device=dev) for i in range(n_img)]
    batch = pipeline.FrameBatch.from_frames(fr, dev)
    sel = pipeline.select(batch, 8, seed=3, return_all=True)
    dc = sel.all.all_cubes[:, 0].contiguous()
    N = dc.shape[0]
    dK = batch.K[batch.obj_frame.long()].contiguous()
    db, dm = batch.boxes, batch.masks
    im_t = torch.tensor([(W, H)] * N, dtype=torch.int32, device=dev)
    idx_t = torch.arange(N, dtype=torch.int32, device=dev)

    def step():
        x = dc.clone().requires_grad_(True)
        o = WL.project_boxes(x, dK, im_t, gt_boxes=db)
        seg = WL.segment_loss(dm, o["seg_corners"], idx_t)
        z = WL.z_loss(db, x, dK, im_t, o["boxes"])
        (o["loss_iou"].mean() + seg.mean() + z.mean()).backward()
        return x.grad
    ms = timed(step, steps, flush)
    ab = 2 * N * H * W                                   # every mask read once forward and once backward
    return {"config": "weak-loss geometry (8f-4), C2 frame shape", "images": n_img, "cubes": N, "ms_per_step": ms,
            "value": N / (ms * 1e-3), "unit": "cubes/s (forward + backward)", "algorithmic_bytes": ab,
            "hbm_frac_step": ab / (ms * 1e-3) / 1e9 / peak(), "launches_per_step": "8 libtdod kernels + torch glue",
            "l2": "512 MB buffer rewritten between timed iterations"}


def run_c5(steps, dev, flush, frames=None):
    c = synth.CONFIGS["C5"]
    F = frames or c["frames"]
    H, W = c["H"], c["W"]
    depth = torch.empty((F, H, W), dtype=torch.float32, device=dev)
    ground = torch.empty((F, H, W), dtype=torch.bool, device=dev)
    K = torch.empty((F, 3, 3), dtype=torch.float32, device=dev)
    for i in range(F):
        f = synth.make_frame(i, H, W, 6, device=dev, with_masks=False)
        depth[i], ground[i], K[i] = f.depth, f.ground, f.K
    P = ((H + 4) // 5) * ((W + 4) // 5)
    pk = peak()
    out = []
    for flavour, nm in ((L.BACKPROJECT_NUMPY, "numpy/pyransac3d fp64"), (L.BACKPROJECT_TORCH, "torch fit_parallel fp32")):
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]

        def step():
            e[0].record()
            pts, _, cnt = api.backproject(depth, ground, K, 5, flavour, want_other=False)
            e[1].record()
            r = api.ransac_plane(pts, cnt[:, 0].contiguous(), None, seed=7, iters=1000, thresh=0.05)
            e[2].record()
            return r
        ms = timed(step, steps, flush)
        torch.cuda.synchronize()
        bp, rs = e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2])
        strict = F * (5 * P + 16)
        sector = F * ((H + 4) // 5) * W * 5            # every 5th row fully touched (32-byte sectors)
        out.append({"config": "C5", "flavour": nm, "frames": F, "ms_per_step": ms, "value": F / (ms * 1e-3),
                    "unit": "frames/s", "backproject_ms": bp, "ransac_ms": rs,
                    "algorithmic_bytes_strict": strict, "bytes_sector_granular": sector,
                    "backproject_hbm_frac_strict": strict / (bp * 1e-3) / 1e9 / pk,
                    "backproject_hbm_frac_sector": sector / (bp * 1e-3) / 1e9 / pk,
                    "ransac_bound": "FP32 FMA issue (1000 x P_ground x 4 instructions per point-hypothesis pair)"})
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("configs", nargs="*", default=["C1", "C1B", "C3", "C5", "WL"])
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--frames", type=int, default=None)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    L.lib()
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    for name in a.configs:
        if name == "C1B":
            print(json.dumps(run_mode_b(a.steps, dev, flush)), flush=True)
        elif name == "WL":
            print(json.dumps(run_weak_loss(a.steps, dev, flush)), flush=True)
        elif name == "C5":
            for line in run_c5(a.steps, dev, flush, a.frames):
                print(json.dumps(line), flush=True)
        else:
            print(json.dumps(run_select(name, a.steps, dev, flush, a.frames)), flush=True)


if __name__ == "__main__":
    main()
