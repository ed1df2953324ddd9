"""bench.py -- scored cube proposals / second (BASELINE.json `metric`) on BASELINE.json configs[1]:
an Omni3D / SUN-RGBD-shaped batch of 256 synthetic frames 730x530, ~15 objects per frame x 1000 proposals.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--frames F]

A step = one pass of the hot path over the batch: K-A (back-projection + RANSAC ground plane + yaw table),
K-B (mask -> min-area rectangle, fused front + tail kernels) and the fused K-C (propose + project + score + select).
  value     device-resident inputs, CUDA events on the launching stream, max over ranks (weak scaling: every rank
            owns its own 256 frames)
  e2e       the same batch through tdod.pipeline.Pipeline.run_host: pinned HOST inputs (bool masks), H2D copies and the
            D2H read of the winners inside the timed region; e2e.optional_packed_masks: host masks at 1 bit per pixel
  roofline  the WHOLE STEP against the measured HBM peak (algorithmic bytes of all stages / step time); the longest stage
            is named with its own fraction and binding resource, every stage is listed
  value_parity_rng  the proposal set defined by the reference: torch's generators in its draw order inside the timed region
  configs   BASELINE configs 0, 2, 3, 4 as sub-records (C1, C3, C4 strong-scaled over the ranks, C5) + mode B
  cpu_baseline  the oracle (CPU restatement of the reference's torch/NumPy/OpenCV/SciPy path) on a bounded sample
`--impl reference` times that CPU path on all host cores (the reference is Python: nothing to compile; the
oracle port is the only form in which it travels to the GPU box)."""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))

# The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner to fd 1 when
# NCCL_DEBUG is set on the box), so fd 1 is pointed at stderr for the whole process and the line goes to the saved fd.
_JSON_OUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)
for p in (os.path.join(ROOT, "3dod_b200"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402
import torch  # noqa: E402

WORKLOAD = "C2"
N_PROP = 1000
METRIC = "scored cube proposals/sec"
UNIT = "proposals/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="tdod", choices=["tdod", "reference"])
    ap.add_argument("--frames", type=int, default=256, help="frames per GPU (BASELINE configs[1]: 256)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the C1 / C3 / C5 / mode-B / C4 sub-records")
    ap.add_argument("--no-parity-rng", action="store_true", help="skip the reference-order RNG figure")
    ap.add_argument("--no-graph", action="store_true", help="time eager launches instead of CUDA-graph replays of the step")
    ap.add_argument("--graph-depth", type=int, default=2, help="copies of the step graph replayed round-robin (1: steps back to back)")
    ap.add_argument("--c4-frames", type=int, default=10000, help="frames of BASELINE configs[3] (all ranks together)")
    ap.add_argument("--profile", action="store_true",
                    help="cudaProfilerStart/Stop around the timed region (ncu --profile-from-start off)")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ----------------------------------------------------------------------------------------------------------------------
# clocks during the timed region
# ----------------------------------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock and throttle reasons every ~5 ms through NVML (a `nvidia-smi` process per sample takes longer than the
    whole timed region); falls back to polling nvidia-smi when pynvml is missing."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop, self.t = index, [], threading.Event(), None
        self.nv = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(self._physical_index(index))
            self.max_sm = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            pynvml.nvmlDeviceGetClockInfo(self.h, pynvml.NVML_CLOCK_SM)      # first queries are slow: not inside the timed region
            if hasattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons"):
                pynvml.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        except Exception:
            self.nv = None

    @staticmethod
    def _physical_index(index):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            parts = [p for p in vis.split(",") if p.strip() != ""]
            if index < len(parts) and parts[index].strip().isdigit():
                return int(parts[index])
        return index

    def _run(self):
        nv = self.nv
        while not self.stop.is_set():
            try:
                if nv is not None:
                    sm = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                    r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                        else int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                    self.rows.append((sm, self.max_sm, r))
                else:
                    out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                    c = [x.strip() for x in out.strip().split(",")]
                    bits = sum((1 << k) for k in range(4) if c[2 + k].lower().startswith("active"))
                    self.rows.append((float(c[0]), float(c[1]), -bits))     # negative: nvidia-smi flag order
            except Exception:
                pass
            self.stop.wait(0.005 if nv is not None else 0.1)

    def __enter__(self):
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.t.join(timeout=6)

    def summary(self):
        sm = [r[0] for r in self.rows]
        mx = [r[1] for r in self.rows]
        reasons = []
        if self.nv is not None:
            nv = self.nv
            names = (("hw_slowdown", "nvmlClocksThrottleReasonHwSlowdown"), ("hw_thermal_slowdown", "nvmlClocksThrottleReasonHwThermalSlowdown"),
                     ("sw_thermal_slowdown", "nvmlClocksThrottleReasonSwThermalSlowdown"), ("sw_power_cap", "nvmlClocksThrottleReasonSwPowerCap"))
            for n, attr in names:
                bit = getattr(nv, attr, None)
                if bit is not None and any(r[2] >= 0 and (r[2] & bit) for r in self.rows):
                    reasons.append(n)
        else:
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            reasons = [n for k, n in enumerate(names) if any((-r[2]) & (1 << k) for r in self.rows if r[2] <= 0)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "source": "nvml, 5 ms" if self.nv is not None else "nvidia-smi, 100 ms"}


# ----------------------------------------------------------------------------------------------------------------------
# CPU path: the oracle run the way ROIHeads_Boxer._forward_cube runs the reference (roi_heads.py:345-505), one frame
# ----------------------------------------------------------------------------------------------------------------------
_FRAMES = {}      # frame_id -> synth.Frame, filled before the worker pool forks


def _cpu_frame_timed(frame_id):
    """synthetic-frame generation is not part of the path: time only the oracle calls"""
    import random
    import tdod_oracle as O
    from tdod import synth
    torch.set_num_threads(1)
    c = synth.CONFIGS[WORKLOAD]
    f = _FRAMES.get(frame_id)
    if f is None:
        f = synth.make_frame(frame_id, c["H"], c["W"], c["n_obj"])
    t0 = time.perf_counter()
    pg, _ = O.backproject_np(f.depth.numpy(), f.K.numpy(), f.ground.numpy())
    ids = O.draw_sample_ids(len(pg), 1000, random.Random(frame_id))
    eq, _, _ = O.ransac_np(pg, ids, 0.05)
    nv = O.rectify_normal(np.array(eq[:3]))
    o = O.propose(f.boxes, f.depth, f.prior_mean, f.prior_std, f.K, N_PROP, torch.from_numpy(nv),
                  torch.Generator().manual_seed(frame_id))
    c2d = O.project_corners(O.cube_corners(o["cubes"]), f.K, c["W"], c["H"])
    pb = O.boxes_from_corners(c2d)
    for i in range(f.n_obj()):
        s_iou = O.iou_score(f.boxes[i], pb[i])
        s_dim = O.dims_score(f.prior_mean[i], f.prior_std[i], o["cubes"][i, :, 3:6], f.boxes[i], pb[i])
        s_cor = O.corners_score(f.masks[i], c2d[i])
        O.select(s_iou, s_dim, s_cor)
    return f.n_obj() * N_PROP, time.perf_counter() - t0


def cpu_baseline_single(budget_s=15.0):
    """rank 0, N=1: the scalar port on one core, frames 0.. until ~budget_s of CPU work"""
    torch.set_num_threads(1)
    props, secs, k = 0, 0.0, 0
    while secs < budget_s and k < 64:
        p, t = _cpu_frame_timed(k)
        props += p; secs += t; k += 1
    return {"value": props / secs, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": "%d frames of %s (%d proposals), oracle/tdod_oracle.py single-threaded, %.1f s" % (k, WORKLOAD, props, secs)}


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (its oracle port: the reference is Python and
    cannot travel to the GPU box) on every host core, frames fanned out over a process pool."""
    import multiprocessing as mp
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    from tdod import synth
    c = synth.CONFIGS[WORKLOAD]
    # size the per-step sample from one probe frame so that warmup + steps end within ~2.5 minutes
    _, t_frame = _cpu_frame_timed(0)
    budget = 150.0
    rounds = max(1, int(budget * cores / max(t_frame, 1e-3) / (args.steps + args.warmup) / cores))
    rounds = min(rounds, max(1, 256 // cores))
    frames_per_step = min(256, rounds * cores)
    for k in range(frames_per_step):
        _FRAMES[k] = synth.make_frame(k, c["H"], c["W"], c["n_obj"])
    ctx = mp.get_context("fork")
    step_ms, props_step = [], 0
    with ctx.Pool(cores) as pool:
        for s in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            res = pool.map(_cpu_frame_timed, range(frames_per_step), chunksize=1)
            dt = time.perf_counter() - t0
            if s >= args.warmup:
                step_ms.append(dt * 1e3)
                props_step = sum(r[0] for r in res)
    total_s = sum(step_ms) / 1e3
    value = props_step * args.steps / total_s
    sample = "%d frames of %s per step (%d proposals) over %d worker processes" % (frames_per_step, WORKLOAD, props_step, cores)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": float(np.mean(step_ms)), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "BASELINE configs[1]: 730x530 frames, ~15 objects/frame x 1000 proposals, mode A "
                                   "(IoU x dims x corners + argmax) incl. back-projection + RANSAC; bounded sample", "frames": frames_per_step},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), file=_JSON_OUT, flush=True)


# ----------------------------------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------------------------------
# what binds each stage; ncu evidence under profiles/r02_*
BOUND = {"K-A backproject+ransac": "fp32 fma issue (hypothesis scoring); overlapped with K-B on a second stream",
         "K-B mask_rect": "hbm (every mask byte once; bit stream, runs and labels stay in shared memory)",
         "K-C propose_score_select": "fp32 / sfu issue (ncu: issue-active 65 %)"}
# dram__bytes_read.sum + dram__bytes_write.sum per launch, `ncu --set full` of this workload (profiles/r02_b_ncu_kb_raw.csv):
# kb_front_kernel 1.4881 GB read + 0.0067 GB written, kb_tail_kernel 0.0034 GB
TRAFFIC = {"K-B mask_rect": 1.4881e9 + 0.0067e9 + 0.0034e9}


def algorithmic_bytes(batch, n_prop, stride=5):
    """compulsory HBM traffic per stage for one pass over `batch` (every needed input element read
    once, every required output written once)."""
    F, H, W = batch.depth.shape
    N = batch.n_obj
    P = ((H + stride - 1) // stride) * ((W + stride - 1) // stride)
    ka = F * (P * 4 + P * 1 + 36 + 12 + 16)                       # strided depth + ground samples, K, normal, plane
    kb = N * (H * W + 33)                                         # every mask pixel once; rectangle + flag out
    kc = N * (4 * n_prop + 44 + 32 + 84)                          # diagonal depth gather, box/priors, rect, winner record
    return {"K-A backproject+ransac": ka, "K-B mask_rect": kb, "K-C propose_score_select": kc}


def _events(n):
    return [torch.cuda.Event(enable_timing=True) for _ in range(n)]


def bench_small_configs(dev, steps, peak):
    """BASELINE configs 0, 2 and 4 (C1, C3, C5) on this GPU, device-resident, L2 flushed between iterations (their inputs
    fit in L2).  C1 also replayed as a CUDA graph: ten launches of microseconds each are launch-bound otherwise."""
    from tdod import api, pipeline, synth, lib as L
    out = {}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def timed(fn, k):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(dev)
        tot = 0.0
        for _ in range(k):
            flush.zero_()
            a, b = _events(2)
            a.record(); fn(); b.record()
            torch.cuda.synchronize(dev)
            tot += a.elapsed_time(b)
        return tot / k
    for name in ("C1", "C3"):
        c = synth.CONFIGS[name]
        fr = [synth.make_frame(i, c["H"], c["W"], c["n_obj"], device=dev) for i in range(c["frames"])]
        b = pipeline.FrameBatch.from_frames(fr, dev)
        n = c["n_prop"]
        ms = timed(lambda: pipeline.select(b, n, seed=7), steps)
        ab = sum(algorithmic_bytes(b, n).values())
        rec = {"frames": c["frames"], "objects": b.n_obj, "n_prop": n, "ms_per_step": ms, "value": b.n_obj * n / (ms * 1e-3),
               "unit": UNIT, "hbm_frac_step": ab / (ms * 1e-3) / 1e9 / peak,
               "bound": "launch latency (11 launches)" if name == "C1" else "fp32 issue in K-C (400 CTAs, one wave)"}
        if name == "C1":
            try:
                gs1 = pipeline.GraphedSelect(b, n, seed=7)          # one step as one graph, K-A forked inside the capture
                ms_g = timed(gs1.replay, steps)
                sg = gs1.out
                ref = pipeline.select(b, n, seed=7)
                torch.cuda.synchronize(dev)
                rec["cuda_graph_ms_per_step"] = ms_g
                rec["cuda_graph_value"] = b.n_obj * n / (ms_g * 1e-3)
                rec["cuda_graph_same_winners"] = bool(torch.equal(ref.best_idx, sg.best_idx))
                del gs1
            except Exception as e:                                  # a capture problem must not cost the bench line
                rec["cuda_graph_error"] = repr(e)[:200]
        out[name] = rec
        del fr, b
    # C5: K-A alone, 1080p
    c = synth.CONFIGS["C5"]
    F = 64
    H, W = c["H"], c["W"]
    depth = torch.empty((F, H, W), dtype=torch.float32, device=dev)
    ground = torch.empty((F, H, W), dtype=torch.bool, device=dev)
    K = torch.empty((F, 3, 3), dtype=torch.float32, device=dev)
    for i in range(F):
        f = synth.make_frame(i, H, W, 6, device=dev, with_masks=False)
        depth[i], ground[i], K[i] = f.depth, f.ground, f.K
    ev = _events(3)

    def ka():
        ev[0].record()
        pts, _, cnt = api.backproject(depth, ground, K, 5, L.BACKPROJECT_NUMPY, want_other=False)
        ev[1].record()
        api.ransac_plane(pts, cnt[:, 0].contiguous(), None, seed=7, iters=1000, thresh=0.05)
        ev[2].record()
    ms = timed(ka, steps)
    P = ((H + 4) // 5) * ((W + 4) // 5)
    out["C5"] = {"frames": F, "of": c["frames"], "ms_per_step": ms, "value": F / (ms * 1e-3), "unit": "frames/s",
                 "backproject_ms": ev[0].elapsed_time(ev[1]), "ransac_ms": ev[1].elapsed_time(ev[2]),
                 "hbm_frac_strict": F * (5 * P + 16) / (ms * 1e-3) / 1e9 / peak,
                 "bound": "fp32 fma issue (1000 hypotheses x P points); the back-projection half alone is sector-bound"}
    return out


def bench_mode_b(dev, peak, frames=8):
    """mode B (MABO branch): all six scores + IoU3D + K-D curves for a few C2-shaped frames"""
    from tdod import mabo, pipeline, synth
    c = synth.CONFIGS["C2"]
    fr = synth.make_frames("C2", range(frames), device=dev)
    b = pipeline.FrameBatch.from_frames(fr, dev)
    gt = pipeline.select(b, N_PROP, seed=99).best_cube.clone()
    for _ in range(2):
        mabo.mabo(b, gt, N_PROP, seed=5)
    torch.cuda.synchronize(dev)
    a, e = _events(2)
    a.record()
    r = mabo.mabo(b, gt, N_PROP, seed=5)
    e.record()
    torch.cuda.synchronize(dev)
    ms = a.elapsed_time(e)
    return {"frames": frames, "objects": b.n_obj, "n_prop": N_PROP, "ms_per_step": ms, "value": b.n_obj * N_PROP / (ms * 1e-3),
            "unit": UNIT, "scores": "IoU2D, dims, corners, segmentation, mod-segmentation, points-in-cube, IoU3D, 8 curves, 26 products",
            "bound": "int / fp32 issue (closed-form fillPoly spans per sampled row; point-in-interval compares)",
            "mean_best_iou3d": float(r["iou3d"].max(1).values.mean())}


def bench_c4(args, rank, world, dev):
    """BASELINE configs[3]: pseudo-GT generation over a synthetic 10 000-frame split, STRONG scaling -- rank r owns the
    contiguous block shard.frame_range(total, r, world).  Frames are produced on the device chunk by chunk (never 77 GB on
    the host); per chunk the PRED-branch forward (tdod.pseudo_gt.forward_pred) and the Omni3D annotation records through the
    resumable writer; the winners are gathered at the end.  Philox keys are global frame / object ids, so the winners do
    not depend on the world size."""
    import tempfile
    import torch.distributed as dist
    from tdod import pseudo_gt as PG, shard, synth
    c = synth.CONFIGS["C4"]
    H, W, total = c["H"], c["W"], args.c4_frames
    lo, hi = shard.frame_range(total, rank, world)
    counts = synth.object_counts(range(total), c["n_obj"])
    obj_base = int(np.sum(counts[:lo]))
    chunk = 256
    tmp = tempfile.mkdtemp(prefix="tdod_c4_")
    w = PG.PseudoGTWriter(os.path.join(tmp, "pseudo_gt_rank%d.json" % rank), {"id": 12, "name": "synthetic C4"}, [], resume=False)
    t_gen = t_write = 0.0
    gpu_ms = 0.0
    n_obj = 0
    recs = []
    if hi > lo:       # untimed warm-up pass over the first chunk: kernels of the PRED branch loaded, allocator pools grown
        wb, _, _ = synth.make_batch(range(lo, min(hi, lo + chunk)), H, W, c["n_obj"], dev)
        for _ in range(2):
            PG.forward_pred(wb, N_PROP, seed=7, obj_id_base=obj_base, frame_id_base=lo)
        torch.cuda.synchronize(dev)
        del wb
    t_all0 = time.perf_counter()
    for f0 in range(lo, hi, chunk):
        f1 = min(hi, f0 + chunk)
        t0 = time.perf_counter()
        batch, off, cls = synth.make_batch(range(f0, f1), H, W, c["n_obj"], dev)
        torch.cuda.synchronize(dev)
        t_gen += time.perf_counter() - t0
        a, e = _events(2)
        a.record()
        fields, sel = PG.forward_pred(batch, N_PROP, seed=7, obj_id_base=obj_base + n_obj, frame_id_base=f0)
        e.record()
        recs.append(shard.pack_winners(sel))
        t0 = time.perf_counter()
        cl = cls.tolist()
        meta = [dict(id=fid, width=W, height=H, file_path="synthetic/%06d.jpg" % fid, K=batch.K[k].tolist())
                for k, fid in enumerate(range(f0, f1))]
        PG.write_frames(w, fields, off, meta, cl, ["class%d" % k for k in cl], gt_boxes=batch.boxes, net_hw=(H, W))
        t_write += time.perf_counter() - t0
        gpu_ms += a.elapsed_time(e)
        if os.environ.get("TDOD_BENCH_DEBUG"):
            sys.stderr.write("[c4 rank %d] chunk %d..%d: %.3f ms device, gen %.3f s, write %.3f s\n" % (rank, f0, f1, a.elapsed_time(e), t_gen, t_write))
        n_obj += batch.n_obj
        del batch, fields, sel
    t0 = time.perf_counter()
    w.finalize(return_doc=False)
    n_ann = w.next_id - 1
    t_write += time.perf_counter() - t0
    wall = time.perf_counter() - t_all0
    local = torch.cat(recs) if recs else torch.zeros((0, shard.WINNER_COLS), device=dev)
    t0 = time.perf_counter()
    allw = shard.gather_winners(local)
    torch.cuda.synchronize(dev)
    t_gather = time.perf_counter() - t0
    v = torch.tensor([gpu_ms, wall - t_gen, t_write, t_gen, float(n_obj), float(hi - lo)], dtype=torch.float64, device=dev)
    if world > 1:
        vmax = v.clone(); dist.all_reduce(vmax, op=dist.ReduceOp.MAX)
        vsum = v.clone(); dist.all_reduce(vsum, op=dist.ReduceOp.SUM)
    else:
        vmax = vsum = v
    props = float(vsum[4]) * N_PROP
    import shutil
    shutil.rmtree(tmp, ignore_errors=True)
    return {"frames": total, "frames_rank0": hi - lo, "objects": int(vsum[4]), "scaling": "strong",
            "path_ms_max_rank": float(vmax[0]), "value": props / (float(vmax[0]) * 1e-3), "unit": UNIT,
            "frames_per_s_path": float(vsum[5]) / (float(vmax[0]) * 1e-3),
            "with_writer_s_max_rank": float(vmax[1]), "value_with_writer": props / float(vmax[1]),
            "writer_s_max_rank": float(vmax[2]), "writer_share": float(vmax[2]) / max(float(vmax[1]), 1e-9),
            "synthetic_generation_s_max_rank": float(vmax[3]), "annotations_rank0": n_ann,
            "final_gather": {"records": int(allw.shape[0]), "ms": t_gather * 1e3},
            "note": "value = proposals of all ranks / device time of the slowest rank (CUDA events around the per-chunk forward); "
                    "with_writer adds the Python json records (host-bound); frame generation is outside both"}


def run_tdod(args):
    import torch.distributed as dist
    from tdod import api, pipeline, synth, lib as L
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a GPU: libtdod has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L.lib()
    c = synth.CONFIGS[WORKLOAD]
    H, W = c["H"], c["W"]
    F = args.frames
    # weak scaling: rank r owns global frames [r*F, (r+1)*F); no data-path collective
    batch, off, _ = synth.make_batch(range(rank * F, (rank + 1) * F), H, W, c["n_obj"], dev)
    N = batch.n_obj
    props = N * N_PROP
    obj_base = rank * F * 64          # any injective global id works for the Philox keying
    abytes = algorithmic_bytes(batch, N_PROP)
    stream = torch.cuda.current_stream(dev)
    stage_names = list(abytes)
    stage_marks = [("ka0", "ka1"), ("kb0", "kb1"), ("kc0", "kc1")]

    def step(marks=None, overlap=True):
        """one pass of the hot path (tdod.pipeline.select): K-A -> yaw table on the auxiliary stream, K-B on the current
        stream, then K-C; overlap=False keeps everything on one stream (stage times without contention)"""
        return pipeline.select(batch, N_PROP, seed=7, obj_id_base=obj_base, frame_id_base=rank * F, overlap=overlap,
                               marks=marks)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(max(args.warmup, 3)):
        o = step()
    barrier()
    # The timed steps replay ONE CUDA graph of the step (pipeline.GraphedSelect: the same 11 kernels, K-A forked onto the
    # auxiliary stream inside the capture): the GPU never waits for Python, so a host hiccup cannot leak into a
    # 20 ms timed region.  The same K steps launched eagerly are timed right after and reported beside it.
    gs, graph_note = None, "eager launches (capture unavailable)"
    if not args.no_graph:
        try:
            gs = pipeline.GraphedSelect(batch, N_PROP, warmup=max(args.warmup, 3), depth=args.graph_depth, seed=7, obj_id_base=obj_base,
                                        frame_id_base=rank * F)
            graph_note = ("one CUDA graph per step (%d kernel nodes, K-A on a forked stream), replayed; %d copies of the graph on %d "
                          "streams round-robin, so that step i+1's mask kernel runs beside step i's proposal kernel (every step "
                          "complete, own outputs)" % (gs.launches, gs.depth, gs.depth))
        except Exception as e:                      # a capture problem must not cost the bench line
            gs, graph_note = None, "eager launches (graph capture failed: %s)" % repr(e)[:120]
    barrier()
    t_start, t_end, g_start, g_end = _events(4)
    with ClockSampler(local) as clk:
        barrier()
        if args.profile:
            torch.cuda.profiler.start()
        t_start.record(stream)
        for k in range(args.steps):
            o = gs.replay() if gs is not None else step()
        if gs is not None:
            gs.join()
        t_end.record(stream)
        barrier()
        if args.profile:
            torch.cuda.profiler.stop()
        total_ms = t_start.elapsed_time(t_end)
        eager_ms = total_ms
        single_ms = None
        if gs is not None and gs.depth > 1:            # the same K replays of ONE graph on one stream: steps back to back
            q0, q1 = _events(2)
            q0.record(stream)
            for k in range(args.steps):
                gs.graphs[0].replay()
            q1.record(stream)
            barrier()
            single_ms = q0.elapsed_time(q1)
        if gs is not None:
            g_start.record(stream)
            for k in range(args.steps):
                o = step()
            g_end.record(stream)
            barrier()
            eager_ms = g_start.elapsed_time(g_end)
        # the same K steps once more on ONE stream with events between the stages: stage times free of contention (inside the
        # timed region K-A runs on an auxiliary stream beside K-B).  Still under the clock sampler.
        marks_serial = [dict() for _ in range(args.steps)]
        s0, s1 = _events(2)
        barrier()
        s0.record(stream)
        for k in range(args.steps):
            step(marks_serial[k], overlap=False)
        s1.record(stream)
        barrier()
    serial_ms_step = s0.elapsed_time(s1) / args.steps
    stage_ms_serial = [float(np.mean([m[a].elapsed_time(m[b]) for m in marks_serial])) for a, b in stage_marks]
    # np.argmax returns the first NaN (roi_heads.py:501): objects whose proposals include a zero-height projected box
    # have a NaN winner in the reference too; everything else must be finite
    winners_nan = int(torch.isnan(o.best_score).sum().item())
    winners_ok = bool((torch.isfinite(o.best_score) | torch.isnan(o.best_score)).all().item())

    # ---- the proposal set DEFINED BY THE REFERENCE: torch's generators in its draw order, inside the timed region ----
    parity = None
    if not args.no_parity_rng:
        gd_, gc_ = torch.Generator(device=dev).manual_seed(1000 + rank), torch.Generator().manual_seed(2000 + rank)
        p_steps = 3
        barrier()
        a0, a1 = _events(2)
        t0 = time.perf_counter()
        a0.record(stream)
        for _ in range(p_steps):
            nz, yw = pipeline.draw_reference_noise(batch, off, N_PROP, generator=gd_, cpu_generator=gc_)
            pipeline.select(batch, N_PROP, noise=nz, yaw_idx=yw, frame_id_base=rank * F, std_mode=L.STD_TORCH_CUDA)
        a1.record(stream)
        barrier()
        p_ms = a0.elapsed_time(a1) / p_steps
        parity = {"value": props / (p_ms * 1e-3), "unit": UNIT, "ms_per_step": p_ms,
                  "host_ms_per_step": (time.perf_counter() - t0) * 1e3 / p_steps,
                  "what": "normals from torch's CUDA generator and yaw bins from its CPU generator in the reference's program "
                          "order (per frame, whole-tensor redraw loops with their host synchronisation), fed through HOST mode "
                          "with std_mode TORCH_CUDA: the proposal set is bit-equal to the reference's on CUDA tensors "
                          "(tests/test_gpu_round2.py); the draws, not the kernels, take the time"}

    # ---- final gather of the 84-byte winner records (the only exchange; outside the timed region) ----
    from tdod import shard
    t0 = time.perf_counter()
    gathered = shard.gather_winners(shard.pack_winners(o))
    torch.cuda.synchronize(dev)
    gather_ms = (time.perf_counter() - t0) * 1e3

    # ---- e2e: pinned host inputs -> chunks -> winners back on the host, everything inside the timed region ----
    e2e = None
    if not args.no_e2e:
        hb = batch.to("cpu").pin()
        pipe = pipeline.Pipeline(H, W, N_PROP, chunk_frames=32, device=dev, seed=7)
        out = torch.empty((N, 21), dtype=torch.float32).pin_memory()
        for _ in range(2):
            pipe.run_host(hb, off, out=out, obj_id_base=obj_base, frame_id_base=rank * F)
        pipe.h2d_bytes = pipe.d2h_bytes = pipe.launches = 0
        e_steps = max(3, min(args.steps, 10))
        barrier()
        t0 = time.perf_counter()
        for _ in range(e_steps):
            pipe.run_host(hb, off, out=out, obj_id_base=obj_base, frame_id_base=rank * F)
        barrier()
        e_ms = (time.perf_counter() - t0) * 1e3 / e_steps
        e2e = {"ms": e_ms, "h2d": pipe.h2d_bytes // e_steps, "d2h": pipe.d2h_bytes // e_steps,
               "same_winners": bool(torch.equal(out[:, 20].int(), o.best_idx.cpu()))}
        del pipe
        # the same host batch (bool masks), but every chunk's masks are packed to 1 bit/px on the host cores while the previous
        # chunk's copy is in flight (Pipeline(host_pack=True) -> tdod_host_pack_masks): the packing is inside the timed region
        pipe = pipeline.Pipeline(H, W, N_PROP, chunk_frames=64, device=dev, seed=7, host_pack=True)
        out3 = torch.empty((N, 21), dtype=torch.float32).pin_memory()
        for _ in range(2):
            pipe.run_host(hb, off, out=out3, obj_id_base=obj_base, frame_id_base=rank * F)
        pipe.h2d_bytes = pipe.d2h_bytes = 0
        barrier()
        t0 = time.perf_counter()
        for _ in range(e_steps):
            pipe.run_host(hb, off, out=out3, obj_id_base=obj_base, frame_id_base=rank * F)
        barrier()
        e2e["hp_ms"] = (time.perf_counter() - t0) * 1e3 / e_steps
        e2e["hp_h2d"] = pipe.h2d_bytes // e_steps
        e2e["hp_threads"] = pipe.pack_threads
        e2e["hp_same_winners"] = bool(torch.equal(out3[:, 20].int(), o.best_idx.cpu()))
        del pipe
        # the OPTIONAL host format with the masks at 1 bit per pixel (pipeline.pack_masks; tdod_mask_rect_bitmap): reported
        # beside the headline, never instead of it -- the reference's callers hold bool masks
        hp = pipeline.FrameBatch(hb.depth, hb.ground, hb.K, hb.boxes, hb.prior_mean, hb.prior_std, hb.obj_frame, None,
                                 hb.frame_ids, hb.has_ground, pipeline.pack_masks(hb.masks).pin_memory())
        pipe = pipeline.Pipeline(H, W, N_PROP, chunk_frames=32, device=dev, seed=7, packed_masks=True)
        out2 = torch.empty((N, 21), dtype=torch.float32).pin_memory()
        for _ in range(2):
            pipe.run_host(hp, off, out=out2, obj_id_base=obj_base, frame_id_base=rank * F)
        pipe.h2d_bytes = pipe.d2h_bytes = 0
        barrier()
        t0 = time.perf_counter()
        for _ in range(e_steps):
            pipe.run_host(hp, off, out=out2, obj_id_base=obj_base, frame_id_base=rank * F)
        barrier()
        e2e["packed_ms"] = (time.perf_counter() - t0) * 1e3 / e_steps
        e2e["packed_h2d"] = pipe.h2d_bytes // e_steps
        e2e["packed_same_winners"] = bool(torch.equal(out2[:, 20].int(), o.best_idx.cpu()))
        del hb, hp, pipe

    # ---- the other BASELINE configs (sub-records; C2 stays the headline) ----
    peak, peak_src = peaks()
    configs = {}
    if not args.no_configs:
        del batch
        torch.cuda.empty_cache()
        if rank == 0:
            try:
                configs.update(bench_small_configs(dev, 10, peak))
                configs["C2 mode B"] = bench_mode_b(dev, peak)
            except Exception as e:
                configs["error"] = repr(e)[:300]
        torch.cuda.empty_cache()
        barrier()
        try:
            configs["C4"] = bench_c4(args, rank, world, dev)
        except Exception as e:
            configs["C4"] = {"error": repr(e)[:300]}

    # ---- max over ranks ----
    t = torch.tensor([total_ms, e2e["ms"] if e2e else 0.0, float(props), parity["ms_per_step"] if parity else 0.0,
                      e2e["packed_ms"] if e2e else 0.0, eager_ms, e2e["hp_ms"] if e2e else 0.0], dtype=torch.float64, device=dev)
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        total_ms, e_ms_max, props_all, p_ms_max, ep_ms_max = float(tmax[0]), float(tmax[1]), float(tsum[2]), float(tmax[3]), float(tmax[4])
        eager_ms, hp_ms_max = float(tmax[5]), float(tmax[6])
    else:
        e_ms_max, props_all, p_ms_max, ep_ms_max, hp_ms_max = float(t[1]), float(props), float(t[3]), float(t[4]), float(t[6])
    if rank == 0:
        ms_step = total_ms / args.steps
        step_bytes = sum(abytes.values())
        j = int(np.argmax(stage_ms_serial))                # the roofline object describes the stage that takes longest
        dom = stage_names[j]
        ach = step_bytes / (ms_step * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": props_all / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "BASELINE configs[1]: %d frames %dx%d per GPU, %d objects (~15/frame) x %d proposals, mode A "
                                   "(back-projection + RANSAC plane, mask rectangle, fused propose/project/IoU x dims x corners/argmax)"
                                   % (F, W, H, N, N_PROP),
                       "frames_per_gpu": F, "objects_rank0": N, "rng": "in-kernel Philox4x32-10 (replayed bit for bit through the "
                       "reference-pinned HOST mode in tests/test_gpu_round2.py)",
                       "l2": "inputs (1.98 GB per pass) are larger than the 126 MB L2",
                       "launch": graph_note},
            "eager": {"ms_per_step": eager_ms / args.steps, "value": props_all / (eager_ms / args.steps * 1e-3), "unit": UNIT,
                      "what": "the same K steps as eleven eager launches per step from Python (pipeline.select)"},
            "graph_back_to_back": None if single_ms is None else {"ms_per_step": single_ms / args.steps,
                                                                   "what": "K replays of one graph on one stream (no overlap between steps), this rank"},
            "gpu_launches": pipeline.launches_per_select(N) * args.steps,
            "stages_ms_serial": dict(zip(stage_names, stage_ms_serial)), "ms_per_step_serial": serial_ms_step,
            "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                         "traffic": sum(TRAFFIC.values()), "peak_source": peak_src,
                         "what": "the whole step: SURVEY 8(d) algorithmic bytes of all stages / measured step time (98 % of the "
                                 "bytes are mask bytes, read once by K-B)",
                         "step_algorithmic_bytes": step_bytes,
                         "kernel": dom, "kernel_ms": stage_ms_serial[j], "kernel_bound": BOUND[dom],
                         "kernel_hbm_frac": abytes[dom] / (stage_ms_serial[j] * 1e-3) / 1e9 / peak,
                         "kernel_algorithmic_bytes_per_launch": abytes[dom], "kernel_traffic": TRAFFIC.get(dom),
                         "timed": "stages_ms_serial: the same K steps once more on one stream, CUDA events between the stages",
                         "all_stages": {n: {"ms": m, "bound": BOUND[n], "hbm_frac": abytes[n] / (m * 1e-3) / 1e9 / peak}
                                        for n, m in zip(stage_names, stage_ms_serial)}},
            "clocks": clk.summary(), "winners_ok": winners_ok, "winners_nan": winners_nan,
            "final_gather": {"records": int(gathered.shape[0]), "bytes": int(gathered.numel() * 4), "ms": gather_ms},
        }
        if parity:
            parity["value"] = props_all / (p_ms_max * 1e-3)
            line["value_parity_rng"] = parity
        if e2e:
            # headline: the faster of the two ways to move the SAME host batch (bool masks); both are reported
            use_hp = hp_ms_max < e_ms_max
            best_ms = hp_ms_max if use_hp else e_ms_max
            line["e2e"] = {"value": props_all / (best_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": e2e["hp_h2d"] if use_hp else e2e["h2d"],
                           "d2h_bytes_per_step": e2e["d2h"], "ms_per_step": best_ms,
                           "same_winners_as_resident": e2e["hp_same_winners"] if use_hp else e2e["same_winners"],
                           "host_masks": "torch.bool, 1 byte per pixel (what the reference's callers hold)",
                           "transfer": ("masks packed to 1 bit/px on %d host threads per chunk inside the timed region, copied packed "
                                        "(Pipeline(host_pack=True))" % e2e["hp_threads"]) if use_hp else "every tensor copied as it is",
                           "variants": {"bool masks copied as they are": {"ms_per_step": e_ms_max, "value": props_all / (e_ms_max * 1e-3),
                                                                            "h2d_bytes_per_step": e2e["h2d"]},
                                        "bool masks packed on the host cores": {"ms_per_step": hp_ms_max, "value": props_all / (hp_ms_max * 1e-3),
                                                                                "h2d_bytes_per_step": e2e["hp_h2d"], "threads": e2e["hp_threads"]}},
                           "optional_packed_masks": {"value": props_all / (ep_ms_max * 1e-3), "unit": UNIT, "ms_per_step": ep_ms_max,
                                                     "h2d_bytes_per_step": e2e["packed_h2d"],
                                                     "same_winners_as_resident": e2e["packed_same_winners"],
                                                     "what": "host batch with masks at 1 bit per pixel (pipeline.pack_masks, "
                                                             "packed outside the timed region); labelled, not the headline"}}
        if configs:
            line["configs"] = configs
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_single()
        print(json.dumps(line), file=_JSON_OUT, flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_tdod(a)
