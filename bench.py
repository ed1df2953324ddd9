"""bench.py -- scored cube proposals / second (BASELINE.json `metric`) on BASELINE.json configs[1]:
an Omni3D / SUN-RGBD-shaped batch of 256 synthetic frames 730x530, ~15 objects per frame x 1000 proposals.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--frames F]

A step = one pass of the hot path over the batch: K-A (back-projection + RANSAC ground plane + yaw table),
K-B (mask -> min-area rectangle) and the fused K-C (propose + project + score + select) for every object.
  value     device-resident inputs, CUDA events on the launching stream, max over ranks (weak scaling: every rank
            owns its own 256 frames)
  e2e       the same batch through tdod.pipeline.Pipeline.run_host: pinned HOST inputs, H2D copies and the D2H read
            of the winners inside the timed region
  roofline  the stage that dominates the step, algorithmic bytes / its CUDA-event time over the timed region
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
# dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` capture of this workload
# (profiles/r01_f_ncu_full_raw_c2.csv): pack_tma_kernel 1.4878 GB + 0.1818 GB, rect_kernel passes 0.0955 GB + 0.0041 GB
TRAFFIC = {"K-B pack": 1.4878e9 + 0.1818e9, "K-B rect": 0.0955e9 + 0.0041e9}
# what binds each stage: only the pack kernel moves gigabytes; the others are issue- / latency-bound
BOUND = {"K-A backproject+ransac": "fp32 fma issue (hypothesis scoring)", "K-B pack": "hbm",
         "K-B rect": "latency (serial hull / calipers chains on L2-resident bitmaps)",
         "K-C propose_score_select": "fp32 / sfu issue"}
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
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop, self.t = index, [], threading.Event(), None

    def _run(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            self.stop.wait(0.1)

    def __enter__(self):
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.t.join(timeout=6)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(r) >= 6 and r[2 + k].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


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
def algorithmic_bytes(batch, n_prop, stride=5):
    """compulsory HBM traffic per stage for one pass over `batch` (every needed input element read
    once, every required output written once)."""
    F, H, W = batch.depth.shape
    N = batch.n_obj
    P = ((H + stride - 1) // stride) * ((W + stride - 1) // stride)
    ka = F * (P * 4 + P * 1 + 36 + 12 + 16)                       # strided depth + ground samples, K, normal, plane
    kb1 = N * (H * W)                                             # every mask pixel once (the bitmap is an intermediate)
    kb2 = N * 33                                                  # rectangle + flag out
    kc = N * (4 * n_prop + 44 + 32 + 84)                          # diagonal depth gather, box/priors, rect, winner record
    return {"K-A backproject+ransac": ka, "K-B pack": kb1, "K-B rect": kb2, "K-C propose_score_select": kc}


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
    frame_ids = range(rank * F, (rank + 1) * F)
    frames = [synth.make_frame(i, H, W, c["n_obj"], device=dev) for i in frame_ids]
    batch = pipeline.FrameBatch.from_frames(frames, dev)
    off = [0]
    for f in frames:
        off.append(off[-1] + f.n_obj())
    del frames
    N = batch.n_obj
    props = N * N_PROP
    obj_base = rank * F * 64          # any injective global id works for the Philox keying
    abytes = algorithmic_bytes(batch, N_PROP)

    stream = torch.cuda.current_stream(dev)
    stage_names = list(abytes)
    stage_marks = [("ka0", "ka1"), ("kb0", "kb_mid"), ("kb_mid", "kb1"), ("kc0", "kc1")]

    def step(marks=None, overlap=True):
        """one pass of the hot path (tdod.pipeline.select): K-A -> yaw table on the auxiliary stream, K-B on the current
        stream, then K-C; overlap=False keeps everything on one stream (stage times without contention)"""
        return pipeline.select(batch, N_PROP, seed=7, obj_id_base=obj_base, frame_id_base=rank * F, overlap=overlap,
                               marks=marks)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def stage_times(all_marks):
        return [float(np.mean([m[a].elapsed_time(m[b]) for m in all_marks])) for a, b in stage_marks]

    for _ in range(max(args.warmup, 3)):
        o = step()
    barrier()
    t_start, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        barrier()
        if args.profile:
            torch.cuda.profiler.start()
        t_start.record(stream)
        for k in range(args.steps):
            o = step()
        t_end.record(stream)
        barrier()
        if args.profile:
            torch.cuda.profiler.stop()
    total_ms = t_start.elapsed_time(t_end)
    # the same K steps once more on ONE stream with events between the stages: stage times free of contention (inside the
    # timed region K-A runs on an auxiliary stream and two object ranges are in flight on two more, so kernels overlap)
    marks_serial = [dict() for _ in range(args.steps)]
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    s0.record(stream)
    for k in range(args.steps):
        step(marks_serial[k], overlap=False)
    s1.record(stream)
    barrier()
    serial_ms_step = s0.elapsed_time(s1) / args.steps
    stage_ms_serial = stage_times(marks_serial)
    # np.argmax returns the first NaN (roi_heads.py:501): objects whose proposals include a zero-height projected box
    # have a NaN winner in the reference too; everything else must be finite
    winners_nan = int(torch.isnan(o.best_score).sum().item())
    winners_ok = bool((torch.isfinite(o.best_score) | torch.isnan(o.best_score)).all().item())

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

    # ---- max over ranks ----
    t = torch.tensor([total_ms, e2e["ms"] if e2e else 0.0, float(props)], dtype=torch.float64, device=dev)
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        total_ms, e_ms_max, props_all = float(tmax[0]), float(tmax[1]), float(tsum[2])
    else:
        e_ms_max, props_all = float(t[1]), float(props)
    if rank == 0:
        peak, peak_src = peaks()
        ms_step = total_ms / args.steps
        # the roofline object describes the kernel that moves the step's bytes (98 % of the compulsory traffic): the
        # HBM-bound stage with the longest duration; every stage's time, bound and HBM fraction is listed beside it
        hbm = [k for k, nme in enumerate(stage_names) if BOUND[nme] == "hbm"]
        j = max(hbm, key=lambda k: stage_ms_serial[k])
        dom = stage_names[j]
        ach = abytes[dom] / (stage_ms_serial[j] * 1e-3) / 1e9
        step_bytes = sum(abytes.values())
        line = {
            "metric": METRIC, "value": props_all / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "BASELINE configs[1]: %d frames %dx%d per GPU, %d objects (~15/frame) x %d proposals, mode A "
                                   "(back-projection + RANSAC plane, mask rectangle, fused propose/project/IoU x dims x corners/argmax)"
                                   % (F, W, H, N, N_PROP),
                       "frames_per_gpu": F, "objects_rank0": N, "rng": "in-kernel Philox4x32-10",
                       "l2": "inputs (%.2f GB per pass) are larger than the 126 MB L2" % (batch.input_bytes() / 1e9)},
            "gpu_launches": pipeline.launches_per_select(N) * args.steps,
            "stages_ms_serial": dict(zip(stage_names, stage_ms_serial)), "ms_per_step_serial": serial_ms_step,
            "roofline": {"kernel": dom, "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                         "traffic": TRAFFIC.get(dom), "peak_source": peak_src, "algorithmic_bytes_per_launch": abytes[dom],
                         "timed": "stages_ms_serial: the same K steps once more on one stream, CUDA events between the stages "
                                  "(in the timed region the stages overlap on several streams)",
                         "step_algorithmic_bytes": step_bytes, "step_frac": step_bytes / (ms_step * 1e-3) / 1e9 / peak,
                         "all_stages": {n: {"ms": m, "bound": BOUND[n], "hbm_frac": abytes[n] / (m * 1e-3) / 1e9 / peak}
                                        for n, m in zip(stage_names, stage_ms_serial)}},
            "clocks": clk.summary(), "winners_ok": winners_ok, "winners_nan": winners_nan,
            "final_gather": {"records": int(gathered.shape[0]), "bytes": int(gathered.numel() * 4), "ms": gather_ms},
        }
        if e2e:
            line["e2e"] = {"value": props_all / (e_ms_max * 1e-3), "unit": UNIT, "h2d_bytes_per_step": e2e["h2d"],
                           "d2h_bytes_per_step": e2e["d2h"], "ms_per_step": e_ms_max, "same_winners_as_resident": e2e["same_winners"]}
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
