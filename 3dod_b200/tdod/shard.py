"""Multi-GPU plumbing: frames are independent units, so they are block-partitioned over ranks with
no collective on the data path; the only exchange is the final gather of the 84-byte winner records, the analogue of
the reference's `comm.gather(inference_json, dst=0)` (cubercnn/evaluation/omni3d_evaluation.py:635-636).

Random numbers are keyed by GLOBAL frame / object id (Philox), so the winners do not depend on the world size:
rank r passes `frame_id_base = frame_range(...)[0]` and `obj_id_base = obj_offsets[frame_id_base]`.
One process per GPU; NCCL on GPUs, gloo in the CPU tests (the helpers only touch torch.distributed)."""
import torch
import torch.distributed as dist

WINNER_COLS = 21   # cube(15) + box(4) + score + index, fp32 (the index is exact below 2^24)


def frame_range(n_frames, rank, world):
    """contiguous block partition: the first `n_frames % world` ranks own one frame more"""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank / world")
    q, r = divmod(int(n_frames), world)
    lo = rank * q + min(rank, r)
    return lo, lo + q + (1 if rank < r else 0)


def balanced_frame_ranges(work_per_frame, world):
    """contiguous ranges balanced by per-frame work (e.g. objects x proposals): greedy prefix split.
    Returns [(lo, hi)] * world covering [0, F) in order."""
    w = [float(x) for x in work_per_frame]
    F, total = len(w), sum(w)
    out, lo, acc = [], 0, 0.0
    for r in range(world):
        target = total * (r + 1) / world
        hi = lo
        while hi < F and (acc + w[hi] <= target + 1e-9 or hi == lo) and (F - hi) > (world - 1 - r):
            acc += w[hi]
            hi += 1
        if r == world - 1:
            hi = F
        out.append((lo, hi))
        lo = hi
    return out


def pack_winners(sel):
    """Selection / PssOutputs -> [N,21] fp32 on the same device"""
    n = sel.best_idx.shape[0]
    p = torch.empty((n, WINNER_COLS), dtype=torch.float32, device=sel.best_idx.device)
    p[:, 0:15] = sel.best_cube
    p[:, 15:19] = sel.best_box
    p[:, 19] = sel.best_score
    p[:, 20] = sel.best_idx.to(torch.float32)
    return p


def gather_winners(local, group=None):
    """local: [N_r,21] records of this rank's frames (rank order == frame order).  Returns the concatenation over
    ranks, in rank order, on every rank ([sum N_r, 21]); a no-op without an initialised process group."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    n = torch.tensor([local.shape[0]], dtype=torch.int64, device=local.device)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n, group=group)
    counts = [int(c.item()) for c in counts]
    cap = max(max(counts), 1)
    pad = torch.zeros((cap, local.shape[1]), dtype=local.dtype, device=local.device)
    pad[:local.shape[0]] = local
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    return torch.cat([p[:c] for p, c in zip(parts, counts)], dim=0)
