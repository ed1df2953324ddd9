"""Multi-GPU host logic on CPU (gloo, world_size 2): frame partition, winner gather.
The data path itself has no collective; what is checked here is that the gathered records are the single-process
records in frame order, for ragged per-rank object counts."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tdod import shard, synth


def test_frame_range_covers_everything_once():
    for F in (0, 1, 7, 256, 10000):
        for world in (1, 2, 3, 4, 8):
            got = []
            for r in range(world):
                lo, hi = shard.frame_range(F, r, world)
                assert 0 <= lo <= hi <= F
                got += list(range(lo, hi))
            assert got == list(range(F))
            sizes = [shard.frame_range(F, r, world) for r in range(world)]
            assert max(h - l for l, h in sizes) - min(h - l for l, h in sizes) <= 1
    with pytest.raises(ValueError):
        shard.frame_range(4, 2, 2)


def test_balanced_ranges_follow_the_work():
    rs = np.random.RandomState(0)
    work = np.clip(rs.poisson(15, 256), 1, 40) * 1000
    for world in (1, 2, 4, 8):
        rr = shard.balanced_frame_ranges(work, world)
        assert rr[0][0] == 0 and rr[-1][1] == 256
        assert all(a[1] == b[0] for a, b in zip(rr, rr[1:]))
        loads = [work[l:h].sum() for l, h in rr]
        assert max(loads) <= work.sum() / world + work.max()
    assert shard.balanced_frame_ranges([5.0], 2) in ([(0, 1), (1, 1)], [(0, 0), (0, 1)])


def _records_for_frames(lo, hi):
    """stand-in winner records derived from the frames' (host-side, world-size independent) object lists"""
    c = synth.CONFIGS["C2"]
    rows = []
    for fid in range(lo, hi):
        f = synth.make_frame(fid, 64, 96, c["n_obj"], with_masks=False)
        for i in range(f.n_obj()):
            r = torch.zeros(shard.WINNER_COLS)
            r[0:4] = f.boxes[i]
            r[4:7] = f.prior_mean[i]
            r[19] = float(fid)
            r[20] = float(i)
            rows.append(r)
    return torch.stack(rows) if rows else torch.zeros((0, shard.WINNER_COLS))


def _worker(rank, world, port, n_frames, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = shard.frame_range(n_frames, rank, world)
        local = _records_for_frames(lo, hi)
        allw = shard.gather_winners(local)
        q.put((rank, allw.numpy()))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_frames", [5, 1])
def test_gather_winners_gloo_world2(n_frames):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_frames, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    whole = _records_for_frames(0, n_frames).numpy()
    assert np.array_equal(got[0], whole)
    assert np.array_equal(got[1], whole)


def test_gather_is_identity_without_a_process_group():
    t = torch.arange(42, dtype=torch.float32).reshape(2, 21)
    assert shard.gather_winners(t) is t
