"""Host-to-device bandwidth from pinned memory for subsets of the box's GPUs (one process, one stream per GPU): is the
end-to-end limit at 8 GPUs a per-pair limit (a shared PCIe switch uplink) or an aggregate one (host memory)?"""
import subprocess, sys, time, torch
n = torch.cuda.device_count()
SZ = 512 << 20
host = [torch.empty(SZ, dtype=torch.uint8).pin_memory() for _ in range(n)]
devb = [torch.empty(SZ, dtype=torch.uint8, device="cuda:%d" % i) for i in range(n)]
streams = [torch.cuda.Stream("cuda:%d" % i) for i in range(n)]
def run(ids, reps=6):
    for i in ids:
        with torch.cuda.stream(streams[i]): devb[i].copy_(host[i], non_blocking=True)
    for i in ids: streams[i].synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        for i in ids:
            with torch.cuda.stream(streams[i]): devb[i].copy_(host[i], non_blocking=True)
    for i in ids: streams[i].synchronize()
    dt = time.perf_counter() - t0
    return len(ids) * reps * SZ / dt / 1e9
sets = [[0], [1], [0, 1], [0, 2], [0, 4], [2, 3], [0, 1, 2, 3], [0, 2, 4, 6], [4, 5, 6, 7], list(range(n))]
for ids in sets:
    if max(ids) < n:
        g = run(ids)
        print("GPUs %-26s aggregate %7.1f GB/s   per GPU %6.1f GB/s" % (ids, g, g / len(ids)), flush=True)
for cmd in (["nvidia-smi", "topo", "-m"], ["numactl", "-H"], ["lscpu"]):
    try:
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=20).stdout
        print("$ " + " ".join(cmd)); print("\n".join(out.splitlines()[:40]))
    except Exception as e:
        print("$ %s: %r" % (" ".join(cmd), e))
