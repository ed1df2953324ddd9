"""Depth / ground-map ingest: what the reference's DatasetMapper does per frame on the CPU before
the path -- `np.load(...npz)` then `PIL.Image.resize(size, NEAREST)` to the network resolution
(cubercnn/data/dataset_mapper.py:125-147) -- with the resize on the device: the small decoded maps are copied up
(pinned, asynchronous) and expanded there, so the full-resolution maps never cross PCIe."""
import numpy as np
import torch

from . import lib as L

_MAPS = {}


def pil_nearest_map(n_in, n_out):
    """source index of every output pixel as Pillow's nearest resize computes it: a double accumulator that starts at
    half a step and advances by n_in / n_out per pixel, truncated (validated against PIL 12.2 in tests/test_ingest.py)"""
    key = (int(n_in), int(n_out))
    if key not in _MAPS:
        s = float(n_in) / float(n_out)
        steps = np.full(n_out, s, dtype=np.float64)
        steps[0] = s * 0.5
        idx = np.floor(np.cumsum(steps)).astype(np.int64)        # cumsum adds sequentially, like the C loop
        _MAPS[key] = np.clip(idx, 0, n_in - 1).astype(np.int32)
    return _MAPS[key]


def resize_nearest(src, H, W):
    """src [F,h,w] (or [h,w]) CUDA tensor of float32 / uint8 / bool -> [F,H,W] (or [H,W]), PIL NEAREST semantics"""
    squeeze = src.dim() == 2
    s = src[None] if squeeze else src
    if not s.is_cuda:
        raise ValueError("src must be a CUDA tensor (no CPU path)")
    if s.dtype not in (torch.float32, torch.uint8, torch.bool, torch.int32):
        raise ValueError("src must be float32, int32, uint8 or bool")
    s = s.contiguous()
    F, h, w = s.shape
    xmap = torch.from_numpy(pil_nearest_map(w, W)).to(s.device, non_blocking=True)
    ymap = torch.from_numpy(pil_nearest_map(h, H)).to(s.device, non_blocking=True)
    dst = torch.empty((F, H, W), dtype=s.dtype, device=s.device)
    L.check(L.lib().tdod_resize_nearest(s.data_ptr(), s.element_size(), F, h, w, xmap.data_ptr(), ymap.data_ptr(),
                                        dst.data_ptr(), H, W, L.stream_ptr()))
    return dst[0] if squeeze else dst


# ---- prefetch: .npz decode in worker threads -> pinned staging -> asynchronous copy -> device resize -------------------

def load_maps(depth_path, ground_path=None):
    """host half of dataset_mapper.py:125-147: the arrays stored by generate_depth_maps.py:86 (`depth=`, float32) and
    generate_ground_segmentations.py:200 (`mask=`, bool).  zlib inflate releases the GIL, so this runs in threads."""
    with np.load(depth_path) as z:
        depth = np.ascontiguousarray(z["depth"], dtype=np.float32)
    ground = None
    if ground_path is not None:
        with np.load(ground_path) as z:
            ground = np.ascontiguousarray(z["mask"]).astype(np.uint8, copy=False)
    return depth, ground


class MapPrefetcher:
    """Iterates over frames given as (depth_path, ground_path | None): yields (depth [H,W] float32, ground [H,W] uint8 |
    None) on the device at the network resolution (H, W), PIL-NEAREST semantics, `ahead` frames decoded and copied in
    advance.  Decode runs in `workers` threads, the stored-resolution maps go through pinned staging buffers (a ring of
    `ahead` slots) and one copy stream; the full-resolution maps are produced on the device and never cross PCIe.
    The yielded tensors are ordered after their copy on the consumer's current stream (wait_event), no host sync."""

    def __init__(self, items, H, W, device, workers=4, ahead=4):
        from concurrent.futures import ThreadPoolExecutor
        self.items, self.H, self.W = list(items), int(H), int(W)
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise ValueError("MapPrefetcher needs a CUDA device (no CPU path)")
        self.ahead = max(1, int(ahead))
        self.pool = ThreadPoolExecutor(max_workers=max(1, int(workers)))
        self.copy_stream = torch.cuda.Stream(self.device)
        self.slots = [dict(depth=None, ground=None, free=None) for _ in range(self.ahead)]
        self.futures = {}
        self.next_submit = 0
        self.bytes_h2d = 0

    def _submit(self):
        while self.next_submit < len(self.items) and len(self.futures) < self.ahead:
            dp, gp = self.items[self.next_submit]
            self.futures[self.next_submit] = self.pool.submit(load_maps, dp, gp)
            self.next_submit += 1

    def _stage(self, slot, key, arr, dtype):
        buf = slot[key]
        if buf is None or buf.numel() < arr.size:
            buf = slot[key] = torch.empty(arr.size, dtype=dtype, pin_memory=True)
        view = buf[:arr.size].view(arr.shape)
        view.copy_(torch.from_numpy(arr))
        self.bytes_h2d += arr.nbytes
        return view

    def __len__(self):
        return len(self.items)

    def __iter__(self):
        self._submit()
        for k in range(len(self.items)):
            depth, ground = self.futures.pop(k).result()
            self._submit()
            slot = self.slots[k % self.ahead]
            if slot["free"] is not None:
                slot["free"].synchronize()                       # the copy that last read this staging slot has finished
            with torch.cuda.stream(self.copy_stream):
                d = self._stage(slot, "depth", depth, torch.float32).to(self.device, non_blocking=True)
                g = None if ground is None else self._stage(slot, "ground", ground, torch.uint8).to(self.device, non_blocking=True)
                slot["free"] = torch.cuda.Event()
                slot["free"].record(self.copy_stream)
                d = resize_nearest(d, self.H, self.W) if d.shape != (self.H, self.W) else d
                if g is not None and g.shape != (self.H, self.W):
                    g = resize_nearest(g, self.H, self.W)
                ready = torch.cuda.Event()
                ready.record(self.copy_stream)
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(ready)
            d.record_stream(cur)
            if g is not None:
                g.record_stream(cur)
            yield d, g
        self.pool.shutdown(wait=False)
