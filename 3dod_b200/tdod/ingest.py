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
