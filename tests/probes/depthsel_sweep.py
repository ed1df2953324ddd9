"""One-off stress comparison (not a pytest): tdod_depth_select (medians and masked quantile ranges of depth crops,
roi_heads.py:1196-1231, :1279-1305) against torch.median / torch.quantile of the same crops on the device, on depth maps
with heavy ties, negative values, signed zeros left out, and NaNs; boxes that straddle, leave or degenerate inside the
image.    python tests/probes/depthsel_sweep.py [n_rounds]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "3dod_b200"))
import torch

from tdod import weak_loss as WL

rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 10
d = torch.device("cuda:0")
H, W, F_, n = 530, 730, 4, 96
tot = dict(boxes=0, median_bad=0, median_nan=0, fallback=0, fallback_bad=0, empty=0, range_items=0, range_bad=0,
           range_bitequal=0, range_worst_ulp=0.0, range_fallback=0)
yy, xx = torch.meshgrid(torch.arange(H, device=d), torch.arange(W, device=d), indexing="ij")
for k in range(rounds):
    g = torch.Generator().manual_seed(100 + k)
    depth = torch.rand((F_, H, W), generator=g) * 9 - 0.5
    depth[1] = torch.round(depth[1] * 3) / 3 + 0.125                  # ~30 distinct values
    depth[2] = torch.round(depth[2] * 50) / 50 + 0.01                 # ~450 distinct values
    depth[3, 100:110, 200:260] = float("nan")
    depth = depth.to(d)
    x1 = torch.rand(n, generator=g) * (W + 60) - 30
    y1 = torch.rand(n, generator=g) * (H + 60) - 30
    w = torch.rand(n, generator=g) ** 2 * W * 0.7
    h = torch.rand(n, generator=g) ** 2 * H * 0.7
    boxes = torch.stack((x1, y1, x1 + w, y1 + h), 1)
    boxes[0] = torch.tensor([W + 3.0, 5.0, W + 30.0, 60.0])           # no area after clipping
    boxes[1] = torch.tensor([10.2, 10.1, 10.9, 50.0])                 # area, empty slice
    frame = torch.randint(0, F_, (n,), generator=g)
    got, flag = WL.depth_select(depth, frame, boxes.to(d), 0)
    got, flag = got.cpu(), flag.cpu()
    for i in range(n):
        b = boxes[i]
        c = torch.stack((b[0].clamp(0, W), b[1].clamp(0, H), b[2].clamp(0, W), b[3].clamp(0, H)))
        tot["boxes"] += 1
        if not (c[2] - c[0]) * (c[3] - c[1]) > 0:
            tot["fallback"] += 1
            x = torch.clamp((c[0] + c[2]) / 2, 10, W - 11).long(); y = torch.clamp((c[1] + c[3]) / 2, 10, H - 11).long()
            ref = depth[frame[i], y, x].item()
            ok = flag[i] == 0 and (got[i].item() == ref or (ref != ref and got[i].item() != got[i].item()))
            tot["fallback_bad"] += (not ok)
            continue
        crop = depth[frame[i], int(c[1]):int(c[3]), int(c[0]):int(c[2])]
        if crop.numel() == 0:
            tot["empty"] += 1
            tot["median_bad"] += not (got[i].item() != got[i].item())
            continue
        ref = torch.median(crop).item()
        if ref != ref:
            tot["median_nan"] += 1
            tot["median_bad"] += not (got[i].item() != got[i].item())
        else:
            tot["median_bad"] += (got[i].item() != ref or flag[i] != 1)
    # masked quantile ranges: discs of all sizes, some empty (fall back to the box slice)
    cx = torch.rand(n, generator=g) * W; cy = torch.rand(n, generator=g) * H
    rad = torch.rand(n, generator=g) ** 2 * 200
    rad[::11] = 0
    masks = ((xx[None] - cx.to(d)[:, None, None]) ** 2 + (yy[None] - cy.to(d)[:, None, None]) ** 2) < (rad.to(d)[:, None, None] ** 2)
    inb = boxes.clone()
    inb[:, 0::2] = inb[:, 0::2].clamp(0, W); inb[:, 1::2] = inb[:, 1::2].clamp(0, H)
    got, flag = WL.depth_select(depth, frame, inb.to(d), 2, masks=masks.contiguous(), mask_idx=torch.arange(n))
    got, flag = got.cpu(), flag.cpu()
    for i in range(n):
        v = depth[frame[i]][masks[i]]
        if v.numel() == 0:
            tot["range_fallback"] += 1
            bl = inb[i].long()
            v = depth[frame[i], bl[1]:bl[3], bl[0]:bl[2]].reshape(-1)
            if flag[i] != 0:
                tot["range_bad"] += 1
        tot["range_items"] += 1
        if v.numel() == 0 or bool(torch.isnan(v).any()):
            tot["range_bad"] += not (got[i].item() != got[i].item())
            continue
        hi, lo = torch.quantile(v, 0.9), torch.quantile(v, 0.1)
        ref = (hi - lo).item()
        ulp = max(abs(hi.item()), 1e-30) * 2.0 ** -23
        err = abs(got[i].item() - ref)
        tot["range_bitequal"] += (err == 0)
        tot["range_worst_ulp"] = max(tot["range_worst_ulp"], err / ulp)
        tot["range_bad"] += (err > 2 * ulp)
print({k: (float(v) if k == "range_worst_ulp" else int(v)) for k, v in tot.items()})
