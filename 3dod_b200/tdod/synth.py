"""Synthetic frames of the shapes BASELINE.json names: depth, ground mask, 2D boxes,
classes -> size priors, object masks, intrinsics.  No datasets exist offline, so every config runs on
these.  Scalar parameters are drawn with NumPy (host, seeded by frame id => identical on any device and
independent of world size); the rasterisation runs with torch ops on the requested device."""
import os

import numpy as np
import torch

_PRIORS = None

CONFIGS = {
    # name: (H, W, objects per frame (int or "poisson15"), proposals per object, frames)
    "C1": dict(H=480, W=640, n_obj=5, n_prop=1000, frames=1),
    "C2": dict(H=530, W=730, n_obj="poisson15", n_prop=1000, frames=256),
    "C3": dict(H=768, W=1024, n_obj=50, n_prop=10000, frames=1),
    "C4": dict(H=530, W=730, n_obj="poisson15", n_prop=1000, frames=10000),
    "C5": dict(H=1080, W=1920, n_obj=0, n_prop=0, frames=512),
}


def priors_table():
    """[38,2,3] fp32: valid rows of the reference's tools/priors.pkl `priors_dims_per_cat`
    (frozen by tools/extract_priors.py); [:,0] mean, [:,1] std, order (w,h,l)."""
    global _PRIORS
    if _PRIORS is None:
        a = np.load(os.path.join(os.path.dirname(__file__), "data", "priors_dims_per_cat.npy"))
        valid = ~np.all(a.reshape(a.shape[0], -1) == 1, axis=1)
        _PRIORS = a[valid].astype(np.float32)
    return _PRIORS


class Frame:
    """One synthetic frame; tensors live on `device`."""
    __slots__ = ("frame_id", "H", "W", "depth", "ground", "boxes", "classes", "prior_mean",
                 "prior_std", "masks", "K")

    def n_obj(self):
        return int(self.boxes.shape[0])


def _num_objects(spec, rs):
    if spec == "poisson15":
        return int(np.clip(rs.poisson(15), 1, 40))
    return int(spec)


def make_frame(frame_id, H, W, n_obj, device="cpu", with_masks=True, seed_base=1234):
    rs = np.random.RandomState(seed_base + frame_id)
    N = _num_objects(n_obj, rs)
    dev = torch.device(device)
    f = Frame()
    f.frame_id, f.H, f.W = frame_id, H, W
    vv = torch.arange(H, device=dev, dtype=torch.float32)[:, None]
    uu = torch.arange(W, device=dev, dtype=torch.float32)[None, :]

    # boxes: width in [0.08,0.5] W, height in [0.08,0.6] H, fully inside the image
    bw = rs.uniform(0.08, 0.5, N) * W
    bh = rs.uniform(0.08, 0.6, N) * H
    x0 = rs.uniform(0, 1, N) * (W - 1 - bw)
    y0 = rs.uniform(0, 1, N) * (H - 1 - bh)
    boxes = np.stack([x0, y0, x0 + bw, y0 + bh], 1).astype(np.float32)

    # depth: tilted floor z = z0 + a (1 - v/H), box-shaped steps under the objects, noise, clip
    z0 = rs.uniform(1.0, 2.0)
    a = rs.uniform(2.0, 4.0)
    depth = (z0 + a * (1.0 - vv / H)).expand(H, W).clone()
    step = rs.uniform(0.2, 1.0, N).astype(np.float32)
    for i in range(N):
        xa, ya, xb, yb = [int(t) for t in boxes[i]]
        depth[ya:yb + 1, xa:xb + 1] -= float(step[i])
    g = torch.Generator(device="cpu").manual_seed(seed_base + frame_id)
    noise = torch.randn((H, W), generator=g) * 0.02
    depth = (depth + noise.to(dev)).clamp_(0.3, 20.0)
    f.depth = depth.contiguous()

    # ground: rows below a random horizon, minus object boxes
    horizon = rs.uniform(0.55, 0.70)
    ground = (vv >= horizon * H).expand(H, W).clone()
    for i in range(N):
        xa, ya, xb, yb = [int(t) for t in boxes[i]]
        ground[ya:yb + 1, xa:xb + 1] = False
    f.ground = ground.contiguous()

    pt = priors_table()
    cls = rs.randint(0, pt.shape[0], N)
    f.classes = torch.as_tensor(cls, device=dev)
    f.prior_mean = torch.as_tensor(pt[cls, 0], device=dev)
    f.prior_std = torch.as_tensor(pt[cls, 1], device=dev)
    f.boxes = torch.as_tensor(boxes, device=dev)
    f.K = torch.tensor([[0.9 * W, 0.0, W / 2.0], [0.0, 0.9 * W, H / 2.0], [0.0, 0.0, 1.0]],
                       dtype=torch.float32, device=dev)

    f.masks = None
    if with_masks and N > 0:
        # ellipse inscribed in the box, a rotated-rectangle notch, and one small detached blob
        masks = torch.zeros((N, H, W), dtype=torch.bool, device=dev)
        ang = rs.uniform(0, np.pi, N)
        nfx = rs.uniform(-0.5, 0.5, N)
        nfy = rs.uniform(-0.5, 0.5, N)
        bfx = rs.uniform(0.0, 1.0, N)
        bfy = rs.uniform(0.0, 1.0, N)
        for i in range(N):
            cx, cy = (boxes[i, 0] + boxes[i, 2]) / 2, (boxes[i, 1] + boxes[i, 3]) / 2
            rx, ry = (boxes[i, 2] - boxes[i, 0]) / 2, (boxes[i, 3] - boxes[i, 1]) / 2
            ex = (uu - float(cx)) / float(rx)
            ey = (vv - float(cy)) / float(ry)
            m = (ex * ex + ey * ey) <= 1.0
            # notch: rotated rectangle centred inside the ellipse
            ncx, ncy = cx + nfx[i] * rx, cy + nfy[i] * ry
            ca, sa = float(np.cos(ang[i])), float(np.sin(ang[i]))
            px = (uu - float(ncx)) * ca + (vv - float(ncy)) * sa
            py = -(uu - float(ncx)) * sa + (vv - float(ncy)) * ca
            notch = (px.abs() <= 0.35 * float(rx)) & (py.abs() <= 0.12 * float(ry))
            m = m & ~notch
            # detached blob: 5x5 square somewhere in the image, away from the ellipse box
            bx = int(bfx[i] * (W - 6)); by = int(bfy[i] * (H - 6))
            if not (boxes[i, 0] - 7 <= bx <= boxes[i, 2] + 2 and boxes[i, 1] - 7 <= by <= boxes[i, 3] + 2):
                m[by:by + 5, bx:bx + 5] = True
            masks[i] = m
        f.masks = masks
    return f


def make_frames(cfg, frame_ids, device="cpu", with_masks=True):
    c = CONFIGS[cfg] if isinstance(cfg, str) else cfg
    return [make_frame(i, c["H"], c["W"], c["n_obj"], device, with_masks) for i in frame_ids]
