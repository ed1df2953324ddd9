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


def object_counts(frame_ids, n_obj, seed_base=1234):
    """objects per frame of `make_frame` / `make_batch` without building the frames (first draw of each frame's stream)"""
    return [_num_objects(n_obj, np.random.RandomState(seed_base + int(i))) for i in frame_ids]


def make_batch(frame_ids, H, W, n_obj, device, seed_base=1234, sub=32):
    """The frames `make_frame` builds one by one, built for a whole list of frame ids with batched device operations
    (boxes, classes, priors, ground and masks are identical to make_frame's; the depth noise comes from a device
    generator seeded per frame, so a frame does not depend on how the list is chunked or sharded).  Returns a
    tdod.pipeline.FrameBatch on `device` and the CSR object offsets.  Used where thousands of frames are needed
    (BASELINE configs[3]: 10 000 frames)."""
    from .pipeline import FrameBatch
    dev = torch.device(device)
    F = len(frame_ids)
    P = dict(bw=[], bh=[], x0=[], y0=[], step=[], cls=[], ang=[], nfx=[], nfy=[], bfx=[], bfy=[])
    z0s, as_, hor, counts = [], [], [], []
    for fid in frame_ids:
        rs = np.random.RandomState(seed_base + int(fid))
        N = _num_objects(n_obj, rs)
        counts.append(N)
        P["bw"].append(rs.uniform(0.08, 0.5, N) * W); P["bh"].append(rs.uniform(0.08, 0.6, N) * H)
        P["x0"].append(rs.uniform(0, 1, N)); P["y0"].append(rs.uniform(0, 1, N))
        z0s.append(rs.uniform(1.0, 2.0)); as_.append(rs.uniform(2.0, 4.0))
        P["step"].append(rs.uniform(0.2, 1.0, N).astype(np.float32))
        hor.append(rs.uniform(0.55, 0.70))
        P["cls"].append(rs.randint(0, priors_table().shape[0], N))
        for k in ("ang",):
            P[k].append(rs.uniform(0, np.pi, N))
        P["nfx"].append(rs.uniform(-0.5, 0.5, N)); P["nfy"].append(rs.uniform(-0.5, 0.5, N))
        P["bfx"].append(rs.uniform(0.0, 1.0, N)); P["bfy"].append(rs.uniform(0.0, 1.0, N))
    cat = {k: np.concatenate(v) if len(v) else np.zeros(0) for k, v in P.items()}
    bw, bh = cat["bw"], cat["bh"]
    x0 = cat["x0"] * (W - 1 - bw); y0 = cat["y0"] * (H - 1 - bh)
    boxes = np.stack([x0, y0, x0 + bw, y0 + bh], 1).astype(np.float32)
    Ntot = boxes.shape[0]
    off = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    obj_frame = torch.repeat_interleave(torch.arange(F, dtype=torch.int32), torch.tensor(counts)).to(dev)
    vv = torch.arange(H, device=dev, dtype=torch.float32)[None, :, None]
    uu = torch.arange(W, device=dev, dtype=torch.float32)[None, None, :]
    z0t = torch.tensor(z0s, dtype=torch.float32, device=dev)[:, None, None]
    at = torch.tensor(as_, dtype=torch.float32, device=dev)[:, None, None]
    depth = (z0t + at * (1.0 - vv / H)).expand(F, H, W).clone()
    ground = (vv >= torch.tensor(hor, dtype=torch.float32, device=dev)[:, None, None] * H).expand(F, H, W).clone()
    masks = torch.empty((Ntot, H, W), dtype=torch.bool, device=dev)
    bi = torch.from_numpy(boxes.astype(np.int64)).to(dev)          # int() truncation of the box, as make_frame slices
    bf = torch.from_numpy(boxes).to(dev)
    tt = lambda k: torch.from_numpy(cat[k].astype(np.float64)).to(dev)
    step_t = torch.from_numpy(cat["step"]).to(dev)
    ang, nfx, nfy, bfx, bfy = tt("ang"), tt("nfx"), tt("nfy"), tt("bfx"), tt("bfy")
    for f0 in range(0, F, sub):
        f1 = min(F, f0 + sub)
        o0, o1 = int(off[f0]), int(off[f1])
        if o1 == o0:
            continue
        b = bi[o0:o1]
        inbox = ((vv >= b[:, 1, None, None]) & (vv <= b[:, 3, None, None]) & (uu >= b[:, 0, None, None]) & (uu <= b[:, 2, None, None]))
        of = obj_frame[o0:o1].long() - f0
        depth[f0:f1].index_add_(0, of, -(inbox.to(torch.float32) * step_t[o0:o1, None, None]))
        anyin = torch.zeros((f1 - f0, H, W), dtype=torch.float32, device=dev).index_add_(0, of, inbox.to(torch.float32))
        ground[f0:f1] &= anyin == 0
        # masks: ellipse inscribed in the box, minus a rotated-rectangle notch, plus a detached 5 x 5 blob
        bb = bf[o0:o1].double()
        cx, cy = (bb[:, 0] + bb[:, 2]) / 2, (bb[:, 1] + bb[:, 3]) / 2
        rx, ry = (bb[:, 2] - bb[:, 0]) / 2, (bb[:, 3] - bb[:, 1]) / 2
        f32 = lambda t: t.to(torch.float32)[:, None, None]
        ex = (uu - f32(cx)) / f32(rx); ey = (vv - f32(cy)) / f32(ry)
        m = (ex * ex + ey * ey) <= 1.0
        ncx, ncy = cx + nfx[o0:o1] * rx, cy + nfy[o0:o1] * ry
        ca, sa = torch.cos(ang[o0:o1]), torch.sin(ang[o0:o1])
        px = (uu - f32(ncx)) * f32(ca) + (vv - f32(ncy)) * f32(sa)
        py = -(uu - f32(ncx)) * f32(sa) + (vv - f32(ncy)) * f32(ca)
        m &= ~((px.abs() <= f32(0.35 * rx.to(torch.float32).double())) & (py.abs() <= f32(0.12 * ry.to(torch.float32).double())))
        bx = (bfx[o0:o1] * (W - 6)).long(); by = (bfy[o0:o1] * (H - 6)).long()
        away = ~((bb[:, 0] - 7 <= bx) & (bx <= bb[:, 2] + 2) & (bb[:, 1] - 7 <= by) & (by <= bb[:, 3] + 2))
        blob = ((uu >= bx[:, None, None]) & (uu < bx[:, None, None] + 5) & (vv >= by[:, None, None]) & (vv < by[:, None, None] + 5))
        masks[o0:o1] = m | (blob & away[:, None, None])
    for k, fid in enumerate(frame_ids):
        g = torch.Generator(device=dev).manual_seed(seed_base + int(fid))
        depth[k] += torch.randn((H, W), generator=g, device=dev) * 0.02
    depth.clamp_(0.3, 20.0)
    pt = priors_table()
    cls = cat["cls"].astype(np.int64)
    K = torch.tensor([[0.9 * W, 0.0, W / 2.0], [0.0, 0.9 * W, H / 2.0], [0.0, 0.0, 1.0]], dtype=torch.float32, device=dev)
    batch = FrameBatch(depth, ground, K[None].expand(F, 3, 3).contiguous(), bf, torch.as_tensor(pt[cls, 0], device=dev),
                       torch.as_tensor(pt[cls, 1], device=dev), obj_frame, masks, [int(i) for i in frame_ids])
    batch_classes = torch.as_tensor(cls, device=dev)
    return batch, off.tolist(), batch_classes
