"""TEST INFRASTRUCTURE ONLY -- CPU restatement (oracle) of the reference's proposal-and-scoring path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import
this module.  The product path (3dod_b200/) never does, and has no CPU fallback.

What it is: a restatement of luchsonice/3dod `ProposalNetwork/` on raw arrays (no Cubes / Boxes
classes), keeping the reference's algorithmic structure (including its per-proposal cv2 / cKDTree
loops, so that timing it is representative of the reference) and using the same numerical
libraries the reference itself calls: torch-CPU (RNG stream, atan2/sin/sqrt/median/std), NumPy,
OpenCV (findContours / minAreaRect / convexHull / fillPoly) and SciPy (cKDTree).  OpenCV,
SciPy, detectron2 and pyransac3d are unpinned third-party dependencies of the reference
(requirements.txt:9,13,14) that are not vendored under /root/reference; versions used here are
recorded in every golden file (tests/golden/*.npz: key `versions`).

Parity pin: the reference ships no golden vectors for this path, so the oracle
is pinned against OUTPUTS OF THE REFERENCE ITSELF, imported unmodified in the build container
(oracle/ref_import.py) and frozen by tests/golden/make_golden.py into tests/golden/*.npz; the
reference's own known answers (caixa.ply RANSAC, the 3-point point-cloud count) are included.
tests/test_oracle_golden.py re-checks the oracle against those fixtures on every run.

All file:line citations are relative to /root/reference.
"""
import math

import numpy as np
import torch

try:  # third-party deps of the reference (scorefunction.py:3,6)
    import cv2
    from scipy.spatial import cKDTree
except Exception:  # pragma: no cover - the image has both
    cv2 = None
    cKDTree = None

MIN_PROP_S = 0.05  # proposals.py:10
N_YAW = 36         # proposals.py:405


def versions():
    import scipy
    return "torch=%s numpy=%s cv2=%s scipy=%s" % (
        torch.__version__, np.__version__, getattr(cv2, "__version__", "none"), scipy.__version__)


# ----------------------------------------------------------------------------------------------
# sampling  (ProposalNetwork/proposals/proposals.py:338-424, ProposalNetwork/utils/utils.py:42-177)
# ----------------------------------------------------------------------------------------------
def linspace_rows(lo, hi, n):
    """utils.py:170-177 vectorized_linspace: arange(n) * ((hi-lo)/(n-1)) + lo, fp32, per row."""
    step = (hi - lo) / (n - 1)
    ramp = torch.arange(0, n, dtype=lo.dtype, device=lo.device)   # device-agnostic: the GPU tests run this on CUDA tensors
    return ramp[None, :] * step[:, None] + lo[:, None]


def centre_seeds(boxes, depth, K, n):
    """proposals.py:354-375.  boxes [N,4] xyxy fp32, depth [H,W] fp32, K [3,3] fp32.
    Returns x, y, z_tmp, each [N,n] fp32: metric centre seeds along the box diagonal."""
    bw = boxes[:, 2] - boxes[:, 0]
    bh = boxes[:, 3] - boxes[:, 1]
    gx = linspace_rows(boxes[:, 0] + bw / 4, boxes[:, 2] - bw / 4, n).long()
    gy = linspace_rows(boxes[:, 1] + bh / 4, boxes[:, 3] - bh / 4, n).long()
    d = depth[gy, gx]
    ox = gx - K[0, 2]
    oy = gy - K[1, 2]
    f = K[0, 0]                       # fx for both axes (proposals.py:368)
    ang_x = torch.atan2(ox, f.expand_as(ox))
    r = torch.sqrt(ox ** 2 + f ** 2)
    ang_d = torch.atan2(oy, r)
    y = d * torch.sin(ang_d)
    dx = torch.sqrt(d ** 2 - y ** 2)
    x = dx * torch.sin(ang_x)
    z_tmp = torch.sqrt(dx ** 2 - x ** 2)
    return x, y, z_tmp


def draw_normal_rows(mean, std, n, lo=None, hi=None, generator=None):
    """utils.py:42-60 sample_normal_in_range.  mean,std [N]; hi [N] or None; lo scalar or None.
    torch.normal(mean_t, std_t) == randn*std+mean with separate mul/add; the
    standard-normal stream is consumed identically, which lets us also return the ACCEPTED
    standard normal per element (what the CUDA parity mode is fed).  Returns (samples, z, tries)."""
    N = mean.shape[0]
    dev = mean.device                      # CUDA tensors + a CUDA generator: the reference's production flavour
    z = torch.randn((N, n), generator=generator, device=dev)
    s = z * std[:, None] + mean[:, None]
    tries = 0
    if lo is not None and hi is not None:
        hi_b = hi[:, None].expand_as(s)
        while bool(torch.any((s < lo) | (s > hi_b))):
            bad = (s < lo) | (s > hi_b)
            z2 = torch.randn((N, n), generator=generator, device=dev)
            s2 = z2 * std[:, None] + mean[:, None]
            s = torch.where(bad, s2, s)
            z = torch.where(bad, z2, z)
            tries += 1
            if tries == 10000:
                break
    return s, z, tries


def perp_vector(nrm):
    """utils.py:112-118 vec_perp_t (+ normalize_vector :10-16)."""
    a, b, c = nrm[0], nrm[1], nrm[2]
    if a == 0:
        return torch.stack([torch.zeros_like(a), c, -b])
    v = torch.stack([b, -a, torch.zeros_like(a)])
    mag = torch.sqrt((v ** 2).sum())
    # utils.py:12: torch.max(0-d tensor, fp32 tensor([1e-8])) promotes to the DIMENSIONED operand's
    # dtype, so an fp64 magnitude is rounded to fp32 before the divide (reproduced on purpose)
    mag = torch.max(mag, torch.tensor([1e-8]))
    return v / mag


def yaw_table(normal, n_yaw=N_YAW):
    """utils.py:120-146 orthobasis_from_normal_t over linspace(0, pi, 36) (proposals.py:405-406).
    `normal` keeps its dtype (fp64 when it comes from the NumPy plane fit, roi_heads.py:428);
    the angles are fp32 so cos/sin are evaluated in fp32 and then promoted.  Returns [36,3,3]
    with columns (x_rot, normal, normal x x_rot)."""
    ang = torch.linspace(0, np.pi, n_yaw)
    c, s = torch.cos(ang), torch.sin(ang)
    v = perp_vector(normal)
    kxv = torch.linalg.cross(normal, v)
    kdv = normal @ v
    xr = v[:, None] * c + kxv[:, None] * s + (normal * kdv)[:, None] * (1 - c)   # [3,36]
    yr = torch.linalg.cross(normal[:, None].expand_as(xr), xr, dim=0)              # [3,36]
    tab = torch.stack([xr.t(), normal[None, :].expand(n_yaw, 3).to(xr.dtype), yr.t()], dim=2)
    return tab  # [36, 3(row), 3(col = x, n, y)]


def random_rotations(N, n, generator=None):
    """utils.py:62-69 randn_orthobasis_torch (CPU generator)."""
    z = torch.randn((N, n, 3, 3), generator=generator)
    z = z / torch.norm(z, p=2, dim=-1, keepdim=True)
    z[:, :, 0] = torch.linalg.cross(z[:, :, 1], z[:, :, 2], dim=-1)
    z[:, :, 0] = z[:, :, 0] / torch.norm(z[:, :, 0], dim=-1, keepdim=True)
    z[:, :, 1] = torch.linalg.cross(z[:, :, 2], z[:, :, 0], dim=-1)
    z[:, :, 1] = z[:, :, 1] / torch.norm(z[:, :, 1], dim=-1, keepdim=True)
    return z


def cubes_from_noise(boxes, depth, prior_mean, prior_std, K, n, noise, normal=None, yaw_idx=None, rot=None):
    """The arithmetic of proposals.py:338-413 with the random numbers GIVEN: noise [6,N,n] accepted standard normals
    in draw order (w,h,l,x,y,z); yaw_idx [N,n] (with `normal`) or rot [N,n,3,3].  torch.normal(mean_t, std_t) ==
    z*std + mean with separately rounded multiply and add."""
    x0, y0, z0 = centre_seeds(boxes, depth, K, n)
    sc = lambda z, m, sd: z * sd[:, None] + m[:, None]
    w = sc(noise[0], prior_mean[:, 0], prior_std[:, 0])
    h = sc(noise[1], prior_mean[:, 1], prior_std[:, 1] * 1.1)
    l = sc(noise[2], prior_mean[:, 2], prior_std[:, 2])
    cx = torch.tensor([1.15, 0.0]); cy = torch.tensor([1.1, 0.0]); cz = torch.tensor([0.85, 0.35])
    x = sc(noise[3], cx[0] * torch.median(x0, dim=1).values + cx[1], torch.std(x0, dim=1) * 1.2)
    y = sc(noise[4], cy[0] * torch.median(y0, dim=1).values + cy[1], torch.std(y0, dim=1) * 0.8)
    zt = z0 + l / 2
    z = sc(noise[5], cz[0] * torch.median(zt, dim=1).values + cz[1], torch.std(zt, dim=1) * 1.2)
    R = rot if normal is None else yaw_table(normal.cpu())[yaw_idx.long().cpu()].to(boxes.device)
    xyzwhl = torch.stack([x, y, z, w, h, l], 2)
    return torch.cat((xyzwhl, R.flatten(start_dim=2)), dim=2).to(torch.float32)          # spaces.py:143


def propose(boxes, depth, prior_mean, prior_std, K, n, normal=None, generator=None, host_generator=None):
    """proposals.py:338-424 `propose` on raw tensors.  `generator` feeds the six normal draws (on the tensors' device);
    `host_generator` (default: `generator`) the draws the reference makes on the HOST whatever the device: the yaw bins
    (torch.randint without a device, :407) and the random rotations (utils.py:62-69).  Returns dict with
    cubes [N,n,15] fp32, noise [6,N,n] accepted standard normals in draw order (w,h,l,x,y,z),
    yaw_idx [N,n] int64 (or None), rot [N,n,3,3] (or None), tries [3]."""
    N = boxes.shape[0]
    x0, y0, z0 = centre_seeds(boxes, depth, K, n)
    mw, mh, ml = prior_mean[:, 0], prior_mean[:, 1], prior_mean[:, 2]
    sw, sh, sl = prior_std[:, 0], prior_std[:, 1], prior_std[:, 2]
    w, zw, tw = draw_normal_rows(mw, sw, n, MIN_PROP_S, mw + 2 * sw, generator)
    h, zh, th = draw_normal_rows(mh, sh * 1.1, n, MIN_PROP_S, mh + 2.2 * sh, generator)
    l, zl, tl = draw_normal_rows(ml, sl, n, MIN_PROP_S, ml + 2 * sl, generator)
    # lin_fun(median, [a, b]) = a*median + b with fp32 coefficient tensors (proposals.py:386-396)
    cx = torch.tensor([1.15, 0.0]); cy = torch.tensor([1.1, 0.0]); cz = torch.tensor([0.85, 0.35])
    x, zx, _ = draw_normal_rows(cx[0] * torch.median(x0, dim=1).values + cx[1],
                                torch.std(x0, dim=1) * 1.2, n, generator=generator)
    y, zy, _ = draw_normal_rows(cy[0] * torch.median(y0, dim=1).values + cy[1],
                                torch.std(y0, dim=1) * 0.8, n, generator=generator)
    zt = z0 + l / 2
    z, zz, _ = draw_normal_rows(cz[0] * torch.median(zt, dim=1).values + cz[1],
                                torch.std(zt, dim=1) * 1.2, n, generator=generator)
    xyzwhl = torch.stack([x, y, z, w, h, l], 2)
    yaw_idx = None
    rot = None
    hg = generator if host_generator is None else host_generator
    if normal is None:
        R = rot = random_rotations(N, n, hg).to(boxes.device)
    else:
        tab = yaw_table(normal.cpu())
        yaw_idx = torch.randint(N_YAW, (N, n), generator=hg)
        R = tab[yaw_idx].to(boxes.device)
    cubes = torch.cat((xyzwhl, R.flatten(start_dim=2)), dim=2).to(torch.float32)  # spaces.py:143
    return dict(cubes=cubes, noise=torch.stack([zw, zh, zl, zx, zy, zz], 0), yaw_idx=yaw_idx, rot=rot,
                tries=(tw, th, tl), seeds=(x0, y0, z0))


# ----------------------------------------------------------------------------------------------
# ablation samplers (proposals.py:20-336): same output contract as `propose`, different marginals.
# Each returns dict(cubes [N,n,15], noise [6,N,n] in (w,h,l,x,y,z) order -- standard normals or U[0,1)
# draws as the variant uses them, zeros for channels it does not draw --, rot [N,n,3,3], aux [N,2] | None)
# ----------------------------------------------------------------------------------------------
RATIOS = [0.33, 0.66, 1, 1.33, 1.67, 2, 3]   # proposals.py:232


def _rescale(u):
    """proposals.py:12-14 rescale_interval(u, MIN_PROP_S, 2)"""
    return (MIN_PROP_S - 2) * u + 2


def normalised_ramps(boxes, im_shape, n):
    """proposals.py:51-62 / :98-107: the inner half of the box in normalised image space
    (conversions.py:50-67: (p - 0.5*old) * (new/old); x with (im_w, 3), y with (im_h, 2)), then vectorized_linspace."""
    bw = boxes[:, 2] - boxes[:, 0]
    bh = boxes[:, 3] - boxes[:, 1]

    def norm(lo, hi, old, new):
        p = torch.stack([lo, hi], dim=1).to(torch.float32)
        for i in range(2):
            p[:, i] -= 0.5 * old
            p[:, i] *= new / old
        return p
    xt = norm(boxes[:, 0] + bw / 4, boxes[:, 2] - bw / 4, im_shape[0], 3)
    yt = norm(boxes[:, 1] + bh / 4, boxes[:, 3] - bh / 4, im_shape[1], 2)
    return linspace_rows(xt[:, 0], xt[:, 1], n), linspace_rows(yt[:, 0], yt[:, 1], n)


def _seeded_centres(boxes, depth, K, n, l, generator):
    """proposals.py:179-191 (= :386-396): median / std of the seeds, three untruncated normal draws"""
    x0, y0, z0 = centre_seeds(boxes, depth, K, n)
    cx = torch.tensor([1.15, 0.0]); cy = torch.tensor([1.1, 0.0]); cz = torch.tensor([0.85, 0.35])
    x, zx, _ = draw_normal_rows(cx[0] * torch.median(x0, dim=1).values + cx[1], torch.std(x0, dim=1) * 1.2, n,
                                generator=generator)
    y, zy, _ = draw_normal_rows(cy[0] * torch.median(y0, dim=1).values + cy[1], torch.std(y0, dim=1) * 0.8, n,
                                generator=generator)
    zt = z0 + l / 2
    z, zz, _ = draw_normal_rows(cz[0] * torch.median(zt, dim=1).values + cz[1], torch.std(zt, dim=1) * 1.2, n,
                                generator=generator)
    return (x, y, z), (zx, zy, zz)


def _pack(x, y, z, w, h, l, R, noise, aux=None):
    xyzwhl = torch.stack([x, y, z, w, h, l], 2)
    cubes = torch.cat((xyzwhl, R.flatten(start_dim=2)), dim=2).to(torch.float32)
    return dict(cubes=cubes, noise=torch.stack(noise, 0), rot=R, aux=aux)


def propose_variant(name, boxes, depth, prior_mean, prior_std, K, im_shape, n, generator=None):
    """name in random | xy | z | dim | aspect | rotation (the `proposal_function` strings of roi_heads.py:286-299)."""
    N = boxes.shape[0]
    U = lambda: torch.rand((N, n), generator=generator)
    zero = torch.zeros((N, n))
    if name == "random":                                    # proposals.py:20-45
        ux, uy, uz = U(), U(), U()
        x, y, z = ux * 4 - 2, uy * 2 - 1, uz * 4 + 1
        uw, uh, ul = U(), U(), U()
        R = random_rotations(N, n, generator)
        return _pack(x, y, z, _rescale(uw), _rescale(uh), _rescale(ul), R, [uw, uh, ul, ux, uy, uz])
    if name == "xy":                                        # :47-91
        x, y = normalised_ramps(boxes, im_shape, n)
        uz = U()
        z = uz * 4 + 1
        uw, uh, ul = U(), U(), U()
        R = random_rotations(N, n, generator)
        return _pack(x, y, z, _rescale(uw), _rescale(uh), _rescale(ul), R, [uw, uh, ul, zero, zero, uz])
    if name == "z":                                         # :93-135
        x, y = normalised_ramps(boxes, im_shape, n)
        z = torch.zeros_like(x)
        for i in range(N):
            patch = depth[int(boxes[i, 1]):int(boxes[i, 3]), int(boxes[i, 0]):int(boxes[i, 2])]
            q = torch.quantile(patch, torch.tensor([0.1, 0.9]), dim=None)
            z[i] = torch.linspace(q[0], q[1], n)
        uw, uh, ul = U(), U(), U()
        R = random_rotations(N, n, generator)
        return _pack(x, y, z, _rescale(uw), _rescale(uh), _rescale(ul), R, [uw, uh, ul, zero, zero, z])
    if name == "dim":                                       # :137-197
        uw, uh, ul = U(), U(), U()
        w, h, l = _rescale(uw), _rescale(uh), _rescale(ul)
        (x, y, z), (zx, zy, zz) = _seeded_centres(boxes, depth, K, n, l, generator)
        R = random_rotations(N, n, generator)
        return _pack(x, y, z, w, h, l, R, [uw, uh, ul, zx, zy, zz])
    if name == "aspect":                                    # :199-269
        uw = U()
        w = _rescale(uw)
        h = torch.zeros_like(w); l = torch.zeros_like(w)
        aux = torch.zeros((N, 2))
        for i in range(N):
            r1 = torch.randperm(len(RATIOS), generator=generator)[0]
            r2 = torch.randperm(len(RATIOS), generator=generator)[0]
            h[i] = w[i] * RATIOS[r1]
            l[i] = w[i] * RATIOS[r2]
            aux[i, 0], aux[i, 1] = RATIOS[r1], RATIOS[r2]
        (x, y, z), (zx, zy, zz) = _seeded_centres(boxes, depth, K, n, l, generator)
        R = random_rotations(N, n, generator)
        return _pack(x, y, z, w, h, l, R, [uw, zero, zero, zx, zy, zz], aux)
    if name == "rotation":                                  # :272-336 == propose without a ground normal
        o = propose(boxes, depth, prior_mean, prior_std, K, n, None, generator)
        return dict(cubes=o["cubes"], noise=o["noise"], rot=o["rot"], aux=None)
    raise ValueError(name)


# ----------------------------------------------------------------------------------------------
# geometry  (cubercnn/util/math_util.py:142-245, ProposalNetwork/utils/spaces.py:192-245,
#            ProposalNetwork/utils/conversions.py:25-48)
# ----------------------------------------------------------------------------------------------
# local vertex signs for v0..v7: X <-> l (neg for 0,3,4,7), Y <-> h (neg for 0,1,4,5),
# Z <-> w (neg for 0,1,2,3)   (math_util.py:198-207)
_SX = torch.tensor([-1, 1, 1, -1, -1, 1, 1, -1], dtype=torch.float32)
_SY = torch.tensor([-1, -1, 1, 1, -1, -1, 1, 1], dtype=torch.float32)
_SZ = torch.tensor([-1, -1, -1, -1, 1, 1, 1, 1], dtype=torch.float32)


def cube_corners(cubes):
    """[..., 15] -> [..., 8, 3] camera-space corners (R @ local + centre)."""
    c = cubes[..., 0:3]
    w, h, l = cubes[..., 3], cubes[..., 4], cubes[..., 5]
    R = cubes[..., 6:15].reshape(cubes.shape[:-1] + (3, 3))
    sx, sy, sz = _SX.to(cubes.device), _SY.to(cubes.device), _SZ.to(cubes.device)
    local = torch.stack([sx * (l / 2)[..., None], sy * (h / 2)[..., None], sz * (w / 2)[..., None]], -2)
    v = R @ local                                  # [...,3,8]
    v = v + c[..., :, None]
    return v.transpose(-1, -2)


def project_corners(corners3d, K, W, H):
    """spaces.py:224-245: K @ X, divide by z (no z>0 guard), clamp to
    x in [int(-W/2+1), int(2W-1)], y in [int(-H/2+1), int(2H-1)].  -> [...,8,2]"""
    p = corners3d @ K.t()                          # [...,8,3]  == (K @ X^T)^T
    xy = p[..., :2] / p[..., 2:3]
    x = torch.clamp(xy[..., 0], int(-W / 2 + 1), int(W - 1 + W))
    y = torch.clamp(xy[..., 1], int(-H / 2 + 1), int(H - 1 + H))
    return torch.stack((x, y), dim=-1)


def boxes_from_corners(c2d):
    """conversions.py:25-48: tight xyxy box over the 8 clamped corners. [...,8,2] -> [...,4]"""
    return torch.stack((c2d[..., 0].min(-1).values, c2d[..., 1].min(-1).values,
                        c2d[..., 0].max(-1).values, c2d[..., 1].max(-1).values), dim=-1)


# ----------------------------------------------------------------------------------------------
# scores  (ProposalNetwork/scoring/scorefunction.py, ProposalNetwork/utils/utils.py:186-250)
# ----------------------------------------------------------------------------------------------
def iou_score(ref_box, boxes):
    """scorefunction.py:47-49 -> utils.py:186-192 -> detectron2 pairwise_iou (public formula):
    inter = prod(clamp(min(br) - max(tl), 0)); iou = inter > 0 ? inter / (a1 + a2 - inter) : 0."""
    a1 = (ref_box[2] - ref_box[0]) * (ref_box[3] - ref_box[1])
    a2 = (boxes[:, 2] - boxes[:, 0]) * (boxes[:, 3] - boxes[:, 1])
    wh = torch.min(ref_box[None, 2:], boxes[:, 2:]) - torch.max(ref_box[None, :2], boxes[:, :2])
    wh = wh.clamp(min=0)
    inter = wh.prod(dim=1)
    return torch.where(inter > 0, inter / (a1 + a2 - inter), torch.zeros(1, dtype=inter.dtype, device=inter.device))


def dims_score(prior_mean, prior_std, dims, ref_box, boxes):
    """scorefunction.py:144-160.  prior_* [3], dims [n,3] (w,h,l), ref_box [4], boxes [n,4]."""
    g = torch.exp(-1 / 2 * ((dims - prior_mean) / prior_std) ** 2).mean(1)
    r_ref = (ref_box[2] - ref_box[0]) / (ref_box[3] - ref_box[1])
    r = (boxes[:, 2] - boxes[:, 0]) / (boxes[:, 3] - boxes[:, 1])
    diff = torch.abs(r_ref - r)
    return (1 - diff / torch.max(diff)) * g


def mask_rect(mask_u8):
    """scorefunction.py:59-68: largest external contour -> minAreaRect -> boxPoints.
    mask_u8 [H,W] uint8 (0/1).  Returns float32 [4,2] or None when the mask has no contour."""
    contours, _ = cv2.findContours(mask_u8, cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_SIMPLE)
    if not contours:
        return None
    biggest = max(contours, key=cv2.contourArea)
    return cv2.boxPoints(cv2.minAreaRect(biggest))


def fallback_rect(c2d):
    """scorefunction.py:69-75.  The callers pass bube_corners as [1,n,8,2] (roi_heads.py:465,495), so
    `bube_corners[:,:,0]` is VERTEX 0 of every proposal ([1,n,2]), `.min(1)[0]` the minimum over the n
    PROPOSALS of its (x, y), and `.mean()` the mean of those two numbers; likewise vertex 1 for the
    'y' values.  Reproduced as is (the comment there says 'mean of the bube corners'; the arithmetic is not).
    c2d [n,8,2] -> float32 [4,2]."""
    b = c2d[None]
    x0 = b[:, :, 0].min(1)[0].mean().numpy(); x1 = b[:, :, 0].max(1)[0].mean().numpy()
    y0 = b[:, :, 1].min(1)[0].mean().numpy(); y1 = b[:, :, 1].max(1)[0].mean().numpy()
    return np.array([[x0, y0], [x1, y0], [x1, y1], [x0, y1]])


def corners_score(mask, c2d, rect=None):
    """scorefunction.py:58-85 + :51-56.  mask [H,W] bool/uint8 tensor, c2d [n,8,2] fp32.
    Per proposal: mean over the 4 rectangle points of the (fp64, cKDTree) distance to the nearest
    projected corner, stored into an fp32 tensor; then 1 - d / max(d)."""
    if rect is None:
        rect = mask_rect(mask.numpy().astype(np.uint8))
    if rect is None:
        rect = fallback_rect(c2d)
    d = torch.zeros(len(c2d))
    pts = c2d.numpy()
    for i in range(len(pts)):
        dist, _ = cKDTree(pts[i]).query(rect)
        d[i] = np.mean(dist)
    return 1 - d / torch.max(d)


def _poly_mask(shape, pts8):
    """scorefunction.py:97-103: cv2.convexHull(8 fp32 pts) -> int32 truncation -> fillPoly."""
    canvas = np.zeros(shape, dtype="uint8")
    hull = cv2.convexHull(np.array(pts8))
    cv2.fillPoly(canvas, np.array([hull], dtype=np.int32), 1)
    return canvas


def segmentation_score(mask, c2d, power5=False):
    """scorefunction.py:88-106 (power5=False, utils.py:230-239 mask_iou) and :108-126
    (power5=True, utils.py:241-250 mod_mask_iou with int64 inter**5 wraparound), both on the
    [::4, ::4] grid.  mask [H,W] bool tensor; c2d [n,8,2]."""
    out = torch.zeros(len(c2d))
    m4 = mask[::4, ::4]
    for i in range(len(c2d)):
        poly = torch.tensor(_poly_mask(tuple(mask.shape), c2d[i])[::4, ::4])
        inter = (m4 * poly).sum()
        if inter == 0:
            out[i] = 0.0
            continue
        union = torch.logical_or(m4, poly).to(torch.int).sum()
        out[i] = (inter ** 5 / union) if power5 else (inter / union)
    return out


def point_cloud_score(points, cubes):
    """scorefunction.py:9-43.  The reference indexes `verts[:, 0].min(1)` on verts [n,8,3]: that is
    VERTEX 0's (x,y,z) triple reduced over its three coordinates, not the x coordinate over the 8
    vertices.  So the bounds actually used are: x in (min(v0), max(v0)), y in (min(v1), max(v1)),
    z in (min(v2), max(v2)), where min/max run over the xyz of one vertex.  Reproduced as is (the
    docstring there says AABB; the arithmetic does not).  Strict inequalities, int64 count.
    points [P,3] (fp32 or fp64), cubes [n,15] -> int64 [n]."""
    v = cube_corners(cubes)                                 # [n,8,3]
    lo = v[:, 0:3, :].min(2).values                         # [n,3]: per vertex v0,v1,v2
    hi = v[:, 0:3, :].max(2).values
    px, py, pz = points[:, 0:1], points[:, 1:2], points[:, 2:3]
    inside = ((px > lo[:, 0]) & (px < hi[:, 0]) & (py > lo[:, 1]) & (py < hi[:, 1])
              & (pz > lo[:, 2]) & (pz < hi[:, 2]))
    return inside.sum(0)


def select(iou, dim, cor):
    """roi_heads.py:469-472 / :499-502: fp32 NumPy product (iou*dim)*corners, np.argmax
    (first maximum; a NaN is a maximum, so the first NaN wins)."""
    combined = np.array(iou) * np.array(dim) * np.array(cor)
    return combined, int(np.argmax(combined))


# ----------------------------------------------------------------------------------------------
# back-projection + RANSAC plane + up-vector rectification
#   (cubercnn/modeling/roi_heads/roi_heads.py:345-428 NumPy flavour, :1084-1136 torch flavour;
#    ProposalNetwork/utils/plane.py)
# ----------------------------------------------------------------------------------------------
def backproject_np(depth, K, ground=None, stride=5):
    """roi_heads.py:346-384 (fp64 NumPy flavour): u,v index the SUBSAMPLED grid while fx,fy,cx,cy
    are the full-resolution intrinsics.  depth [H,W] fp32 ndarray, K [3,3] ndarray.
    Returns (plane_points [Pg,3], points_no_ground [P-Pg,3]) as float64."""
    dp = np.asarray(depth)[::stride, ::stride]
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    u, v = np.meshgrid(np.arange(dp.shape[1]), np.arange(dp.shape[0]))
    z = np.array(dp)
    x = (u - cx) * z / fx
    y = (v - cy) * z / fy
    if ground is None:
        pts = np.stack((x, y, z), axis=-1).reshape(-1, 3)
        return pts, pts
    g = np.asarray(ground)[::stride, ::stride]
    on = g > 0
    off = g == 0
    pg = np.stack((x[on], y[on], z[on]), axis=-1).reshape(-1, 3)
    pn = np.stack((x[off], y[off], z[off]), axis=-1).reshape(-1, 3)
    return pg, pn


def backproject_torch(depth, K, ground=None, stride=5):
    """roi_heads.py:1084-1112 (fp32 torch flavour): fx,fy = K // stride, cx,cy = centre of the
    subsampled grid.  Returns plane points [Pg,3] fp32 tensor."""
    dp = depth[::stride, ::stride]
    fx = K[0, 0] // stride
    fy = K[1, 1] // stride
    h, w = dp.shape
    cx, cy = w / 2, h / 2
    u = torch.arange(w, dtype=torch.float32)[None, :].expand(h, w)
    v = torch.arange(h, dtype=torch.float32)[:, None].expand(h, w)
    z = dp
    x = (u - cx) * z / fx
    y = (v - cy) * z / fy
    if ground is not None:
        on = ground[::stride, ::stride] > 0
        return torch.stack((x[on], y[on], z[on]), dim=-1).reshape(-1, 3)
    return torch.stack((x, y, z), dim=-1).reshape(-1, 3)


def draw_sample_ids(n_points, iters, rng):
    """plane.py:95 / :171: python `random.sample(range(n_points), 3)` per hypothesis."""
    return np.array([rng.sample(range(0, n_points), 3) for _ in range(iters)], dtype=np.int64)


def ransac_np(pts, ids, thresh=0.05):
    """plane.py:154-209 Plane_np.fit (== pyransac3d.Plane.fit used at roi_heads.py:374-376), with
    the sample ids given.  fp64 (or whatever dtype `pts` has); the sequential form keeps the first
    hypothesis with a strictly larger inlier count.  Returns (eq[4], inlier_idx, best_iter)."""
    best_eq, best_in, best_it = [], [], -1
    for it in range(len(ids)):
        s = pts[ids[it]]
        c = np.cross(s[1] - s[0], s[2] - s[0])
        c = c / np.linalg.norm(c)
        k = -np.sum(np.multiply(c, s[1]))
        dist = (c[0] * pts[:, 0] + c[1] * pts[:, 1] + c[2] * pts[:, 2] + k) / np.sqrt(
            c[0] ** 2 + c[1] ** 2 + c[2] ** 2)
        inl = np.where(np.abs(dist) <= thresh)[0]
        if len(inl) > len(best_in):
            best_eq, best_in, best_it = [c[0], c[1], c[2], k], inl, it
    return best_eq, best_in, best_it


def ransac_parallel(pts, ids, thresh=0.05):
    """plane.py:79-134 Plane.fit_parallel (fp32 torch), sample ids given.  Returns
    (-eq[4], inlier_idx, best_iter, counts[iters])."""
    s = pts[torch.as_tensor(ids)]
    c = torch.linalg.cross(s[:, 1] - s[:, 0], s[:, 2] - s[:, 0], dim=-1)
    c = c / torch.norm(c, p=2, dim=1, keepdim=True)
    k = -torch.sum(c * s[:, 1], dim=1)
    dist = (c[:, 0:1] * pts[:, 0] + c[:, 1:2] * pts[:, 1] + c[:, 2:3] * pts[:, 2] + k[:, None]) / \
        torch.sqrt(c[:, 0] ** 2 + c[:, 1] ** 2 + c[:, 2] ** 2)[:, None]
    inl = torch.abs(dist) <= thresh
    counts = inl.sum(1)
    b = int(torch.argmax(counts))
    eq = torch.stack([c[b, 0], c[b, 1], c[b, 2], k[b]])
    return -eq, torch.where(inl[b])[0], b, counts


def rectify_normal(nv):
    """roi_heads.py:411-428 (NumPy) == :1120-1136 (torch): make the plane normal y-up."""
    nv = np.array(nv, dtype=np.asarray(nv).dtype)
    if abs(nv[2]) > abs(nv[1]):
        nv = np.array([nv[0], nv[2], -nv[1]])
    if abs(nv[0]) > abs(nv[1]):
        nv = np.array([-nv[2], nv[0], nv[1]])
    if nv[1] < 0:
        nv = nv * -1
    return nv


# ---- IoU3D / MABO ordering (MABO branch, roi_heads.py:277-281, :518-582) ----------------------------------------------
_FACES = ((0, 1, 2, 3), (4, 5, 6, 7), (0, 1, 5, 4), (1, 2, 6, 5), (2, 3, 7, 6), (3, 0, 4, 7))  # math_util.py:178-207


def _box_halfspaces(v):
    """8 corners [8,3] (fp64) -> rows [a,b,c,d] with a x + b y + c z + d <= 0 inside (outward unit normals)"""
    c = v.mean(0)
    rows = []
    for f in _FACES:
        p0 = v[f[0]]
        n = np.cross(v[f[1]] - p0, v[f[3]] - p0)
        n = n / np.linalg.norm(n)
        if n @ (c - p0) > 0:
            n = -n
        rows.append([n[0], n[1], n[2], -(n @ p0)])
    return np.array(rows)


def box3d_iou_np(gt_corners, corners):
    """iou_3d (utils.py:194-210): IoU3D of one box [8,3] against n boxes [n,8,3].  The reference calls
    pytorch3d.ops.box3d_overlap (absent here: unpinned third-party, exact convex intersection of two boxes); this is
    the same quantity in fp64 through SciPy's half-space intersection (Qhull) -- pinned by the reference's known answer
    ProposalNetwork/utils/tests/test_iou.py:3-27 (0.9944)."""
    from scipy.optimize import linprog
    from scipy.spatial import ConvexHull, HalfspaceIntersection
    g = np.asarray(gt_corners, dtype=np.float64)
    P = np.asarray(corners, dtype=np.float64)
    vol_g = abs(np.dot(g[1] - g[0], np.cross(g[3] - g[0], g[4] - g[0])))
    hg = _box_halfspaces(g)
    out = np.zeros(len(P))
    for i, b in enumerate(P):
        vol_b = abs(np.dot(b[1] - b[0], np.cross(b[3] - b[0], b[4] - b[0])))
        hs = np.vstack([hg, _box_halfspaces(b)])
        # Chebyshev centre: max r s.t. n.x + d + r <= 0
        res = linprog([0, 0, 0, -1], A_ub=np.hstack([hs[:, :3], np.ones((12, 1))]), b_ub=-hs[:, 3],
                      bounds=[(None, None)] * 3 + [(0, None)])
        vol = 0.0
        if res.status == 0 and res.x[3] > 1e-9:
            try:
                vol = ConvexHull(HalfspaceIntersection(hs, res.x[:3]).intersections).volume
            except Exception:
                vol = 0.0
        out[i] = vol / (vol_g + vol_b - vol)
    return out


def accumulate_scores(scores, iou3d):
    """roi_heads.py:277-281: IoU3D reordered by descending score, then its running maximum"""
    idx = np.argsort(scores)[::-1]
    return np.maximum.accumulate(np.array([iou3d[i] for i in idx]))


# ---- pseudo-GT record (tools/eval_boxes.py:637-655) ---------------------------------------------------------------
def postprocess_boxes(boxes, net_hw, out_h, out_w):
    """GeneralizedRCNN._postprocess -> detectron2 detector_postprocess as tools/eval_boxes.py:637-640 applies it to the
    GT boxes (detectron2 is absent here: public behaviour restated): x *= out_w / net_w, y *= out_h / net_h in fp32,
    clip to [0, out_w] x [0, out_h], then boxes with no positive area are DROPPED (`results[output_boxes.nonempty()]`)."""
    b = _postprocess_scaled(boxes, net_hw, out_h, out_w)
    return b[postprocess_keep(boxes, net_hw, out_h, out_w)]


def _postprocess_scaled(boxes, net_hw, out_h, out_w):
    b = torch.as_tensor(boxes, dtype=torch.float32).clone()
    sx, sy = out_w / net_hw[1], out_h / net_hw[0]
    b[:, 0::2] *= sx
    b[:, 1::2] *= sy
    b[:, 0].clamp_(min=0, max=out_w); b[:, 2].clamp_(min=0, max=out_w)
    b[:, 1].clamp_(min=0, max=out_h); b[:, 3].clamp_(min=0, max=out_h)
    return b


def postprocess_keep(boxes, net_hw, out_h, out_w):
    b = _postprocess_scaled(boxes, net_hw, out_h, out_w)
    return ((b[:, 2] - b[:, 0]) > 0) & ((b[:, 3] - b[:, 1]) > 0)
