"""ProposalNetwork/utils/utils.py -- the helpers that sit on the proposal-and-scoring path.

The random draws stay torch calls in the reference's program order (the proposal set is defined by
torch's generator streams); everything arithmetic goes through libtdod.  Every name a caller of the reference imports
from this module (roi_heads.py:34, cube_head.py:20, tools/eval_boxes.py:37, proposals.py:3, scorefunction.py:7;
list frozen in tests/golden/mirror_imports.json) resolves here: iou_3d is the exact box-box intersection kernel of
libtdod (the reference calls pytorch3d.ops.box3d_overlap, utils.py:194-210), the differentiable hull of the weak loss
(jarvis_march, fill_polygon, convex_hull, utils.py:424-507) is served by tdod.weak_loss, the overlay helper
show_mask2 is host NumPy like the reference's."""
import numpy as np
import torch

from tdod import api
from tdod.weak_loss import jarvis_march, fill_polygon, convex_hull  # noqa: F401  (utils.py:424-507)

N_YAW = 36


def _accepted_normals(means, stds, count, lo=None, hi=None):
    """The draw loop of sample_normal_in_range (utils.py:42-60) on standard normals: returns (samples, z) where
    samples = z*std+mean (separate mul, add == torch.normal(mean_t, std_t)) and z are the ACCEPTED standard
    normals.  torch.normal(mean_t, std_t) fills a [N,count] tensor with normal_(0,1) and then scales it, so
    torch.randn of the same shape consumes the generator identically."""
    dev = means.device
    N = means.shape[0]
    z = torch.randn((N, count), device=dev, dtype=means.dtype)
    s = z * stds[:, None] + means[:, None]
    if lo is not None and hi is not None:
        hi_b = hi[:, None].expand_as(s)
        tries = 0
        while True:
            bad = (s < lo) | (s > hi_b)
            if not bool(torch.any(bad)):
                break
            z2 = torch.randn((N, count), device=dev, dtype=means.dtype)
            s = torch.where(bad, z2 * stds[:, None] + means[:, None], s)
            z = torch.where(bad, z2, z)
            tries += 1
            if tries == 10000:      # utils.py:57: give up, leave the out-of-range samples in place
                break
    return s, z


def sample_normal_in_range(means, stds, count, threshold_low=None, threshold_high=None):
    """utils.py:42-60"""
    return _accepted_normals(means, stds, count, threshold_low, threshold_high)[0]


def randn_orthobasis_torch(num_samples=1, num_instances=1):
    """utils.py:62-69.  Drawn from torch's CPU generator on the host exactly like the reference (it has no device
    argument there); returns a CPU tensor [num_instances, num_samples, 3, 3] the caller moves to the device."""
    z = torch.randn(num_instances, num_samples, 3, 3)
    z = z / torch.norm(z, p=2, dim=-1, keepdim=True)
    z[:, :, 0] = torch.linalg.cross(z[:, :, 1], z[:, :, 2], dim=-1)
    z[:, :, 0] = z[:, :, 0] / torch.norm(z[:, :, 0], dim=-1, keepdim=True)
    z[:, :, 1] = torch.linalg.cross(z[:, :, 2], z[:, :, 0], dim=-1)
    z[:, :, 1] = z[:, :, 1] / torch.norm(z[:, :, 1], dim=-1, keepdim=True)
    return z


def orthobasis_from_normal_t(normal, yaw_angles=0):
    """utils.py:120-132 for the 36 yaw bins linspace(0, pi, 36) the sampler uses (proposals.py:405): [36,3,3] with
    columns (x_rot, normal, normal x x_rot), computed in the normal's dtype (fp64 from the NumPy plane fit)."""
    if not torch.is_tensor(yaw_angles) or yaw_angles.numel() != N_YAW:
        raise ValueError("the CUDA yaw table is built for the sampler's 36 bins linspace(0, pi, 36)")
    if not normal.is_cuda:
        raise ValueError("normal must be a CUDA tensor (no CPU path)")
    out_dtype = torch.promote_types(normal.dtype, torch.float32)
    return api.yaw_table(normal.reshape(1, 3).contiguous())[0].to(out_dtype)


def vectorized_linspace(start_tensor, end_tensor, number_of_steps):
    """utils.py:170-177: arange(n) * ((end-start)/(n-1)) + start per row (NOT torch.linspace)."""
    spacing = (end_tensor - start_tensor) / (number_of_steps - 1)
    ramp = torch.arange(0, number_of_steps, dtype=start_tensor.dtype, device=start_tensor.device)
    return ramp[None, :] * spacing[:, None] + start_tensor[:, None]


def gt_in_norm_range(range, gt):
    """utils.py:149-153"""
    return (gt - range[0]) / abs(range[1] - range[0])


def iou_2d(gt_box, proposal_boxes):
    """utils.py:186-192 (detectron2 pairwise_iou of one box against n): -> [n]"""
    g, p = gt_box.tensor, proposal_boxes.tensor
    if g.shape[0] != 1:
        raise ValueError("iou_2d on the CUDA path scores ONE reference box against n proposals")
    return api.score_iou(g.to(torch.float32).contiguous(), p.to(torch.float32).contiguous())


def mask_iou(segmentation_mask, bube_mask):
    """utils.py:230-239 on device tensors (the per-proposal rasterised form lives in score_segmentation)."""
    bube_mask = torch.as_tensor(bube_mask, device=segmentation_mask.device)
    inter = (segmentation_mask * bube_mask).sum()
    if inter == 0:
        return torch.tensor(0.0)
    return inter / torch.logical_or(segmentation_mask, bube_mask).to(torch.int).sum()


def mod_mask_iou(segmentation_mask, bube_mask):
    """utils.py:241-250 (int64 inter**5 wraps, as in the reference)"""
    bube_mask = torch.as_tensor(bube_mask, device=segmentation_mask.device)
    inter = (segmentation_mask * bube_mask).sum()
    if inter == 0:
        return torch.tensor(0.0)
    return inter ** 5 / torch.logical_or(segmentation_mask, bube_mask).to(torch.int).sum()


def mask_iou_loss(segmentation_mask, bube_mask):
    """utils.py:252-260: mask_iou on two tensors (no torch.tensor() re-wrap, so autograd flows through bube_mask);
    an empty intersection returns a fresh CPU 0.0 like the reference."""
    inter = (segmentation_mask * bube_mask).sum()
    if inter == 0:
        return torch.tensor(0.0)
    return inter / torch.logical_or(segmentation_mask, bube_mask).to(torch.int).sum()


def iou_3d(gt_cube, proposal_cubes):
    """utils.py:194-210: IoU3D of ONE ground-truth cube (Cubes [1,1,15]) with the n proposals of an object
    (Cubes [1,n,15]) -> [n].  The reference hands the two corner sets to pytorch3d.ops.box3d_overlap; here the
    exact intersection volume of the two oriented boxes comes from libtdod (tdod_box3d_iou)."""
    gt = gt_cube.get_all_corners()[0]
    pr = proposal_cubes.get_all_corners()[0]
    if gt.shape[0] != 1:
        raise ValueError("iou_3d scores ONE ground-truth cube against n proposals")
    return api.box3d_iou(gt[0].contiguous(), pr.contiguous())


# ---- small vector helpers of the sampler (utils.py:10-40, 112-146); torch ops on whatever device the input is on ----
def normalize_vector(v):
    """utils.py:10-16: v / max(|v|, 1e-8)"""
    mag = torch.sqrt(v.pow(2).sum())
    mag = torch.max(mag, torch.tensor([1e-8], device=v.device))
    return (v / mag.view(1, 1).expand(1, v.shape[0]))[0]


def cross_product(u, v):
    """utils.py:18-24"""
    return torch.stack((u[1] * v[2] - u[2] * v[1], u[2] * v[0] - u[0] * v[2], u[0] * v[1] - u[1] * v[0]))


def compute_rotation_matrix_from_ortho6d(poses):
    """utils.py:26-40: Gram-Schmidt of the two 3-vectors in `poses` [6] -> [3,3] with columns (x, y, z)"""
    x = normalize_vector(poses[0:3])
    z = normalize_vector(cross_product(x, poses[3:6]))
    y = cross_product(z, x)
    return torch.stack((x, y, z), dim=1)


def vec_perp_t(vec):
    """utils.py:112-118: a vector perpendicular to `vec`; not re-normalised on the vec[0] == 0 branch"""
    a, b, c = vec
    if a == 0:
        return torch.tensor([0, c, -b], device=vec.device)
    return normalize_vector(torch.tensor([b, -a, 0], device=vec.device))


def rotate_vector_t(v, k, theta):
    """utils.py:134-146: Rodrigues rotation of v about the unit vector k by the angles theta [m] -> [m,3]"""
    cos_t, sin_t = torch.cos(theta), torch.sin(theta)
    t1 = v * cos_t.unsqueeze(1)
    t2 = torch.cross(k, v.to(k.dtype), dim=0) * sin_t.unsqueeze(1)
    t3 = k * (k @ v.to(k.dtype)) * (1 - cos_t).unsqueeze(1)
    return t1 + t2 + t3


def custom_mapping(x, beta=1.7):
    """utils.py:212-228: S-shaped remap of values in [0,1]"""
    return [0.0 if xi <= 0 else 1 / (1 + (xi / (1 - xi)) ** (-beta)) for xi in x]


def show_mask2(masks, im, random_color=False):
    """utils.py:335-361 (tools/eval_boxes.py:37 imports it for the qualitative dumps): alpha-blend the masks
    [m,h,w] over the image [h,w,3] (0..255) -> [h,w,4] in 0..1.  Host NumPy, as in the reference."""
    h, w = im.shape[:2]
    rgba = np.concatenate((im, np.full((h, w, 1), 255.0)), axis=-1) / 255
    layer = np.zeros((h, w, 4))
    for i, m in enumerate(masks):
        colour = random_color[i] if isinstance(random_color, list) else np.concatenate([np.random.random(3), [0.6]])
        layer = layer + m.reshape(m.shape[-2], m.shape[-1], 1) * np.asarray(colour).reshape(1, 1, -1)
    hit = layer > 0
    return (rgba * ~hit + (0.5 * layer + 0.5 * (rgba * hit))).clip(0, 1)


def show_mask(mask, ax, random_color=False):
    """utils.py:326-333 (matplotlib axis supplied by the caller)"""
    colour = np.concatenate([np.random.random(3), [0.6]]) if random_color else np.array([30 / 255, 144 / 255, 1.0, 0.6])
    h, w = mask.shape[-2:]
    ax.imshow(mask.reshape(h, w, 1) * colour.reshape(1, 1, -1))
