"""ProposalNetwork/utils/utils.py -- the helpers that sit on the proposal-and-scoring path.

The random draws stay torch calls in the reference's program order (the proposal set is defined by
torch's generator streams); everything arithmetic goes through libtdod.  Plotting helpers and iou_3d (pytorch3d,
evaluation only) are out of scope and absent; the differentiable hull of the weak loss (jarvis_march, fill_polygon,
convex_hull, utils.py:424-507) is served by tdod.weak_loss."""
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
