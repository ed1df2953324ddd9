"""ProposalNetwork/scoring/scorefunction.py on the GPU: every score is one libtdod launch over the n proposals of
an object (the reference loops over proposals in Python with cv2 / cKDTree on the CPU)."""
import torch

from tdod import api
from ProposalNetwork.utils.utils import iou_2d


def _cuda_f32(t, name):
    if not torch.is_tensor(t) or not t.is_cuda:
        raise ValueError("%s must be a CUDA tensor (no CPU path)" % name)
    return t.to(torch.float32).contiguous()


def _corners(bube_corners):
    c = _cuda_f32(bube_corners, "bube_corners")
    if c.dim() == 4:
        if c.shape[0] != 1:
            raise ValueError("bube_corners must be [1,n,8,2] or [n,8,2]")
        c = c[0]
    return c


def _mask(segmentation_mask):
    m = segmentation_mask
    if not torch.is_tensor(m) or m.dim() != 2:
        raise ValueError("segmentation_mask must be a [H,W] tensor")
    if not m.is_cuda:
        # the reference hands score_* a CPU copy of a mask that was produced on the GPU (roi_heads.py:431); a caller
        # that keeps doing so pays one H2D copy here -- pass the CUDA mask to avoid it
        m = m.cuda()
    if m.dtype not in (torch.bool, torch.uint8):
        m = m != 0
    return m.contiguous()


def score_point_cloud(point_cloud, cubes, K=None, segmentation_mask=None):
    """scorefunction.py:9-43 -> int64 [n]"""
    t = cubes.tensor
    if t.shape[0] != 1:
        raise ValueError("score_point_cloud scores the proposals of ONE object (Cubes [1,n,15])")
    if not point_cloud.is_cuda:
        raise ValueError("point_cloud must be a CUDA tensor (no CPU path)")
    if point_cloud.dtype not in (torch.float32, torch.float64):
        point_cloud = point_cloud.to(torch.float32)
    return api.score_point_cloud(point_cloud.contiguous(), _cuda_f32(t[0], "cubes"))


def score_iou(gt_box, proposal_box):
    """scorefunction.py:47-49"""
    return iou_2d(gt_box, proposal_box)


def score_corners(segmentation_mask, bube_corners):
    """scorefunction.py:58-85: mask -> min-area rectangle of the largest contour (K-B), then the modified chamfer
    distance of every proposal's 8 projected corners to its 4 points, normalised by the per-object maximum."""
    c = _corners(bube_corners)
    rects, has = api.mask_rect(_mask(segmentation_mask)[None])
    code = int(has[0])
    if code == 3:
        raise RuntimeError("mask too fragmented for the shared-memory contour kernel (more than 3072 runs)")
    return api.score_corners(rects[0] if code == 1 else None, c)


def score_segmentation(segmentation_mask, bube_corners):
    """scorefunction.py:88-106"""
    return api.score_segmentation(_mask(segmentation_mask), _corners(bube_corners), power5=False)


def score_mod_segmentation(segmentation_mask, bube_corners):
    """scorefunction.py:108-126"""
    return api.score_segmentation(_mask(segmentation_mask), _corners(bube_corners), power5=True)


def score_dimensions(category, dimensions, gt_boxes, pred_boxes):
    """scorefunction.py:144-160.  category = (prior_mean [3], prior_std [3]); dimensions [n,3]"""
    mean, std = category
    d = dimensions
    if not d.is_cuda:
        raise ValueError("dimensions must be a CUDA tensor (no CPU path)")
    if d.dtype != torch.float32 or d.stride(-1) != 1:
        d = d.to(torch.float32).contiguous()
    return api.score_dimensions(_cuda_f32(mean, "prior_mean"), _cuda_f32(std, "prior_std"), d,
                                _cuda_f32(gt_boxes.tensor[0], "gt_boxes"), _cuda_f32(pred_boxes.tensor, "pred_boxes"))


def score_ratios(gt_box, pred_boxes):
    """scorefunction.py:164-169 (imported by the callers, never called): device torch ops"""
    diff = torch.abs(pred_boxes.tensor - gt_box.tensor[0]).sum(dim=1)
    return 1 - diff / torch.max(diff)
