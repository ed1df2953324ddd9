"""Weak-loss geometry of ROIHeads3DScore on the device: the differentiable consumers of
Cubes / cubes_to_box in the training path of the reference (cubercnn/modeling/roi_heads/roi_heads.py).

    project_boxes   Cubes -> 8 corners -> K -> clamp -> box (+ GIoU)         roi_heads.py:1571-1583, :1599
    z_loss          50-step depth sweep                                        :1151-1194
    segment_loss    jarvis_march + fill_polygon + focal / bce / dice           :1030-1052, utils.py:424-507
    normal_vector_from_maps  batched Plane.fit_parallel + wall rectification   :1076-1149
    pseudo_gt_z_box_loss / pseudo_gt_z_point_loss / depth_range_loss          :1196-1231, :1256-1305

Each is one or two kernel launches for the whole batch of cubes (the reference loops over cubes in Python and lets
autograd record hundreds of small kernels per cube); gradients come from hand-written backward kernels wrapped in
`torch.autograd.Function`, so the results plug into the reference's loss code unchanged.  CUDA tensors only."""
import math

import numpy as np
import torch

from . import lib as L
from . import pipeline

f32, f64, i32 = torch.float32, torch.float64, torch.int32


def _im_wh(im_sizes, n, dev):
    """[(w, h)] * n, one (w, h) pair, or an [n,2] tensor -> int32 [n,2] on the device"""
    if isinstance(im_sizes, torch.Tensor):
        t = im_sizes.to(device=dev, dtype=i32)
    else:
        a = np.asarray(im_sizes, dtype=np.float64).reshape(-1, 2)
        t = torch.as_tensor(a.astype(np.int32), device=dev)       # int(w): the reference truncates (spaces.py:241)
    if t.shape[0] == 1 and n != 1:
        t = t.expand(n, 2)
    if t.shape != (n, 2):
        raise ValueError("im_sizes must give one (width, height) per cube")
    return t.contiguous()


def _per_object_K(K, n):
    K = K.to(f32)
    if K.dim() == 2:
        K = K[None].expand(n, 3, 3)
    if K.shape != (n, 3, 3):
        raise ValueError("K must be [3,3] or [N,3,3]")
    return K.contiguous()


class _CubeBox(torch.autograd.Function):
    @staticmethod
    def forward(ctx, cubes, K, im_wh, gt):
        n, dev = cubes.shape[0], cubes.device
        cubes = cubes.detach().contiguous()
        corners = torch.empty((n, 8, 2), dtype=f32, device=dev)
        seg = torch.empty((n, 8, 2), dtype=f32, device=dev)
        boxes = torch.empty((n, 4), dtype=f32, device=dev)
        arg = torch.empty((n, 4), dtype=i32, device=dev)
        giou = torch.empty((n,), dtype=f32, device=dev) if gt is not None else None
        dgiou = torch.empty((n, 4), dtype=f32, device=dev) if gt is not None else None
        L.check(L.lib().tdod_cube_box_fwd(L.ptr(cubes, f32, "cubes"), L.ptr(K, f32, "K"), L.ptr(im_wh, i32, "im_sizes"),
                                          L.ptr(gt, f32, "gt_boxes"), n, L.ptr(corners), L.ptr(seg), L.ptr(boxes),
                                          L.ptr(arg), L.ptr(giou), L.ptr(dgiou), L.stream_ptr()))
        ctx.save_for_backward(cubes, K, im_wh, arg, dgiou)
        ctx.has_gt = gt is not None
        if gt is None:
            giou = torch.zeros((0,), dtype=f32, device=dev)
        ctx.mark_non_differentiable(arg)
        return corners, seg, boxes, giou, arg

    @staticmethod
    def backward(ctx, g_corners, g_seg, g_boxes, g_giou, _g_arg):
        cubes, K, im_wh, arg, dgiou = ctx.saved_tensors
        n = cubes.shape[0]
        c = lambda g: None if g is None else g.to(f32).contiguous()
        g_corners, g_seg, g_boxes = c(g_corners), c(g_seg), c(g_boxes)
        g_giou = c(g_giou) if ctx.has_gt else None
        out = torch.empty((n, 15), dtype=f32, device=cubes.device)
        L.check(L.lib().tdod_cube_box_bwd(L.ptr(cubes), L.ptr(K), L.ptr(im_wh), L.ptr(arg), n, L.ptr(g_corners),
                                          L.ptr(g_seg), L.ptr(g_boxes), L.ptr(g_giou), L.ptr(dgiou), L.ptr(out),
                                          L.stream_ptr()))
        return out, None, None, None


def project_boxes(cubes, K, im_sizes, gt_boxes=None):
    """cubes [N,15] (requires_grad allowed), K [N,3,3] or [3,3], im_sizes (w,h) per cube.  Returns a dict:
    corners [N,8,2] (get_bube_corners, spaces.py:224-245), seg_corners [N,8,2] (clamped to the image,
    roi_heads.py:1574-1575), boxes [N,4] (cubes_to_box, conversions.py:25-48) and, when gt_boxes [N,4] is given,
    loss_iou [N] = generalized_box_iou_loss(gt, boxes, reduction='none') (:1599).  All four are differentiable."""
    if cubes.dim() != 2 or cubes.shape[1] != 15:
        raise ValueError("cubes must be [N,15]")
    n, dev = cubes.shape[0], cubes.device
    Kc, im = _per_object_K(K, n), _im_wh(im_sizes, n, dev)
    gt = None if gt_boxes is None else gt_boxes.detach().to(f32).contiguous()
    if gt is not None and gt.shape != (n, 4):
        raise ValueError("gt_boxes must be [N,4]")
    corners, seg, boxes, giou, _ = _CubeBox.apply(cubes.to(f32), Kc, im, gt)
    out = dict(corners=corners, seg_corners=seg, boxes=boxes)
    if gt is not None:
        out["loss_iou"] = giou
    return out


def z_loss(gt_boxes, cubes, K, im_sizes, proj_boxes, max_count=50):
    """ROIHeads3DScore.z_loss (roi_heads.py:1151-1194) for all cubes in one launch.  The reference's value
    |z - (z +- v)| does not depend on z, so its gradient is identically zero; the result is returned detached."""
    n, dev = cubes.shape[0], cubes.device
    Kc, im = _per_object_K(K, n), _im_wh(im_sizes, n, dev)
    values = torch.linspace(0.0, (max_count - 1) / 10, max_count, device=dev)           # :1168
    out = torch.empty((n,), dtype=f32, device=dev)
    L.check(L.lib().tdod_z_sweep(L.ptr(gt_boxes.detach().to(f32).contiguous(), f32, "gt_boxes"),
                                 L.ptr(cubes.detach().to(f32).contiguous(), f32, "cubes"), L.ptr(Kc), L.ptr(im),
                                 L.ptr(proj_boxes.detach().to(f32).contiguous(), f32, "proj_boxes"), L.ptr(values),
                                 int(max_count), n, L.ptr(out), L.stream_ptr()))
    return out


def hull(corners):
    """jarvis_march (utils.py:424-468) of n x 8 points: (hull [n,8,2], meta [n,9] = m, corner index per vertex)"""
    n = corners.shape[0]
    pts = corners.detach().to(f32).contiguous()
    h = torch.empty((n, 8, 2), dtype=f32, device=pts.device)
    meta = torch.empty((n, 9), dtype=i32, device=pts.device)
    L.check(L.lib().tdod_hull(L.ptr(pts, f32, "corners"), n, L.ptr(h), L.ptr(meta), L.stream_ptr()))
    return h, meta


class _HullMoments(torch.autograd.Function):
    @staticmethod
    def forward(ctx, corners, masks, mask_idx, with_sigmoid):
        n, dev = corners.shape[0], corners.device
        H, W = masks.shape[-2:]
        h, meta = hull(corners)
        mom = torch.empty((n, 12), dtype=f64, device=dev)
        L.check(L.lib().tdod_hull_fill_moments(L.ptr(masks), L.ptr(mask_idx, i32, "mask_idx"), L.ptr(h), L.ptr(meta), n,
                                               int(H), int(W), int(bool(with_sigmoid)), L.ptr(mom), L.stream_ptr()))
        ctx.save_for_backward(masks, mask_idx, h, meta)
        return mom

    @staticmethod
    def backward(ctx, g_mom):
        masks, mask_idx, h, meta = ctx.saved_tensors
        n, dev = h.shape[0], h.device
        H, W = masks.shape[-2:]
        g = g_mom.to(f64).contiguous()
        ws = torch.empty((n, 24), dtype=f64, device=dev)
        out = torch.empty((n, 8, 2), dtype=f32, device=dev)
        L.check(L.lib().tdod_hull_fill_grad(L.ptr(masks), L.ptr(mask_idx), L.ptr(h), L.ptr(meta), n, int(H), int(W),
                                            L.ptr(g), L.ptr(ws), L.ptr(out), L.stream_ptr()))
        return out, None, None, None


def hull_fill_moments(masks, corners, mask_idx, with_sigmoid=True):
    """masks [M,H,W] bool / uint8 on the device, corners [n,8,2] (differentiable), mask_idx [n].  Returns [n,12] fp64:
    sum over pixels of t^k, k = 0..4, for mask == 0 then mask == 1, then sum sigmoid(t) for mask == 0, mask == 1, where
    t is the reference's soft hull mask convex_hull(mask, corners) (utils.py:504-507)."""
    if masks.dim() != 3:
        raise ValueError("masks must be [M,H,W]")
    if masks.dtype == torch.bool:
        masks = masks.view(torch.uint8)
    if masks.dtype != torch.uint8 or not masks.is_contiguous():
        raise ValueError("masks must be a contiguous bool / uint8 tensor")
    idx = torch.as_tensor(mask_idx, device=corners.device).to(i32).contiguous()
    if idx.numel() != corners.shape[0]:
        raise ValueError("one mask index per cube")
    return _HullMoments.apply(corners.to(f32), masks, idx, with_sigmoid)


class _DenseFill(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pts, H, W, march, m):
        n, dev = pts.shape[0], pts.device
        if march:
            h, meta = hull(pts)
        else:                                       # the polygon as given: vertex t is point t
            h = pts.detach().to(f32).contiguous()
            meta = torch.full((n, 9), -1, dtype=i32, device=dev)
            meta[:, 0] = m
            meta[:, 1:1 + m] = torch.arange(m, dtype=i32, device=dev)
        out = torch.empty((n, H, W), dtype=f32, device=dev)
        L.check(L.lib().tdod_hull_fill(L.ptr(h), L.ptr(meta), n, int(H), int(W), L.ptr(out), L.stream_ptr()))
        ctx.save_for_backward(h, meta)
        ctx.hw = (int(H), int(W))
        return out

    @staticmethod
    def backward(ctx, g):
        h, meta = ctx.saved_tensors
        n, dev = h.shape[0], h.device
        H, W = ctx.hw
        ws = torch.empty((n, 24), dtype=f64, device=dev)
        out = torch.empty((n, 8, 2), dtype=f32, device=dev)
        L.check(L.lib().tdod_hull_fill_dense_grad(L.ptr(h), L.ptr(meta), n, H, W, L.ptr(g.to(f32).contiguous()),
                                                  L.ptr(ws), L.ptr(out), L.stream_ptr()))
        return out, None, None, None, None


def jarvis_march(points):
    """utils.py:424-468 for one set of 8 points [8,2]: the hull vertices [m,2] in the reference's order (values only)"""
    h, meta = hull(points.reshape(1, 8, 2))
    return h[0, :int(meta[0, 0])]


def fill_polygon(mask, polygon):
    """utils.py:470-502: soft mask [h,w] of a polygon with at most 8 vertices [m,2]; differentiable in the vertices"""
    m = polygon.shape[0]
    if not 1 <= m <= 8:
        raise ValueError("fill_polygon: 1 to 8 vertices")
    pts = torch.cat((polygon.to(f32), polygon.new_zeros((8 - m, 2), dtype=f32)))[None]
    return _DenseFill.apply(pts, mask.shape[-2], mask.shape[-1], False, m)[0]


def convex_hull(mask, coords):
    """utils.py:504-507: fill_polygon(mask, jarvis_march(coords)) for coords [8,2] or a batch [n,8,2] (-> [n,h,w]);
    only the shape of `mask` is used, as in the reference.  segment_loss does not go through this dense form."""
    batch = coords.dim() == 3
    pts = coords.to(f32) if batch else coords.to(f32)[None]
    if pts.shape[1:] != (8, 2):
        raise ValueError("convex_hull: coords must be [8,2] or [n,8,2]")
    out = _DenseFill.apply(pts, mask.shape[-2], mask.shape[-1], True, 0)
    return out if batch else out[0]


def _focal_polynomials(alpha=0.25, gamma=2):
    """sigmoid_focal_loss(inputs=x, targets=t) for x in {0, 1} as polynomials in t (coefficients of t^0..t^4)"""
    P = np.polynomial.polynomial
    out = []
    for x in (0.0, 1.0):
        p = 1.0 / (1.0 + math.exp(-x))
        ce = [x + math.log1p(math.exp(-x)), -x]                  # (1 - t) x + log(1 + exp(-x))
        one_minus_pt = [p, -(2 * p - 1)]                         # 1 - (p t + (1 - p)(1 - t))
        alpha_t = [1 - alpha, 2 * alpha - 1]
        poly = P.polymul(P.polymul(alpha_t, ce), P.polypow(one_minus_pt, gamma))
        out.append(np.pad(poly, (0, 5 - len(poly))))
    return np.stack(out)


_FOCAL = _focal_polynomials()


def segment_loss(gt_mask, bube_corners, at_which_mask_idx, loss='focal'):
    """ROIHeads3DScore.segment_loss (roi_heads.py:1030-1052).  gt_mask: [M,H,W] device tensor, or the reference's list
    of [1,H,W] masks (stacked here); bube_corners [n,8,2] on the device, already clamped to the image (:1574-1575).
    'focal' / 'bce' return [n]; 'dice' one scalar over the batch (dice_loss :1013-1028).  As in the reference the mask
    is the `inputs` (logits) argument and the hull fill the `targets`."""
    if isinstance(gt_mask, (list, tuple)):
        gt_mask = torch.stack([m.reshape(m.shape[-2:]) for m in gt_mask])
    H, W = gt_mask.shape[-2:]
    if loss not in ('focal', 'bce', 'dice'):
        raise ValueError("loss must be 'focal', 'bce' or 'dice'")
    S = hull_fill_moments(gt_mask, bube_corners, at_which_mask_idx, with_sigmoid=(loss == 'dice'))   # [n,12] fp64
    if loss == 'focal':
        c = torch.as_tensor(_FOCAL.reshape(10), dtype=f64, device=S.device)
        return ((S[:, :10] * c).sum(1) / (H * W)).to(f32)
    if loss == 'bce':
        # x = 0: log 2;  x = 1: (1 - t) + log(1 + exp(-1))
        return ((math.log(2.0) * S[:, 0] + (1 + math.log1p(math.exp(-1.0))) * S[:, 5] - S[:, 6]) / (H * W)).to(f32)
    if loss == 'dice':
        inter, s_sum, y_sum = S[:, 11].sum(), S[:, 10].sum() + S[:, 11].sum(), S[:, 5].sum()
        return (1 - (2. * inter + 1) / (s_sum + y_sum + 1)).to(f32)


def _depth_stack(depth_maps):
    """[F,H,W] tensor, or the reference's iterable of equally sized [H,W] maps -> contiguous fp32 [F,H,W]"""
    if not isinstance(depth_maps, torch.Tensor):
        depth_maps = torch.stack([d.reshape(d.shape[-2:]) for d in depth_maps])
    if depth_maps.dim() != 3:
        raise ValueError("depth_maps must be [F,H,W]")
    return depth_maps.detach().to(f32).contiguous()


def depth_select(depth_maps, frame, boxes, mode, masks=None, mask_idx=None):
    """One depth order statistic per box in one launch (tdod_depth_select, include/tdod.h): (values [n], flags [n])."""
    depth = _depth_stack(depth_maps)
    F_, H, W = depth.shape
    dev = depth.device
    boxes = boxes.detach().to(device=dev, dtype=f32).contiguous()
    n = boxes.shape[0]
    frame = torch.as_tensor(frame, device=dev).to(i32).contiguous()
    if boxes.shape != (n, 4) or frame.shape != (n,):
        raise ValueError("boxes must be [n,4] with one frame index per box")
    if masks is not None:
        if masks.dtype == torch.bool:
            masks = masks.view(torch.uint8)
        if masks.dtype != torch.uint8 or not masks.is_contiguous() or tuple(masks.shape[-2:]) != (H, W):
            raise ValueError("masks must be a contiguous bool / uint8 [M,H,W] tensor of the depth maps' size")
        mask_idx = torch.as_tensor(mask_idx, device=dev).to(i32).contiguous()
        if mask_idx.shape != (n,):
            raise ValueError("one mask index per box")
    out = torch.empty((n,), dtype=f32, device=dev)
    flag = torch.empty((n,), dtype=i32, device=dev)
    L.check(L.lib().tdod_depth_select(L.ptr(depth, f32, "depth_maps"), F_, H, W, L.ptr(frame), L.ptr(boxes),
                                      L.ptr(masks), L.ptr(mask_idx), int(mode), n, L.ptr(out), L.ptr(flag),
                                      L.stream_ptr()))
    return out, flag


def _l1(vals, target):
    return torch.nn.functional.smooth_l1_loss(vals, target, reduction='none', beta=0.0)      # roi_heads.py:962-963


def pseudo_gt_z_box_loss(depth_maps, proposal_boxes, pred_z):
    """ROIHeads3DScore.pseudo_gt_z_box_loss (roi_heads.py:1196-1231): |pred_z - median depth under each proposal box|
    for all boxes of all frames in one launch.  depth_maps [F,H,W] (or F maps of one size), proposal_boxes: F tensors
    [Ni,4].  As in the reference the pseudo ground truth of a frame lists the boxes with area first and the boxes
    without (centre samples) after them, whatever their order in proposal_boxes."""
    counts = [int(b.shape[0]) for b in proposal_boxes]
    boxes = torch.cat([b.reshape(-1, 4) for b in proposal_boxes])
    frame = torch.repeat_interleave(torch.arange(len(counts), device=boxes.device),
                                    torch.as_tensor(counts, device=boxes.device))
    gt, has_area = depth_select(depth_maps, frame, boxes, 0)
    order = torch.argsort(frame * 2 + (1 - has_area), stable=True)                           # :1216-1227
    return _l1(pred_z, gt[order])


def pseudo_gt_z_point_loss(depth_maps, pred_xy, pred_z, num_boxes_per_image):
    """ROIHeads3DScore.pseudo_gt_z_point_loss (roi_heads.py:1256-1277): |pred_z - depth at the predicted centre|."""
    counts = torch.as_tensor(list(num_boxes_per_image), device=pred_xy.device)
    frame = torch.repeat_interleave(torch.arange(counts.numel(), device=pred_xy.device), counts)
    xy = pred_xy.detach().to(f32)
    gt, _ = depth_select(depth_maps, frame, torch.cat((xy, xy), 1), 1)
    return _l1(pred_z, gt)


def depth_extent(cubes):
    """max - min depth of the 8 corners of cubes [N,15] (roi_heads.py:1286-1288).  The value comes from the corner
    kernel behind Cubes.get_all_corners; the gradient from the same expression written with torch operators."""
    from ProposalNetwork.utils.spaces import Cubes
    t = cubes.to(f32)
    with torch.no_grad():
        z = Cubes(t.detach().reshape(-1, 1, 15)).get_all_corners()[:, 0, :, 2]
        value = z.max(1)[0] - z.min(1)[0]
    if not t.requires_grad:
        return value
    sx = t.new_tensor([-1, 1, 1, -1, -1, 1, 1, -1])                                          # math_util.py:198-207
    sy = t.new_tensor([-1, -1, 1, 1, -1, -1, 1, 1])
    sz = t.new_tensor([-1, -1, -1, -1, 1, 1, 1, 1])
    d, r2 = t[:, 3:6], t[:, 12:15]
    zt = (r2[:, 0:1] * sx * d[:, 2:3] + r2[:, 1:2] * sy * d[:, 1:2] + r2[:, 2:3] * sz * d[:, 0:1]) / 2 + t[:, 2:3]
    soft = zt.max(1)[0] - zt.min(1)[0]
    return value + (soft - soft.detach())


def depth_range_loss(gt_mask, at_which_mask_idx, depth_maps, cubes, gt_boxes, num_instances):
    """ROIHeads3DScore.depth_range_loss (roi_heads.py:1279-1305): |q90 - q10 of the depth under each object's mask -
    depth extent of its cube|.  gt_mask [M,H,W] (or the reference's list of [1,H,W]) at the depth maps' size (the
    reference's bilinear resize with align_corners=True is then the identity), cubes [N,15] or a Cubes of N x 1,
    gt_boxes [N,4] (tensor or Boxes), num_instances: cubes per frame."""
    if isinstance(gt_mask, (list, tuple)):
        gt_mask = torch.stack([m.reshape(m.shape[-2:]) for m in gt_mask])
    gt_mask = gt_mask.reshape(-1, *gt_mask.shape[-2:])
    t = cubes.tensor if hasattr(cubes, "tensor") else cubes
    t = t.reshape(-1, 15)
    boxes = gt_boxes.tensor if hasattr(gt_boxes, "tensor") else gt_boxes
    counts = torch.as_tensor(list(num_instances), device=t.device)
    frame = torch.repeat_interleave(torch.arange(counts.numel(), device=t.device), counts)
    gt, _ = depth_select(depth_maps, frame, boxes, 2, masks=gt_mask.contiguous(), mask_idx=at_which_mask_idx)
    return _l1(gt, depth_extent(t))


def normal_vector_from_maps(ground_maps, depth_maps, Ks, use_nth=5, seed=0, sample_ids=None):
    """ROIHeads3DScore.normal_vector_from_maps (roi_heads.py:1076-1149) for a stack of equally sized frames in two
    launches: depth_maps [F,H,W], ground_maps [F,H,W] or None, Ks [F,3,3] -> rectified up-vectors [F,3] fp32."""
    return pipeline.ground_normals(depth_maps, ground_maps, Ks, seed=seed, sample_ids=sample_ids, stride=use_nth,
                                   flavour=L.BACKPROJECT_TORCH)[0]
