"""Tensor-level host API over libtdod (batched entry points).  Every function takes CUDA tensors, allocates
the outputs with torch on the same device, launches on torch's current stream and returns without synchronising."""
import ctypes as C

import torch

from . import lib as L

f32, f64, i32, i64, u8, i16 = torch.float32, torch.float64, torch.int32, torch.int64, torch.uint8, torch.int16


def _as_u8(t):
    """bool masks are reinterpreted (no copy) as the 1-byte-per-pixel layout the kernels read"""
    if t is None:
        return None
    if t.dtype == torch.bool:
        return t.view(torch.uint8)
    return t


def cube_corners(cubes):
    """Cubes.get_all_corners: [..., 15] -> [..., 8, 3]"""
    flat = cubes.reshape(-1, 15)
    out = torch.empty((flat.shape[0], 8, 3), dtype=f32, device=cubes.device)
    L.check(L.lib().tdod_cube_corners(L.ptr(flat, f32, "cubes"), flat.shape[0], L.ptr(out), L.stream_ptr()))
    return out.reshape(cubes.shape[:-1] + (8, 3))


def project_cubes(cubes, K, W, H, corners=True, boxes=True):
    """Cubes.get_bube_corners + cubes_to_box: [..., 15] -> ([..., 8, 2] | None, [..., 4] | None)"""
    flat = cubes.reshape(-1, 15)
    M = flat.shape[0]
    c2d = torch.empty((M, 8, 2), dtype=f32, device=cubes.device) if corners else None
    bx = torch.empty((M, 4), dtype=f32, device=cubes.device) if boxes else None
    Kc = K.reshape(9).contiguous()
    L.check(L.lib().tdod_project_cubes(L.ptr(flat, f32, "cubes"), M, L.ptr(Kc, f32, "K"), int(W), int(H),
                                       L.ptr(c2d), L.ptr(bx), L.stream_ptr()))
    lead = cubes.shape[:-1]
    return (c2d.reshape(lead + (8, 2)) if corners else None, bx.reshape(lead + (4,)) if boxes else None)


def project_cubes_indexed(cubes, K, cube_frame, W, H, corners=True, boxes=True):
    """project_cubes for cubes of different frames: cubes [M,15], K [F,3,3], cube_frame [M] int32 -> ([M,8,2] | None, [M,4] | None)"""
    M = cubes.shape[0]
    c2d = torch.empty((M, 8, 2), dtype=f32, device=cubes.device) if corners else None
    bx = torch.empty((M, 4), dtype=f32, device=cubes.device) if boxes else None
    L.check(L.lib().tdod_project_cubes_indexed(L.ptr(cubes, f32, "cubes"), M, L.ptr(K.reshape(-1, 9), f32, "K"),
                                               L.ptr(cube_frame, i32, "cube_frame"), int(W), int(H), L.ptr(c2d), L.ptr(bx),
                                               L.stream_ptr()))
    return c2d, bx


def score_iou(ref_box, boxes):
    n = boxes.shape[0]
    out = torch.empty(n, dtype=f32, device=boxes.device)
    L.check(L.lib().tdod_score_iou(L.ptr(ref_box.reshape(4), f32, "ref_box"), L.ptr(boxes, f32, "boxes"), n,
                                   L.ptr(out), L.stream_ptr()))
    return out


def score_dimensions(prior_mean, prior_std, dims, ref_box, boxes):
    n = boxes.shape[0]
    out = torch.empty(n, dtype=f32, device=boxes.device)
    if dims.dim() != 2 or dims.shape[1] != 3 or dims.stride(1) != 1:
        raise ValueError("dims must be [n,3] with unit inner stride")
    if not dims.is_cuda or dims.dtype != f32:
        raise ValueError("dims must be a CUDA fp32 tensor")
    L.check(L.lib().tdod_score_dimensions(L.ptr(prior_mean, f32, "prior_mean"), L.ptr(prior_std, f32, "prior_std"),
                                          dims.data_ptr(), dims.stride(0), L.ptr(ref_box.reshape(4), f32, "ref_box"),
                                          L.ptr(boxes, f32, "boxes"), n, L.ptr(out), L.stream_ptr()))
    return out


def score_corners(rect, corners2d):
    """rect: [4,2] CUDA fp32 or None (fallback rectangle of scorefunction.py:69-75)"""
    n = corners2d.shape[0]
    out = torch.empty(n, dtype=f32, device=corners2d.device)
    L.check(L.lib().tdod_score_corners(L.ptr(rect, f32, "rect"), 0 if rect is None else 1,
                                       L.ptr(corners2d, f32, "corners2d"), n, L.ptr(out), L.stream_ptr()))
    return out


_KB_WS = {}


def _kb_workspace(n, H, W, dev):
    """K-B scratch, cached per (device, stream, shape): the rectangle kernels allocate nothing, and asking torch for 0.2 GB
    on every call costs more than the small-batch kernels themselves.  A cached block is only ever used on the stream it
    was created for, so reuse is ordered like any other kernel of that stream."""
    wsb = int(L.lib().tdod_mask_rect_workspace_bytes(n, H, W))
    key = (dev.index if dev.index is not None else torch.cuda.current_device(), torch.cuda.current_stream(dev).cuda_stream)
    ws = _KB_WS.get(key)
    if ws is None or ws.numel() < wsb:
        ws = torch.empty(max(wsb, 16), dtype=u8, device=dev)
        _KB_WS[key] = ws
    return ws, wsb


def mask_rect(masks, mark=None, out=None, two_kernel=False):
    """masks [n,H,W] bool/uint8 -> (rects [n,4,2] fp32, has_rect [n] uint8).  One call of tdod_mask_rect: the fused path
    (bit stream in shared memory, extremes out, one warp per hull) when the mask fits, else pack + rectangle passes.
    two_kernel=True forces the latter through its two entry points; `mark("kb_mid")` is then called between them."""
    m = _as_u8(masks)
    n, H, W = m.shape
    if out is not None:
        rects, has = out                      # preallocated [n,4,2] fp32 / [n] uint8 (slices of larger tensors are fine)
    else:
        rects = torch.empty((n, 4, 2), dtype=f32, device=m.device)
        has = torch.empty(n, dtype=u8, device=m.device)
    if n == 0:
        return rects, has
    ws, wsb = _kb_workspace(n, H, W, m.device)
    if not two_kernel:
        L.check(L.lib().tdod_mask_rect(L.ptr(m, u8, "masks"), n, H, W, L.ptr(rects), L.ptr(has), L.ptr(ws), wsb, L.stream_ptr()))
        return rects, has
    L.check(L.lib().tdod_mask_pack(L.ptr(m, u8, "masks"), n, H, W, L.ptr(ws), wsb, L.stream_ptr()))
    if mark is not None:
        mark("kb_mid")
    L.check(L.lib().tdod_mask_rect_packed(n, H, W, L.ptr(rects), L.ptr(has), L.ptr(ws), wsb, L.stream_ptr()))
    return rects, has


def mask_rect_bitmap(bitmap, W, out=None):
    """masks already packed 1 bit/px: bitmap [n,H,ceil(W/32)] int32 (bit c & 31 of word c >> 5 = pixel c) ->
    (rects [n,4,2], has_rect [n])"""
    n, H, Wp = bitmap.shape
    if Wp != (W + 31) // 32 or bitmap.dtype != i32:
        raise ValueError("bitmap must be int32 [n,H,ceil(W/32)]")
    if out is not None:
        rects, has = out
    else:
        rects = torch.empty((n, 4, 2), dtype=f32, device=bitmap.device)
        has = torch.empty(n, dtype=u8, device=bitmap.device)
    if n == 0:
        return rects, has
    ws, wsb = _kb_workspace(n, H, W, bitmap.device)
    L.check(L.lib().tdod_mask_rect_bitmap(L.ptr(bitmap, i32, "bitmap"), n, H, W, L.ptr(rects), L.ptr(has), L.ptr(ws), wsb,
                                          L.stream_ptr()))
    return rects, has


def score_segmentation(mask, corners2d, power5=False):
    m = _as_u8(mask)
    H, W = m.shape
    n = corners2d.shape[0]
    out = torch.empty(n, dtype=f32, device=corners2d.device)
    L.check(L.lib().tdod_score_segmentation(L.ptr(m, u8, "mask"), H, W, L.ptr(corners2d, f32, "corners2d"), n,
                                            1 if power5 else 0, L.ptr(out), L.stream_ptr()))
    return out


def score_point_cloud(points, cubes):
    if points.dtype not in (f32, f64):
        raise ValueError("points must be fp32 or fp64")
    n = cubes.shape[0]
    out = torch.empty(n, dtype=i64, device=cubes.device)
    L.check(L.lib().tdod_score_point_cloud(L.ptr(points, None, "points"), 1 if points.dtype == f64 else 0,
                                           points.shape[0], L.ptr(cubes, f32, "cubes"), n, L.ptr(out), L.stream_ptr()))
    return out


def box3d_iou(gt_corners, corners, want_vol=False):
    """iou_3d: gt_corners [8,3] with corners [n,8,3] -> [n]; or batched [G,8,3] with [G,n,8,3] -> [G,n]"""
    single = gt_corners.dim() == 2
    g = gt_corners.reshape(-1, 8, 3)
    c = corners.reshape(g.shape[0], -1, 8, 3)
    G, n = c.shape[0], c.shape[1]
    out = torch.empty((G, n), dtype=f32, device=c.device)
    vol = torch.empty((G, n), dtype=f32, device=c.device) if want_vol else None
    L.check(L.lib().tdod_box3d_iou(L.ptr(g.contiguous(), f32, "gt_corners"), L.ptr(c.contiguous(), f32, "corners"), G, n,
                                   L.ptr(out), L.ptr(vol), L.stream_ptr()))
    if single:
        out, vol = out[0], (vol[0] if want_vol else None)
    return (out, vol) if want_vol else out


def segmented_order(scores, k=None, want_val=False):
    """K-D: scores [S,n] fp32 -> descending order per row, first k: idx [S,k] int32 (+ values)"""
    S, n = scores.shape
    k = n if k is None else int(k)
    idx = torch.empty((S, k), dtype=i32, device=scores.device)
    val = torch.empty((S, k), dtype=f32, device=scores.device) if want_val else None
    L.check(L.lib().tdod_segmented_order(L.ptr(scores, f32, "scores"), S, n, k, L.ptr(idx), L.ptr(val), L.stream_ptr()))
    return (idx, val) if want_val else idx


def accumulate_scores(scores, iou3d, n_out=None):
    """accumulate_scores (roi_heads.py:277-281): scores [S,C,n] fp64, iou3d [S,n] fp32 -> curves [S,C,n_out] fp32"""
    S, Cn, n = scores.shape
    n_out = n if n_out is None else int(n_out)
    out = torch.empty((S, Cn, n_out), dtype=f32, device=scores.device)
    L.check(L.lib().tdod_accumulate_scores(L.ptr(scores, f64, "scores"), L.ptr(iou3d, f32, "iou3d"), S, Cn, n, n_out,
                                           L.ptr(out), L.stream_ptr()))
    return out


def torch_std(x):
    """torch.std(x, dim=1) of a contiguous fp32 [N,n] CUDA tensor with the rounding of torch's own CUDA kernel"""
    N, n = x.shape
    out = torch.empty(N, dtype=f32, device=x.device)
    L.check(L.lib().tdod_torch_std(L.ptr(x, f32, "x"), N, n, L.ptr(out), L.stream_ptr()))
    return out


def yaw_table(normals):
    """normals [F,3] fp32/fp64 -> [F,36,3,3] fp32"""
    F = normals.shape[0]
    out = torch.empty((F, 36, 3, 3), dtype=f32, device=normals.device)
    L.check(L.lib().tdod_yaw_table(L.ptr(normals, None, "normals"), 1 if normals.dtype == f64 else 0, F, L.ptr(out),
                                   L.stream_ptr()))
    return out


def backproject(depth, ground, K, stride=5, flavour=L.BACKPROJECT_NUMPY, want_other=True):
    """depth [F,H,W] fp32, ground [F,H,W] bool/u8 or None, K [F,3,3] ->
    (plane_pts [F,3,P], other_pts [F,3,P] | None, counts [F,2] int32 (a view of the [2,F] buffer)); fp64 for the
    NumPy flavour."""
    F, H, W = depth.shape
    P = ((H + stride - 1) // stride) * ((W + stride - 1) // stride)
    dt = f64 if flavour == L.BACKPROJECT_NUMPY else f32
    pts = torch.empty((F, 3, P), dtype=dt, device=depth.device)
    other = torch.empty((F, 3, P), dtype=dt, device=depth.device) if (want_other and ground is not None) else None
    counts2 = torch.empty((2, F), dtype=i32, device=depth.device)   # [2,F] in memory; returned as its [F,2] view, so
    counts = counts2.t()                                            # counts[:, 0] is contiguous (n_pts of ransac_plane)
    g = _as_u8(ground)
    L.check(L.lib().tdod_backproject(L.ptr(depth, f32, "depth"), L.ptr(g, u8, "ground"),
                                     L.ptr(K.reshape(F, 9).contiguous(), f32, "K"), F, H, W, stride, flavour,
                                     L.ptr(pts), L.ptr(other), L.ptr(counts2), L.stream_ptr()))
    return pts, other, counts


def ransac_plane(pts, n_pts, sample_ids=None, seed=0, iters=1000, thresh=0.05, want_mask=False, frame_id_base=0,
                 want_counts=False):
    """pts [F,3,P] fp32/fp64 SoA, n_pts [F] int32, sample_ids [F,iters,3] int32 or None (Philox).
    Returns dict(eq [F,4], best_iter [F], count [F], normal [F,3], mask [F,P] | None)."""
    F, _, P = pts.shape
    is64 = pts.dtype == f64
    dev = pts.device
    eq = torch.empty((F, 4), dtype=pts.dtype, device=dev)
    normal = torch.empty((F, 3), dtype=pts.dtype, device=dev)
    best = torch.empty(F, dtype=i32, device=dev)
    cnt = torch.empty(F, dtype=i32, device=dev)
    mask = torch.empty((F, P), dtype=u8, device=dev) if want_mask else None
    wsb = L.lib().tdod_ransac_workspace_bytes(F, iters, 1 if is64 else 0)
    ws = torch.empty(int(wsb), dtype=u8, device=dev)
    L.check(L.lib().tdod_ransac_plane(L.ptr(pts, None, "pts"), 1 if is64 else 0, L.ptr(n_pts, i32, "n_pts"), F, P,
                                      L.ptr(sample_ids, i32, "sample_ids"), int(seed), int(frame_id_base), int(iters),
                                      float(thresh),
                                      L.ptr(eq), L.ptr(best), L.ptr(cnt), L.ptr(normal), L.ptr(mask), L.ptr(ws), wsb,
                                      L.stream_ptr()))
    r = dict(eq=eq, best_iter=best, count=cnt, normal=normal, mask=mask)
    if want_counts:   # per-hypothesis inlier counts live in the workspace right after the (256-byte aligned) hypotheses
        hyp_bytes = (F * iters * (5 * pts.element_size()) + 255) & ~255
        r["counts"] = ws[hyp_bytes:hyp_bytes + F * iters * 4].view(i32).reshape(F, iters)
    return r


class PssOutputs:
    __slots__ = ("best_idx", "best_score", "best_cube", "best_box", "all_cubes", "all_boxes", "all_scores", "stats",
                 "noise", "yaw_idx", "seeds", "all_pc")


def propose_score_select(depth, K, boxes, prior_mean, prior_std, obj_frame, rects, has_rect, n_prop, yaw_tab=None,
                         iou_boxes=None, noise=None, yaw_idx=None, rot=None, seed=None, obj_id_base=0,
                         return_all=False, cluster_size=0, out=None, propose_only=False, centre_mode=L.CENTRE_DEPTH,
                         dims_mode=L.DIMS_PRIOR, aux=None, im_shape=None, want_noise=False, want_seeds=False,
                         std_mode=L.STD_EXACT, masks=None, points=None, n_points=None):
    """The fused hot path (csrc/fused.cu).  depth [F,H,W], K [F,3,3], boxes [N,4], priors [N,3], obj_frame [N] int32,
    rects [N,4,2], has_rect [N] uint8 | None, yaw_tab [F,36,3,3] | None.
    Parity RNG mode: noise [N,6,n] (+ yaw_idx [N,n] int16 | rot [N,n,9]); throughput mode: seed=int (Philox).
    Mode B (MABO branch, roi_heads.py:535-538): masks [N,H,W] adds the segmentation and mod-segmentation scores as
    channels 4, 5 of all_scores; points [F,3,P] (+ n_points [F] int32: the frames' non-ground points as `backproject`
    returns them) adds all_pc [N,n] int32.  Both imply return_all."""
    F, H, W = depth.shape
    N = boxes.shape[0]
    dev = depth.device
    o = out if out is not None else PssOutputs()
    if propose_only or masks is not None or points is not None:
        return_all = True
    n_ch = 6 if masks is not None else 4
    if out is None:
        o.best_idx = o.best_score = o.best_cube = o.best_box = None
        if not propose_only:
            o.best_idx = torch.empty(N, dtype=i32, device=dev)
            o.best_score = torch.empty(N, dtype=f32, device=dev)
            o.best_cube = torch.empty((N, 15), dtype=f32, device=dev)
            o.best_box = torch.empty((N, 4), dtype=f32, device=dev)
        o.all_cubes = torch.empty((N, n_prop, 15), dtype=f32, device=dev) if return_all else None
        o.all_boxes = torch.empty((N, n_prop, 4), dtype=f32, device=dev) if return_all else None
        o.all_scores = torch.empty((N, n_ch, n_prop), dtype=f32, device=dev) if (return_all and not propose_only) else None
        o.stats = torch.empty((N, 6), dtype=f32, device=dev) if return_all else None
    a = L.PssArgs()
    a.depth = L.ptr(depth, f32, "depth")
    a.K = L.ptr(K.reshape(F, 9), f32, "K")
    a.yaw_table = L.ptr(yaw_tab, f32, "yaw_tab")
    a.F, a.H, a.W = F, H, W
    a.n_obj, a.n_prop = N, int(n_prop)
    a.obj_frame = L.ptr(obj_frame, i32, "obj_frame")
    a.boxes = L.ptr(boxes, f32, "boxes")
    a.iou_boxes = L.ptr(iou_boxes, f32, "iou_boxes")
    a.prior_mean = L.ptr(prior_mean, f32, "prior_mean")
    a.prior_std = L.ptr(prior_std, f32, "prior_std")
    a.rects = L.ptr(rects, f32, "rects")
    a.has_rect = L.ptr(has_rect, u8, "has_rect")
    if seed is None:
        a.rng_mode = L.RNG_HOST
        if noise is None or tuple(noise.shape) != (N, 6, n_prop):
            raise ValueError("parity RNG mode needs noise [N,6,n_prop]")
        a.noise = L.ptr(noise, f32, "noise")
        a.yaw_idx = L.ptr(yaw_idx, i16, "yaw_idx")
        a.rot = L.ptr(rot, f32, "rot")
        a.seed = 0
    else:
        a.rng_mode = L.RNG_PHILOX
        a.seed = int(seed)
    a.obj_id_base = int(obj_id_base)
    a.best_idx, a.best_score = L.ptr(o.best_idx), L.ptr(o.best_score)
    a.best_cube, a.best_box = L.ptr(o.best_cube), L.ptr(o.best_box)
    a.all_cubes, a.all_boxes = L.ptr(o.all_cubes), L.ptr(o.all_boxes)
    a.all_scores, a.stats = L.ptr(o.all_scores), L.ptr(o.stats)
    a.cluster_size = int(cluster_size)
    a.propose_only = 1 if propose_only else 0
    # ablation samplers (proposals.py:20-336): where centres / dimensions come from; im_shape = (width, height)
    a.centre_mode, a.dims_mode = int(centre_mode), int(dims_mode)
    a.im_w, a.im_h = (int(im_shape[0]), int(im_shape[1])) if im_shape is not None else (0, 0)
    if aux is not None and tuple(aux.shape) != (N, 2):
        raise ValueError("aux must be [N,2]")
    a.aux = L.ptr(aux, f32, "aux")
    a.std_mode = int(std_mode)
    a.mode_flags = (L.MODEB_SEG if masks is not None else 0) | (L.MODEB_PC if points is not None else 0)
    a.score_channels = n_ch
    m8 = _as_u8(masks)
    if m8 is not None and tuple(m8.shape) != (N, H, W):
        raise ValueError("masks must be [N,H,W]")
    a.masks = L.ptr(m8, u8, "masks")
    o.all_pc = None
    if points is not None:
        if points.dim() != 3 or points.shape[0] != F or points.shape[1] != 3 or points.dtype not in (f32, f64):
            raise ValueError("points must be [F,3,P] fp32 / fp64 (SoA, as backproject returns them)")
        if n_points is None:
            raise ValueError("points need n_points [F] int32")
        a.points, a.n_points = L.ptr(points, None, "points"), L.ptr(n_points, i32, "n_points")
        a.P_cap, a.points_f64 = int(points.shape[2]), 1 if points.dtype == f64 else 0
        o.all_pc = torch.empty((N, n_prop), dtype=i32, device=dev)
        a.all_pc = L.ptr(o.all_pc)
    # audit outputs: the numbers behind the proposal set (HOST-mode layout) and the centre seeds
    o.noise = torch.empty((N, 6, n_prop), dtype=f32, device=dev) if want_noise else None
    o.yaw_idx = torch.empty((N, n_prop), dtype=i16, device=dev) if (want_noise and yaw_tab is not None) else None
    o.seeds = torch.empty((N, 3, n_prop), dtype=f32, device=dev) if want_seeds else None
    a.out_noise, a.out_yaw_idx, a.out_seeds = L.ptr(o.noise), L.ptr(o.yaw_idx), L.ptr(o.seeds)
    L.check(L.lib().tdod_propose_score_select(C.byref(a), L.stream_ptr()))
    return o


def _on_tensor_device(fn):
    """Every public call runs with the device of its first CUDA tensor argument current, so the launch, the stream
    (lib.stream_ptr) and the outputs all belong to that device whatever device the caller had selected."""
    import functools

    @functools.wraps(fn)
    def call(*args, **kw):
        for x in args + tuple(kw.values()):
            if isinstance(x, torch.Tensor) and x.is_cuda:
                if x.device.index == torch.cuda.current_device():
                    break
                with torch.cuda.device(x.device):
                    return fn(*args, **kw)
        return fn(*args, **kw)
    return call


for _name, _fn in list(globals().items()):
    if callable(_fn) and not _name.startswith("_") and getattr(_fn, "__module__", None) == __name__ and not isinstance(_fn, type):
        globals()[_name] = _on_tensor_device(_fn)
del _name, _fn
