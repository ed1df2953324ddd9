"""The MABO / IoU3D analysis branch of ROIHeads_Boxer._forward_cube (cubercnn/modeling/roi_heads/roi_heads.py:512-582,
`PLOT.EVAL MABO`) for a whole batch of frames: every proposal of every object gets all six scores and its IoU3D with the
object's ground-truth cube; each score's proposals are then ranked and the running maximum of IoU3D along that ranking is
the curve the reference plots (`accumulate_scores`, :277-281), for the eight single scores (:543-550) and, first element
only, the 26 score products (:553-579).

Device work: one `pipeline.select(mode_b=True)` (K-A, K-B, the fused K-C with its mode-B pass), one IoU3D launch, one
K-D launch per group of channels; the channel products are a handful of elementwise torch ops in NumPy's promotion rules
(fp32 x fp32 stays fp32, anything times the int64 point count becomes float64)."""
import torch

from . import api, pipeline

# order of the reference's score arrays (roi_heads.py:543-550)
CURVES = ("IoU2D", "point_c", "seg", "dim", "seg_mod", "corner", "combined", "random")
# the 26 products of roi_heads.py:553-579, left-to-right multiplication order as written there
COMBINATIONS = (
    ("IoU2D", "seg_mod"), ("IoU2D", "dim"), ("IoU2D", "corner"), ("IoU2D", "point_c"),
    ("IoU2D", "seg_mod", "dim"), ("IoU2D", "seg_mod", "corner"), ("IoU2D", "seg_mod", "point_c"),
    ("IoU2D", "dim", "corner"), ("IoU2D", "dim", "point_c"), ("IoU2D", "corner", "point_c"),
    ("IoU2D", "seg_mod", "point_c"), ("IoU2D", "seg_mod", "dim", "point_c"), ("IoU2D", "seg_mod", "corner", "point_c"),
    ("IoU2D", "dim", "corner", "point_c"), ("IoU2D", "seg_mod", "dim", "corner", "point_c"),
    ("seg_mod", "dim"), ("seg_mod", "corner"), ("seg_mod", "point_c"), ("seg_mod", "dim", "corner"),
    ("seg_mod", "dim", "point_c"), ("seg_mod", "corner", "point_c"), ("seg_mod", "dim", "corner", "point_c"),
    ("dim", "corner"), ("dim", "point_c"), ("dim", "corner", "point_c"), ("corner", "point_c"),
)


def numpy_product(terms):
    """np.array(a) * np.array(b) * ... left to right with NumPy's promotion: fp32 * fp32 -> fp32, fp32 * int64 -> float64"""
    acc = terms[0]
    for t in terms[1:]:
        if acc.dtype == torch.float32 and t.dtype == torch.float32:
            acc = acc * t
        else:
            acc = acc.to(torch.float64) * t.to(torch.float64)
    return acc


def score_channels(o):
    """PssOutputs of a mode-B select -> dict name -> [N,n] tensor in the dtype the reference's NumPy arrays have"""
    s = o.all_scores
    return {"IoU2D": s[:, 0], "dim": s[:, 1], "corner": s[:, 2], "combined": s[:, 3], "seg": s[:, 4], "seg_mod": s[:, 5],
            "point_c": o.all_pc.to(torch.int64)}


def mabo(batch, gt_cubes, n_prop, seed=0, random_scores=None, **kw):
    """batch: pipeline.FrameBatch; gt_cubes [N,15] (x y z w h l R) of the objects' ground-truth cubes.
    Returns dict: `iou3d` [N,n]; `curves` [N,8,n] in CURVES order; `combinations` [N,26]; `selection` (winners, all cubes /
    scores); `empty_percent` [N] = share of proposals with IoU3D == 0 (:640).  The IoU term is scored against the projected
    GT cube (roi_heads.py:532), as the reference does here."""
    dev = batch.device
    N = batch.n_obj
    K_obj = batch.K[batch.obj_frame.long()]
    H, W = int(batch.depth.shape[1]), int(batch.depth.shape[2])
    # cubes_to_box(gt_cubes[i], K, im_shape) per object (roi_heads.py:532): one launch, the frame of each object picks its K
    gt_boxes = api.project_cubes_indexed(gt_cubes.contiguous(), batch.K.contiguous(), batch.obj_frame, W, H, corners=False)[1]
    sel = pipeline.select(batch, n_prop, seed=seed, iou_boxes=gt_boxes, mode_b=True, **kw)
    o = sel.all
    iou3d = api.box3d_iou(api.cube_corners(gt_cubes.contiguous()), api.cube_corners(o.all_cubes))
    ch = score_channels(o)
    if random_scores is None:
        random_scores = torch.rand((N, n_prop), dtype=torch.float64, device=dev)      # np.random.rand in the reference
    ch["random"] = random_scores
    singles = torch.stack([ch[k].to(torch.float64) for k in CURVES], 1).contiguous()
    curves = api.accumulate_scores(singles, iou3d)
    prods = torch.stack([numpy_product([ch[k] for k in names]).to(torch.float64) for names in COMBINATIONS], 1).contiguous()
    comb = api.accumulate_scores(prods, iou3d, n_out=1)[:, :, 0]
    empty = ((iou3d == 0).sum(1).to(torch.float64) / n_prop * 100).to(torch.int64)     # int(count / size * 100)
    return dict(iou3d=iou3d, curves=curves, combinations=comb, selection=sel, channels=ch, empty_percent=empty,
                _K_obj=K_obj)
