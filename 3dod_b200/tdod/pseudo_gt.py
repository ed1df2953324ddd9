"""Pseudo-ground-truth generation on top of the batched path.

`forward_pseudo` is the batched form of the `pseudo_gt == 'pseudo'` branch of ROIHeads_Boxer._forward_cube
(cubercnn/modeling/roi_heads/roi_heads.py:453-489): one `pipeline.select` call for many frames instead of a Python loop
over objects, returning per frame the fields the reference stores on its `Instances` (:480-487).

`PseudoGTWriter` streams those winners into the Omni3D annotation schema the reference dumps in one piece at the end of
`tools/eval_boxes.py:622-662` (image record :646, annotation record :648-653) -- here frame by frame into a JSON-lines
part file with a resume point, so an interrupted 10k-frame run continues where it stopped; `finalize` writes the same
single JSON document ({info, categories, images, annotations}) the training data loader reads
(cubercnn/data/datasets.py:418-432)."""
import json
import os

import torch

from . import api, pipeline


def forward_pseudo(batch, n_prop, seed=0, iou_boxes=None, box_from_object0=False, obj_id_base=0, frame_id_base=0, **kw):
    """batch: pipeline.FrameBatch on the device.  Returns (fields, selection) where fields is a dict of device tensors
    over all N objects: pred_boxes [N,4], scores [N], pred_bbox3D [N,8,3], pred_center_cam [N,3], pred_dimensions [N,3],
    pred_pose [N,3,3], pred_center_2D [N,2] (roi_heads.py:480-487).

    iou_boxes: [N,4] boxes the IoU term is scored against (the reference projects the GT cube, :466); default = the
    2D boxes.  box_from_object0 reproduces the reference's quirk that the winner's 2D box is read from the proposals of
    the frame's FIRST object at the winning index (`pred_boxes[0][highest_score]`, :476) -- it needs every proposal's
    box in HBM (return_all); the default False returns the winner's own box.  Neither reaches the pseudo-GT file: its
    `bbox2D_proj` is the GT 2D box (`write_frames`)."""
    need_all = bool(box_from_object0)
    sel = pipeline.select(batch, n_prop, seed=seed, obj_id_base=obj_id_base, frame_id_base=frame_id_base,
                          return_all=need_all, iou_boxes=iou_boxes, **kw)
    cubes = sel.best_cube
    boxes = sel.best_box
    if box_from_object0:
        # first object of each object's frame: objects are stored frame by frame, so it is the first index with that frame id
        of = batch.obj_frame.long()
        first = torch.zeros(batch.n_frames, dtype=torch.long, device=of.device)
        idx = torch.arange(of.numel(), device=of.device)
        first.scatter_reduce_(0, of, idx, reduce="amin", include_self=False)
        boxes = sel.all.all_boxes[first[of], sel.best_idx.long()]
    fields = dict(pred_boxes=boxes, scores=sel.best_score, pred_bbox3D=api.cube_corners(cubes),
                  pred_center_cam=cubes[:, 0:3], pred_dimensions=cubes[:, 3:6], pred_pose=cubes[:, 6:15].reshape(-1, 3, 3),
                  pred_center_2D=(boxes[:, :2] + boxes[:, 2:]) / 2)
    return fields, sel


def _fields(cubes, boxes, scores):
    """the fields the reference stores on its `Instances` (roi_heads.py:480-487, :649-659)"""
    return dict(pred_boxes=boxes, scores=scores, pred_bbox3D=api.cube_corners(cubes), pred_center_cam=cubes[:, 0:3],
                pred_dimensions=cubes[:, 3:6], pred_pose=cubes[:, 6:15].reshape(-1, 3, 3),
                pred_center_2D=(boxes[:, :2] + boxes[:, 2:]) / 2)


def forward_pred(batch, n_prop, seed=0, obj_id_base=0, frame_id_base=0, **kw):
    """Batched form of the `use_pred_boxes` (AP evaluation) branch of _forward_cube (roi_heads.py:493-505, :646-659): the
    IoU term is scored against the 2D box itself, the winner's cube and score are kept, and `pred_boxes` is the input box
    (`instances_i.pred_boxes = gt_boxes`, :650).  Returns (fields, selection)."""
    sel = pipeline.select(batch, n_prop, seed=seed, obj_id_base=obj_id_base, frame_id_base=frame_id_base, **kw)
    return _fields(sel.best_cube, batch.boxes, sel.best_score), sel


def forward_learn(batch, n_prop, seed=0, obj_id_base=0, frame_id_base=0, **kw):
    """Batched form of the `pseudo_gt == 'learn'` branch (roi_heads.py:460-464): EVERY proposal is returned, its score
    the IoU of its projected box with the object's 2D box.  Returns (cubes [N,n,15], scores [N,n], selection)."""
    sel = pipeline.select(batch, n_prop, seed=seed, obj_id_base=obj_id_base, frame_id_base=frame_id_base, return_all=True,
                          **kw)
    return sel.all.all_cubes, sel.all.all_scores[:, 0], sel


def annotation(ann_id, image_id, dataset_id, category_id, category_name, bbox2d, center, dims, bbox3d, rot):
    """one record of tools/eval_boxes.py:648-653 (field names, constants and order as the reference writes them)"""
    return {'behind_camera': False, 'truncation': 0, 'bbox2D_proj': bbox2d, 'bbox2D_tight': -1, 'visibility': 1.0,
            'segmentation_pts': -1, 'lidar_pts': -1, 'valid3D': True, 'category_id': category_id,
            'category_name': category_name, 'id': ann_id, 'image_id': image_id, 'dataset_id': dataset_id,
            'depth_error': -1, 'center_cam': center, 'dimensions': dims, 'bbox3D_cam': bbox3d, 'R_cam': rot}


def image_record(image_id, width, height, file_path, K):
    """tools/eval_boxes.py:646"""
    return {'width': width, 'height': height, 'file_path': file_path, 'K': K, 'src_90_rotate': False,
            'src_flagged': False, 'incomplete': False, 'id': image_id, 'dataset_id': 12}


class PseudoGTWriter:
    """Append-only part file `<path>.part.jsonl`, two lines per frame -- the image record (plus the id range of its
    annotations) and the frame's annotation objects already serialised -- flushed and fsync'ed per `flush`; the image lines
    are the resume point (a frame counts only when both of its lines are complete).  Annotation ids continue from the part
    file (reference: one global counter starting at 1, eval_boxes.py:622).  `finalize` assembles the reference's single
    document from the serialised pieces without parsing them again."""

    def __init__(self, path, info, categories, resume=True):
        self.path = path
        self.part = path + ".part.jsonl"
        self.info, self.categories = info, categories
        self.done_images = []
        self.next_id = 1
        if resume and os.path.exists(self.part):
            good = 0
            with open(self.part, "rb") as f:
                while True:
                    l1 = f.readline()
                    l2 = f.readline()
                    if not l1 or not l2 or not l1.endswith(b"\n") or not l2.endswith(b"\n"):
                        break                      # torn tail of an interrupted run
                    try:
                        head = json.loads(l1)
                    except ValueError:
                        break
                    self.done_images.append(head["image"]["id"])
                    self.next_id = head["next_id"]
                    good += len(l1) + len(l2)
            with open(self.part, "r+b") as f:      # drop a torn tail
                f.truncate(good)
        elif os.path.exists(self.part):
            os.remove(self.part)
        self._f = open(self.part, "ab")

    def add_frame(self, image, anns):
        """image: image_record(...); anns: list of dicts from `annotation` WITHOUT ids (assigned here, in order)"""
        for a in anns:
            a['id'] = self.next_id
            self.next_id += 1
        body = json.dumps(anns)[1:-1]              # the objects, comma separated, without the list brackets
        self._f.write((json.dumps({"image": image, "next_id": self.next_id}) + "\n" + body + "\n").encode())
        self.done_images.append(image["id"])

    def flush(self):
        self._f.flush()
        os.fsync(self._f.fileno())

    def finalize(self, return_doc=True):
        """the reference's single document (eval_boxes.py:658-661): {categories, info, images, annotations}"""
        self.flush()
        self._f.close()
        tmp = self.path + ".tmp"
        with open(tmp, "w") as out:
            out.write('{"categories": ' + json.dumps(self.categories) + ', "info": ' + json.dumps(self.info) + ', "images": [')
            first = True
            with open(self.part, "r") as f:
                while True:
                    l1 = f.readline()
                    if not l1:
                        break
                    f.readline()
                    img = json.dumps(json.loads(l1)["image"])
                    out.write(("" if first else ", ") + img)
                    first = False
            out.write('], "annotations": [')
            first = True
            with open(self.part, "r") as f:
                while True:
                    if not f.readline():
                        break
                    body = f.readline().rstrip("\n")
                    if body:
                        out.write(("" if first else ", ") + body)
                        first = False
            out.write("]}")
        os.replace(tmp, self.path)
        os.remove(self.part)
        return json.load(open(self.path)) if return_doc else None


def postprocess_boxes(gt_boxes, net_hw, out_hw):
    """What tools/eval_boxes.py:637-640 does to the GT 2D boxes before they are written as `bbox2D_proj`:
    GeneralizedRCNN._postprocess -> detectron2 detector_postprocess rescales them from the network resolution
    net_hw = (h, w) to the original image size out_hw = (height, width) in fp32 (x *= width / w, y *= height / h),
    clips them to the image and flags boxes that lost their area (detectron2 drops those).
    gt_boxes [n,4] (any device) -> (boxes [n,4] fp32, keep [n] bool)"""
    b = gt_boxes.detach().to(torch.float32).clone()
    out_h, out_w = float(out_hw[0]), float(out_hw[1])
    b[:, 0::2] *= out_w / float(net_hw[1])
    b[:, 1::2] *= out_h / float(net_hw[0])
    b[:, 0::2] = b[:, 0::2].clamp(min=0, max=out_w)
    b[:, 1::2] = b[:, 1::2].clamp(min=0, max=out_h)
    keep = ((b[:, 2] - b[:, 0]) > 0) & ((b[:, 3] - b[:, 1]) > 0)
    return b, keep


def write_frames(writer, fields, obj_off, frame_meta, class_ids, class_names, gt_boxes=None, net_hw=None, dataset_id=12,
                 empty_boxes="reference"):
    """fields: forward_pseudo's dict (device tensors, N objects); obj_off: CSR [F+1]; frame_meta: list of
    dict(id, width, height, file_path, K) per frame; class_ids / class_names: per-object category id / name.

    gt_boxes [N,4] + net_hw = (h, w) of the network input: `bbox2D_proj` is the GROUND-TRUTH 2D box rescaled to the
    original image size and clipped (tools/eval_boxes.py:637-640, written at :648) -- the training loader reads it as the
    2D target when bbox2D_tight == -1 (cubercnn/data/datasets.py:108-109) -- NOT the winner's projected box.
    empty_boxes: a GT box that loses its area in the clip is dropped by detectron2, after which the reference's
    `zip(bboxes, gt_classes, centers, ...)` (:647) pairs the remaining boxes with the FIRST len(bboxes) cubes;
    "reference" reproduces that, "skip" drops the object's whole record instead.
    One device->host copy per field for the whole batch, then pure Python."""
    if gt_boxes is None or net_hw is None:
        raise ValueError("write_frames needs gt_boxes [N,4] and net_hw=(h, w): bbox2D_proj is the rescaled GT 2D box "
                         "(tools/eval_boxes.py:637-640), not the predicted one")
    if empty_boxes not in ("reference", "skip"):
        raise ValueError("empty_boxes must be 'reference' or 'skip'")
    host = {k: v.detach().cpu().tolist() for k, v in fields.items() if k != "pred_boxes"}
    done = set(writer.done_images)
    gt_boxes = gt_boxes.detach().cpu()
    for f, meta in enumerate(frame_meta):
        if meta["id"] in done:
            continue
        o0, o1 = obj_off[f], obj_off[f + 1]
        bx, keep = postprocess_boxes(gt_boxes[o0:o1], net_hw, (meta["height"], meta["width"]))
        if empty_boxes == "reference":
            pairs = list(zip(bx[keep].tolist(), range(o0, o1)))          # zip truncates to the surviving boxes
        else:
            pairs = [(bx[j].tolist(), o0 + j) for j in range(o1 - o0) if bool(keep[j])]
        anns = []
        for box, i in pairs:
            anns.append(annotation(None, meta["id"], dataset_id, class_ids[i], class_names[i], box,
                                   host["pred_center_cam"][i], host["pred_dimensions"][i], host["pred_bbox3D"][i],
                                   host["pred_pose"][i]))
        writer.add_frame(image_record(meta["id"], meta["width"], meta["height"], meta["file_path"], meta["K"]), anns)
    writer.flush()
