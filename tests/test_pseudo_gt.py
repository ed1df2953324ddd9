"""Pseudo-GT wire format and resume: schema of tools/eval_boxes.py:646-661, append / torn-tail /
resume behaviour of the writer on CPU; the batched forward that feeds it is checked on the GPU against the per-object
path (roi_heads.py:453-489)."""
import json
import os

import numpy as np
import pytest
import torch

from tdod import pseudo_gt as PG

ANN_KEYS = ['behind_camera', 'truncation', 'bbox2D_proj', 'bbox2D_tight', 'visibility', 'segmentation_pts', 'lidar_pts',
            'valid3D', 'category_id', 'category_name', 'id', 'image_id', 'dataset_id', 'depth_error', 'center_cam',
            'dimensions', 'bbox3D_cam', 'R_cam']          # eval_boxes.py:648-653, in the reference's order
IMG_KEYS = ['width', 'height', 'file_path', 'K', 'src_90_rotate', 'src_flagged', 'incomplete', 'id', 'dataset_id']


def _fake_fields(n, seed):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.rand(*s, generator=g)
    return dict(pred_boxes=r(n, 4), scores=r(n), pred_bbox3D=r(n, 8, 3), pred_center_cam=r(n, 3),
                pred_dimensions=r(n, 3), pred_pose=r(n, 3, 3), pred_center_2D=r(n, 2))


def _meta(fid):
    return dict(id=1000 + fid, width=730, height=530, file_path="SUNRGBD/%05d.jpg" % fid, K=[[657.0, 0, 365.0], [0, 657.0, 265.0], [0, 0, 1]])


def _run(path, frames, resume, stop_after=None):
    w = PG.PseudoGTWriter(path, {"id": 12, "name": "SUNRGBD Train"}, [{"id": 18, "name": "chair"}], resume=resume)
    for k, fid in enumerate(frames):
        if stop_after is not None and k == stop_after:
            w.flush()
            return None
        n = 2 + fid % 3
        gt = torch.tensor([[10.0 + k, 20.0, 200.0 + k, 300.0] for k in range(n)])
        PG.write_frames(w, _fake_fields(n, fid), [0, n], [_meta(fid)], [18] * n, ["chair"] * n, gt_boxes=gt, net_hw=(530, 730))
    return w.finalize()


def test_schema_matches_the_reference_records(tmp_path):
    doc = _run(str(tmp_path / "a.json"), range(4), resume=False)
    assert list(doc) == ["categories", "info", "images", "annotations"]
    assert all(list(a) == ANN_KEYS for a in doc["annotations"])
    assert all(list(i) == IMG_KEYS for i in doc["images"])
    assert [a["id"] for a in doc["annotations"]] == list(range(1, len(doc["annotations"]) + 1))   # one global counter from 1
    a = doc["annotations"][0]
    assert np.array(a["bbox3D_cam"]).shape == (8, 3) and np.array(a["R_cam"]).shape == (3, 3)
    assert len(a["center_cam"]) == 3 and len(a["dimensions"]) == 3 and len(a["bbox2D_proj"]) == 4
    assert json.load(open(tmp_path / "a.json")) == doc
    assert not os.path.exists(str(tmp_path / "a.json") + ".part.jsonl")


def test_resume_after_interruption_gives_the_same_document(tmp_path):
    whole = _run(str(tmp_path / "w.json"), range(7), resume=False)
    p = str(tmp_path / "r.json")
    assert _run(p, range(7), resume=False, stop_after=3) is None
    with open(p + ".part.jsonl", "ab") as f:                 # a torn line, as a killed process leaves it
        f.write(b'{"image": {"id": 99, "wid')
    w = PG.PseudoGTWriter(p, {}, [], resume=True)
    assert w.done_images == [1000, 1001, 1002]
    w._f.close()
    resumed = _run(p, range(7), resume=True)                  # frames 0-2 are skipped, ids continue
    assert resumed == whole


def test_records_equal_the_reference_built_records(tmp_path):
    """tests/golden/pseudo_record.json holds the image + annotation records built by the UNMODIFIED statements of
    tools/eval_boxes.py:645-655 (make_golden.py pseudo_record): bbox2D_proj is the GT 2D box after
    GeneralizedRCNN._postprocess (:637-640), a box emptied by the clip is dropped and the zip truncates"""
    ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    g = json.load(open(os.path.join(ROOT, "tests", "golden", "pseudo_record.json")))
    cubes = torch.tensor(g["cubes"])
    n = cubes.shape[0]
    fields = dict(pred_boxes=torch.zeros(n, 4), scores=torch.zeros(n), pred_bbox3D=torch.tensor(g["corners"]),
                  pred_center_cam=cubes[:, :3], pred_dimensions=cubes[:, 3:6], pred_pose=cubes[:, 6:].reshape(n, 3, 3),
                  pred_center_2D=torch.zeros(n, 2))
    inp = g["input"]
    meta = dict(id=inp["image_id"], width=inp["width"], height=inp["height"], file_path=inp["file_name"][9:], K=inp["K"])
    cls = g["gt_classes"]
    w = PG.PseudoGTWriter(str(tmp_path / "p.json"), {}, [], resume=False)
    PG.write_frames(w, fields, [0, n], [meta], [100 + c for c in cls], [g["thing_classes"][c] for c in cls],
                    gt_boxes=torch.tensor(g["gt_boxes"]), net_hw=tuple(g["net_hw"]))
    doc = w.finalize()
    assert doc["images"] == [g["image"]]
    assert len(doc["annotations"]) == len(g["annotations"]) == 3          # one GT box is emptied by the clip
    for a, b in zip(doc["annotations"], g["annotations"]):
        assert list(a) == list(b)
        assert a == b, (a, b)


def test_write_frames_refuses_to_write_predicted_boxes():
    with pytest.raises(ValueError):
        PG.write_frames(None, _fake_fields(2, 0), [0, 2], [_meta(0)], [18, 18], ["chair"] * 2)


@pytest.mark.gpu
def test_batched_forward_equals_per_object_path():
    """forward_pseudo over several frames == roi_heads.py:453-489 frame by frame (same Philox keys), including the
    `pred_boxes[0][highest_score]` quirk for the 2D box"""
    from tdod import api, pipeline, synth
    d = torch.device("cuda:0")
    frames = synth.make_frames("C2", range(3))
    batch = pipeline.FrameBatch.from_frames(frames, d)
    off = np.cumsum([0] + [f.n_obj() for f in frames]).tolist()
    fields, sel = PG.forward_pseudo(batch, 1000, seed=4, box_from_object0=True)
    own, _ = PG.forward_pseudo(batch, 1000, seed=4, box_from_object0=False)
    torch.cuda.synchronize()
    assert torch.equal(own["pred_center_cam"], fields["pred_center_cam"])
    for f in range(3):
        b1 = pipeline.FrameBatch.from_frames([frames[f]], d)
        s1 = pipeline.select(b1, 1000, seed=4, obj_id_base=off[f], frame_id_base=f, return_all=True)
        torch.cuda.synchronize()
        sl = slice(off[f], off[f + 1])
        assert torch.equal(s1.best_idx, sel.best_idx[sl])
        assert torch.equal(s1.best_cube[:, :3], fields["pred_center_cam"][sl])
        assert torch.equal(api.cube_corners(s1.best_cube), fields["pred_bbox3D"][sl])
        quirk = s1.all.all_boxes[0, s1.best_idx.long()]                      # object 0's proposals at each winner's index
        assert torch.equal(quirk, fields["pred_boxes"][sl])
        eq = (s1.best_box == own["pred_boxes"][sl]) | (s1.best_box.isnan() & own["pred_boxes"][sl].isnan())
        assert bool(eq.all())


def _oracle_frame(f, n, gen_seed, iou_ref_boxes):
    """the oracle's chain for one frame with host-drawn randomness; iou_ref_boxes [N,4]: what the IoU term is scored
    against (the 2D boxes in the PRED branch, roi_heads.py:496; the projected GT cubes in the pseudo branch, :466)"""
    import random
    import tdod_oracle as O
    pg, _ = O.backproject_np(f.depth.numpy(), f.K.numpy(), f.ground.numpy())
    ids = O.draw_sample_ids(len(pg), 300, random.Random(gen_seed))
    eq, _, _ = O.ransac_np(pg, ids, 0.05)
    nv = torch.from_numpy(O.rectify_normal(np.array(eq[:3])))
    o = O.propose(f.boxes, f.depth, f.prior_mean, f.prior_std, f.K, n, nv, torch.Generator().manual_seed(gen_seed))
    c2d = O.project_corners(O.cube_corners(o["cubes"]), f.K, f.W, f.H)
    pb = O.boxes_from_corners(c2d)
    best, comb, iou = [], [], []
    for i in range(f.n_obj()):
        s_iou = O.iou_score(iou_ref_boxes[i], pb[i])
        s_dim = O.dims_score(f.prior_mean[i], f.prior_std[i], o["cubes"][i, :, 3:6], f.boxes[i], pb[i])
        s_cor = O.corners_score(f.masks[i], c2d[i])
        c, b = O.select(s_iou, s_dim, s_cor)
        best.append(b); comb.append(c); iou.append(s_iou.numpy())
    return o, ids, pb, np.array(best), np.stack(comb), np.stack(iou)


@pytest.mark.gpu
def test_batched_branches_against_the_oracle_chain():
    """forward_pseudo (roi_heads.py:466-489: IoU against the PROJECTED GT CUBE, the `pred_boxes[0][highest_score]` quirk for
    the 2D box), forward_pred (:493-505) and forward_learn (:460-464) against the oracle run object by object"""
    import tdod_oracle as O
    from tdod import pipeline, synth
    d = torch.device("cuda:0")
    n = 256
    f = synth.make_frame(4, 480, 640, 5)
    b = pipeline.FrameBatch.from_frames([f], d)
    # "ground-truth" cubes: plausible cubes in front of the camera, projected like cubes_to_box(gt_cubes[i], K, im_shape)
    g = torch.Generator().manual_seed(1)
    gt = torch.zeros(5, 15)
    gt[:, 0:2] = torch.rand(5, 2, generator=g) - 0.5
    gt[:, 2] = 3.0 + torch.rand(5, generator=g)
    gt[:, 3:6] = 0.5 + torch.rand(5, 3, generator=g)
    gt[:, 6:] = torch.eye(3).reshape(9)
    gt_boxes = O.boxes_from_corners(O.project_corners(O.cube_corners(gt), f.K, f.W, f.H))
    for branch, ref_boxes in (("pseudo", gt_boxes), ("pred", f.boxes)):
        o, ids, pb, best, comb, iou = _oracle_frame(f, n, 7, ref_boxes)
        kw = dict(noise=o["noise"].permute(1, 0, 2).contiguous().to(d), yaw_idx=o["yaw_idx"].to(torch.int16).to(d),
                  sample_ids=torch.from_numpy(ids.astype(np.int32))[None].contiguous().to(d), ransac_iters=300)
        if branch == "pseudo":
            fields, sel = PG.forward_pseudo(b, n, iou_boxes=gt_boxes.to(d), box_from_object0=True, **kw)
        else:
            fields, sel = PG.forward_pred(b, n, **kw)
        torch.cuda.synchronize()
        got = sel.best_idx.cpu().numpy()
        for i in range(5):
            top2 = np.sort(comb[i][~np.isnan(comb[i])])[-2:]
            if np.isnan(comb[i]).any() or (len(top2) == 2 and top2[1] - top2[0] > 1e-5 * abs(top2[1])):
                assert got[i] == best[i], (branch, i)
                assert torch.allclose(fields["pred_center_cam"][i].cpu(), o["cubes"][i, best[i], :3], rtol=1e-5, atol=1e-6)
                assert torch.equal(fields["pred_dimensions"][i].cpu(), o["cubes"][i, best[i], 3:6])
                if branch == "pseudo":        # the 2D box is read from OBJECT 0's proposals at the winner's index (:476)
                    assert torch.allclose(fields["pred_boxes"][i].cpu(), pb[0][best[i]], rtol=1e-5, atol=2e-3)
        if branch == "pred":
            assert torch.equal(fields["pred_boxes"].cpu(), f.boxes)          # instances_i.pred_boxes = gt_boxes (:650)
            cubes, scores, _ = PG.forward_learn(b, n, **kw)
            torch.cuda.synchronize()
            assert cubes.shape == (5, n, 15)
            assert np.allclose(scores.cpu().numpy(), iou, rtol=2e-5, atol=1e-6)
            assert torch.equal(cubes[..., 3:].cpu(), o["cubes"][..., 3:])
