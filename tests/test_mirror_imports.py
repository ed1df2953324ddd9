"""Row (b) of SURVEY.md section 8: every `ProposalNetwork` import statement of the reference's callers
(cubercnn/modeling/roi_heads/roi_heads.py:4,30-34, cube_head.py:17-20, cubercnn/data/generate_normal_vectors.py:54,
tools/eval_boxes.py:37, and the on-path modules' own cross imports) must resolve against the mirror in 3dod_b200/.
The list is frozen from /root/reference by tests/golden/make_mirror_imports.py (ast, no execution)."""
import importlib
import inspect
import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "3dod_b200"))
ROWS = json.load(open(os.path.join(ROOT, "tests", "golden", "mirror_imports.json")))


@pytest.mark.parametrize("row", ROWS, ids=[r["where"] for r in ROWS])
def test_reference_import_statement_resolves(row):
    mod = importlib.import_module(row["module"])
    assert mod.__file__.startswith(os.path.join(ROOT, "3dod_b200")), "resolved outside the mirror: %s" % mod.__file__
    for name in row["names"]:
        assert hasattr(mod, name) or _is_submodule(row["module"], name), \
            "%s: `from %s import %s` fails against the mirror" % (row["where"], row["module"], name)


def _is_submodule(module, name):
    try:
        importlib.import_module(module + "." + name)
        return True
    except ImportError:
        return False


def test_attribute_uses_of_the_proposals_module():
    """roi_heads.py:283-302 (`predict_cubes`) reaches the samplers as attributes of the module imported at :30"""
    import ProposalNetwork.proposals.proposals as proposals
    want = ["propose", "propose_random", "propose_xy_patch", "propose_z", "propose_random_dim",
            "propose_random_rotation", "propose_aspect_ratio"]
    ref_sig = ["reference_box", "depth_image", "priors", "im_shape", "K", "number_of_proposals", "gt_cubes",
               "ground_normal"]
    for name in want:
        fn = getattr(proposals, name)
        assert list(inspect.signature(fn).parameters) == ref_sig, name


def test_signatures_of_the_late_additions():
    """names the first round's mirror lacked: iou_3d / mask_iou_loss (utils.py:194-210, :252-260), cube_to_box
    (conversions.py:5-23), show_mask2 (utils.py:335), Plane.fit (plane.py:22)"""
    from ProposalNetwork.utils import conversions, plane, spaces, utils
    assert list(inspect.signature(utils.iou_3d).parameters) == ["gt_cube", "proposal_cubes"]
    assert list(inspect.signature(utils.mask_iou_loss).parameters) == ["segmentation_mask", "bube_mask"]
    assert list(inspect.signature(utils.show_mask2).parameters) == ["masks", "im", "random_color"]
    assert list(inspect.signature(conversions.cube_to_box).parameters) == ["cube", "K"]
    assert list(inspect.signature(plane.Plane.fit).parameters) == ["self", "pts", "thresh", "minPoints", "maxIteration"]
    assert list(inspect.signature(spaces.Cube.__init__).parameters) == ["self", "tensor", "R", "score", "label"]


def test_show_mask2_blend_matches_the_formula():
    """utils.py:335-361: per CHANNEL, 0.5 * colour + 0.5 * image where the colour layer is positive, image elsewhere
    (a zero colour component leaves that channel untouched); alpha channel 1 outside"""
    import numpy as np
    from ProposalNetwork.utils.utils import show_mask2
    im = np.full((4, 6, 3), 100.0)
    m = np.zeros((1, 4, 6), dtype=bool)
    m[0, 1:3, 2:4] = True
    out = show_mask2(m, im, random_color=[np.array([1.0, 0.0, 0.0, 0.6])])
    assert out.shape == (4, 6, 4)
    assert np.allclose(out[0, 0], [100 / 255, 100 / 255, 100 / 255, 1.0])
    assert np.allclose(out[1, 2], [0.5 + 0.5 * 100 / 255, 100 / 255, 100 / 255, 0.5 * 0.6 + 0.5])


def test_plane_fit_draws_like_the_reference():
    """plane.py:41: one torch.randperm(n) per iteration from torch's default CPU generator, first three kept"""
    import torch
    from ProposalNetwork.utils import plane
    torch.manual_seed(3)
    ids = plane._draw_ids_randperm(500, 7)
    torch.manual_seed(3)
    ref = torch.stack([torch.randperm(500)[:3] for _ in range(7)])
    assert torch.equal(ids.long(), ref)
