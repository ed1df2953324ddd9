"""TEST INFRASTRUCTURE ONLY -- never imported by the product path.

Imports the UNMODIFIED reference hot path from /root/reference (this container only; the
GPU box has no /root/reference) behind four in-memory stubs, as described in SURVEY.md 8(c):

  1. detectron2.structures.{Boxes, pairwise_iou}  -- public formula restated (detectron2 is an
     unpinned dependency of the reference, requirements.txt:9; call site utils.py:191)
  2. pytorch3d.ops.box3d_overlap = None           -- only used by iou_3d (evaluation, off path)
  3. matplotlib.pyplot                            -- empty module (utils.py:3)
  4. cubercnn.util                                -- get_cuboid_verts_faces + mat2euler exec'd
     verbatim *from the file where it lies* (cubercnn/util/math_util.py:133-245, :71-81);
     nothing is copied into this repository.

Used by tests/golden/make_golden.py (fixture generation) and oracle validation scripts.
"""
import importlib
import math
import os
import sys
import types

import numpy as np
import torch

REF_ROOT = os.environ.get("TDOD_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REF_ROOT, "ProposalNetwork"))


class _Boxes:
    """Minimal stand-in for detectron2.structures.Boxes (public behaviour only)."""

    def __init__(self, tensor):
        if not isinstance(tensor, torch.Tensor):
            tensor = torch.as_tensor(tensor, dtype=torch.float32)
        else:
            tensor = tensor.to(torch.float32)
        if tensor.numel() == 0:
            tensor = tensor.reshape((-1, 4)).to(dtype=torch.float32)
        assert tensor.dim() == 2 and tensor.size(-1) == 4, tensor.size()
        self.tensor = tensor

    def area(self):
        b = self.tensor
        return (b[:, 2] - b[:, 0]) * (b[:, 3] - b[:, 1])

    def __len__(self):
        return self.tensor.shape[0]

    def __getitem__(self, item):
        if isinstance(item, int):
            return _Boxes(self.tensor[item].view(1, -1))
        return _Boxes(self.tensor[item])

    def __iter__(self):
        yield from self.tensor

    def to(self, device):
        return _Boxes(self.tensor.to(device=device))

    @property
    def device(self):
        return self.tensor.device

    def get_centers(self):
        return (self.tensor[:, :2] + self.tensor[:, 2:]) / 2

    def clip(self, box_size):
        assert torch.isfinite(self.tensor).all(), "Box tensor contains infinite or NaN!"
        h, w = box_size
        x1, y1 = self.tensor[:, 0].clamp(min=0, max=w), self.tensor[:, 1].clamp(min=0, max=h)
        x2, y2 = self.tensor[:, 2].clamp(min=0, max=w), self.tensor[:, 3].clamp(min=0, max=h)
        self.tensor = torch.stack((x1, y1, x2, y2), dim=-1)

    @classmethod
    def cat(cls, boxes_list):
        if len(boxes_list) == 0:
            return cls(torch.empty(0))
        return cls(torch.cat([b.tensor for b in boxes_list], dim=0))


def _pairwise_iou(boxes1, boxes2):
    area1 = boxes1.area()
    area2 = boxes2.area()
    b1, b2 = boxes1.tensor, boxes2.tensor
    wh = torch.min(b1[:, None, 2:], b2[:, 2:]) - torch.max(b1[:, None, :2], b2[:, :2])
    wh.clamp_(min=0)
    inter = wh.prod(dim=2)
    iou = torch.where(
        inter > 0,
        inter / (area1[:, None] + area2 - inter),
        torch.zeros(1, dtype=inter.dtype, device=inter.device),
    )
    return iou


def _install_stubs():
    if "detectron2" not in sys.modules:
        d2 = types.ModuleType("detectron2")
        st = types.ModuleType("detectron2.structures")
        st.Boxes = _Boxes
        st.pairwise_iou = _pairwise_iou
        d2.structures = st
        sys.modules["detectron2"] = d2
        sys.modules["detectron2.structures"] = st
    if "pytorch3d" not in sys.modules:
        p3 = types.ModuleType("pytorch3d")
        ops = types.ModuleType("pytorch3d.ops")
        ops.box3d_overlap = None
        p3.ops = ops
        sys.modules["pytorch3d"] = p3
        sys.modules["pytorch3d.ops"] = ops
    if "matplotlib" not in sys.modules:
        mpl = types.ModuleType("matplotlib")
        plt = types.ModuleType("matplotlib.pyplot")
        mpl.pyplot = plt
        sys.modules["matplotlib"] = mpl
        sys.modules["matplotlib.pyplot"] = plt
    if "cubercnn" not in sys.modules:
        src = open(os.path.join(REF_ROOT, "cubercnn/util/math_util.py")).read().split("\n")
        ns = {"torch": torch, "np": np, "math": math}
        # math_util.py:133-245 (to_float_tensor + get_cuboid_verts_faces), :71-81 (mat2euler)
        exec("\n".join(src[132:245]), ns)
        exec("\n".join(src[70:81]), ns)
        cc = types.ModuleType("cubercnn")
        util = types.ModuleType("cubercnn.util")
        util.get_cuboid_verts_faces = ns["get_cuboid_verts_faces"]
        util.mat2euler = ns["mat2euler"]
        cc.util = util
        sys.modules["cubercnn"] = cc
        sys.modules["cubercnn.util"] = util


class Ref:
    """Namespace of the reference's hot-path modules."""


def load():
    """Returns a namespace with the reference modules; raises if /root/reference is absent."""
    if not available():
        raise RuntimeError("reference tree not present (expected on the GPU box)")
    _install_stubs()
    # the product mirror is also called `ProposalNetwork`; make sure the reference wins here
    for k in [k for k in sys.modules if k == "ProposalNetwork" or k.startswith("ProposalNetwork.")]:
        del sys.modules[k]
    # the reference's ProposalNetwork has no __init__.py (a namespace package): a regular package of the same name
    # anywhere on sys.path (the mirror under 3dod_b200/) would win whatever the order, so hide those entries meanwhile
    hidden = [p for p in sys.path if p != REF_ROOT and os.path.isfile(os.path.join(p or ".", "ProposalNetwork", "__init__.py"))]
    saved_path = list(sys.path)
    sys.path[:] = [REF_ROOT] + [p for p in sys.path if p not in hidden]
    try:
        r = Ref()
        r.proposals = importlib.import_module("ProposalNetwork.proposals.proposals")
        r.scorefunction = importlib.import_module("ProposalNetwork.scoring.scorefunction")
        r.spaces = importlib.import_module("ProposalNetwork.utils.spaces")
        r.utils = importlib.import_module("ProposalNetwork.utils.utils")
        r.conversions = importlib.import_module("ProposalNetwork.utils.conversions")
        r.plane = importlib.import_module("ProposalNetwork.utils.plane")
        r.Boxes = _Boxes
        r.pairwise_iou = _pairwise_iou
    finally:
        sys.path[:] = saved_path
        # leave the names importable only through the returned namespace
        for k in [k for k in sys.modules if k == "ProposalNetwork" or k.startswith("ProposalNetwork.")]:
            del sys.modules[k]
    return r
