"""`Cubes` container with the interface of the reference's ProposalNetwork/utils/spaces.py:95-328, its two
compute methods (`get_all_corners` :192-204, `get_bube_corners` :224-245) forwarding to libtdod.

Layout: fp32 [num_instances, num_proposals, 15] = (x, y, z, w, h, l, R row-major)."""
import numpy as np
import torch

from tdod import api


class Boxes:
    """Stand-in with the public behaviour of detectron2.structures.Boxes for what this path touches (tensor,
    len, indexing, area, to/device, cat): detectron2 is not a dependency of this package.  When detectron2 is
    importable the mirror returns its Boxes instead (see `boxes_cls`)."""

    def __init__(self, tensor):
        if not isinstance(tensor, torch.Tensor):
            tensor = torch.as_tensor(tensor, dtype=torch.float32)
        else:
            tensor = tensor.to(torch.float32)
        if tensor.numel() == 0:
            tensor = tensor.reshape((-1, 4))
        if tensor.dim() != 2 or tensor.size(-1) != 4:
            raise AssertionError(tensor.size())
        self.tensor = tensor

    def __len__(self):
        return self.tensor.shape[0]

    def __getitem__(self, item):
        if isinstance(item, int):
            return Boxes(self.tensor[item].view(1, -1))
        return Boxes(self.tensor[item])

    def __iter__(self):
        yield from self.tensor

    def area(self):
        t = self.tensor
        return (t[:, 2] - t[:, 0]) * (t[:, 3] - t[:, 1])

    def to(self, device):
        return Boxes(self.tensor.to(device=device))

    @property
    def device(self):
        return self.tensor.device

    @classmethod
    def cat(cls, boxes_list):
        if len(boxes_list) == 0:
            return cls(torch.empty(0))
        return cls(torch.cat([b.tensor for b in boxes_list], dim=0))

    def __repr__(self):
        return "Boxes(" + str(self.tensor) + ")"


def boxes_cls():
    try:
        from detectron2.structures import Boxes as D2Boxes
        return D2Boxes
    except Exception:
        return Boxes


def _require_cuda(t, what):
    if not t.is_cuda:
        raise ValueError("%s must live on a CUDA device: the B200 path has no CPU implementation" % what)
    if t.requires_grad:
        raise ValueError("%s requires grad: the CUDA path is forward-only (weak-loss training is out of scope)" % what)


class Cube:
    """spaces.py:16-93: the single-cube container some callers still import (`cube_to_box`, conversions.py:5-23).
    tensor [6] = (x, y, z, w, h, l), R [3,3]; corners and the unclamped projection come from libtdod."""

    def __init__(self, tensor, R, score=None, label=None):
        self.tensor = tensor
        self.center = tensor[:3]
        self.dimensions = tensor[3:6]
        self.rotation = R
        self.score = score
        self.label = label

    def _row(self):
        _require_cuda(self.tensor, "Cube.tensor")
        return torch.cat((self.tensor[:6].to(torch.float32), self.rotation.to(torch.float32).reshape(9)))[None].contiguous()

    def get_all_corners(self):
        """[8,3] (math_util.py:142-245 vertex order)"""
        return api.cube_corners(self._row())[0]

    def get_bube_corners(self, K):
        """[8,2]: K @ corner, divided by depth, no clamp (spaces.py:69-72)"""
        p = self.get_all_corners() @ K.to(self.tensor.device, torch.float32).t()
        return p[:, :2] / p[:, 2].unsqueeze(1)

    def get_volume(self):
        return self.dimensions.prod().item()

    def __repr__(self):
        return f"Cube({self.center}, {self.dimensions}, {self.rotation})"

    def to_device(self, device):
        self.tensor = self.tensor.to(device)
        self.center, self.dimensions = self.center.to(device), self.dimensions.to(device)
        self.rotation = self.rotation.to(device)
        return self


class Cubes:
    """spaces.py:95-328.  `scores` / `labels` are auxiliary side fields exactly as in the reference."""

    def __init__(self, tensor, scores=None, labels=None):
        if scores is not None:
            assert scores.ndim == 2, f"scores.shape must be (n_instances, n_proposals), but was {scores.shape}"
        self.scores = scores
        self.labels = labels
        if not isinstance(tensor, torch.Tensor):
            tensor = torch.as_tensor(np.asarray(tensor), dtype=torch.float32, device=torch.device("cpu"))
        else:
            tensor = tensor.to(torch.float32)
        if tensor.numel() == 0:
            tensor = tensor.reshape((-1, 15))
        while tensor.dim() < 3:
            tensor = tensor.unsqueeze(0)
        self.tensor = tensor

    # -- views ------------------------------------------------------------------------------------------------
    @property
    def centers(self):
        return self.tensor[:, :, :3]

    @property
    def dimensions(self):
        return self.tensor[:, :, 3:6]

    @property
    def rotations(self):
        s = self.tensor.shape
        return self.tensor[:, :, 6:].reshape(s[0], s[1], 3, 3)

    @property
    def device(self):
        return self.tensor.device

    @property
    def num_instances(self):
        return self.tensor.shape[0]

    @property
    def shape(self):
        return self.tensor.shape

    def __len__(self):
        return self.tensor.shape[0]

    def __repr__(self):
        return f"Cubes({self.tensor})"

    def __iter__(self):
        yield from self.tensor

    def clone(self):
        return Cubes(self.tensor.clone())

    def to(self, device):
        if isinstance(self.scores, torch.Tensor):
            self.scores = self.scores.to(device=device)
        if isinstance(self.labels, torch.Tensor):
            self.labels = self.labels.to(device=device)
        return Cubes(self.tensor.to(device=device), self.scores, self.labels)

    def __getitem__(self, item):
        if isinstance(item, int):
            return Cubes(self.tensor[item].view(1, self.tensor.shape[1], -1))
        if isinstance(item, tuple):
            return Cubes(self.tensor[item[0], item[1]].view(1, 1, -1))
        b = self.tensor[item]
        assert b.dim() == 2, "Indexing on Cubes with {} failed to return a matrix!".format(item)
        return Cubes(b)

    @classmethod
    def cat(cls, cubes_list):
        assert isinstance(cubes_list, (list, tuple))
        if len(cubes_list) == 0:
            return cls(torch.empty(0))
        assert all(isinstance(c, Cubes) for c in cubes_list)
        return cls(torch.cat([c.tensor for c in cubes_list], dim=0))

    def split(self, split_size, dim=1):
        return tuple(Cubes(x) for x in self.tensor.split(split_size, dim=dim))

    def reshape(self, *args):
        return Cubes(self.tensor.reshape(*args), self.scores, self.labels)

    def get_volumes(self):
        return self.dimensions.prod(-1)

    # -- compute (libtdod) ------------------------------------------------------------------------------------
    def get_all_corners(self):
        """[num_instances, n, 8, 3] camera-space corners (math_util.py:142-245 vertex order)."""
        _require_cuda(self.tensor, "Cubes.tensor")
        return api.cube_corners(self.tensor.contiguous())

    def get_bube_corners(self, K, clamp=None):
        """[num_instances, n, 8, 2] projected corners.  `clamp` = (width, height) is mandatory, as it is in effect
        in the reference (spaces.py:240-243 raises NameError without it)."""
        if clamp is None:
            raise NameError("get_bube_corners needs clamp=(width, height) (spaces.py:240-243)")
        _require_cuda(self.tensor, "Cubes.tensor")
        c2d, _ = api.project_cubes(self.tensor.contiguous(), K.to(self.tensor.device, torch.float32), clamp[0], clamp[1],
                                   corners=True, boxes=False)
        return c2d
