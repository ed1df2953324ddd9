"""ProposalNetwork/utils/conversions.py: cubes -> tight 2D boxes on the GPU (+ the two coordinate helpers)."""
import numpy as np
import torch

from tdod import api
from .spaces import _require_cuda, boxes_cls


def cube_to_box(cube, K):
    """conversions.py:5-23: tight box around the 8 projected corners of ONE `Cube` -> Boxes[1] (no clamp)."""
    c = cube.get_bube_corners(K)
    lo, hi = c.min(dim=0)[0], c.max(dim=0)[0]
    return boxes_cls()(torch.stack((lo[0], lo[1], hi[0], hi[1]))[None])


def cubes_to_box(cubes, K, im_shape):
    """conversions.py:25-48.  cubes: Cubes [N,n,15]; K [3,3]; im_shape = (width, height) -> list of N Boxes[n]."""
    _require_cuda(cubes.tensor, "cubes")
    _, bx = api.project_cubes(cubes.tensor.contiguous(), K.to(cubes.tensor.device, torch.float32), im_shape[0],
                              im_shape[1], corners=False, boxes=True)
    B = boxes_cls()
    return [B(bx[i]) for i in range(cubes.num_instances)]


def pixel_to_normalised_space(pixel_coord, im_shape, norm_shape):
    """conversions.py:50-67 (host-side helper of the ablation samplers): list of N tensors -> [len, N] fp32,
    (p - old/2) * new/old per axis."""
    p = torch.stack(pixel_coord, dim=1).to(torch.float32)
    for i in range(p.size(1)):
        p[:, i] -= 0.5 * im_shape[i]
        p[:, i] *= norm_shape[i] / im_shape[i]
    return p


def normalised_space_to_pixel(coords, im_shape, norm_shape):
    """conversions.py:69-78"""
    c = np.array(coords).astype(np.float32)
    for i in range(len(c)):
        c[i] *= im_shape[i] / norm_shape[i]
        c[i] += 0.5 * im_shape[i]
    return c
