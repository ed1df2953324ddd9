"""Seeded inputs of the depth-map losses (pseudo_gt_z_box_loss, pseudo_gt_z_point_loss, depth_range_loss), shared by the
fixture generator (make_golden.py depth_losses) and the tests.  Needs the oracle and tdod.synth only, not the reference."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (os.path.join(ROOT, "oracle"), os.path.join(ROOT, "3dod_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import tdod_oracle as O  # noqa: E402
from tdod import synth  # noqa: E402


def depth_loss_case():
    """inputs of the depth-map losses: three 96x128 frames, their boxes jittered, some pushed out of the image, one
    object without mask pixels; shared by the fixture below and the tests (everything is seeded)"""
    H, W = 96, 128
    frames = [synth.make_frame(31 + i, H, W, 5 + i, with_masks=True) for i in range(3)]
    g = torch.Generator().manual_seed(9)
    depth = torch.stack([f.depth.float() for f in frames])
    boxes, masks, counts, gtb = [], [], [], []
    for i, f in enumerate(frames):
        b = f.boxes.float() + (torch.rand(f.boxes.shape, generator=g) - 0.5) * 6
        b[0] = torch.tensor([W + 5.0, 10.0, W + 40.0, 50.0])          # right of the image: no area after clipping
        if i == 1:
            b[1] = torch.tensor([-30.0, -20.0, -2.0, 40.0])            # left of it
            b[2, 0], b[2, 2] = -8.5, 37.25                             # straddles the left border
            b[3] = torch.tensor([20.3, 30.2, 90.7, 30.9])              # has area, empty slice: NaN here, raises there
        boxes.append(b)
        gtb.append(f.boxes.float())
        masks += [m.bool() for m in f.masks]
        counts.append(f.boxes.shape[0])
    masks = torch.stack(masks)
    masks[4] = False                                                   # segmentation failed: falls back to the box
    N = sum(counts)
    pred_z = torch.rand(N, generator=g) * 5 + 1
    pred_xy = torch.rand((N, 2), generator=g) * torch.tensor([W + 20.0, H + 20.0]) - 10
    cubes = torch.cat([O.propose(f.boxes, f.depth, f.prior_mean, f.prior_std, f.K, 4, normal=torch.tensor([0.05, 0.99, 0.1]),
                                 generator=g)["cubes"][:, 2].float() for f in frames])
    idx = torch.randperm(N, generator=g)
    return dict(depth=depth, boxes=boxes, gt_boxes=torch.cat(gtb), masks=masks, counts=counts, pred_z=pred_z,
                pred_xy=pred_xy, cubes=cubes, mask_idx=idx)
