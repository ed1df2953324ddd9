"""ProposalNetwork/utils/plane.py on the GPU: RANSAC plane fit without the 1000 x P distance matrix.

`Plane` keeps the torch API (fit / fit_parallel return the NEGATED equation, plane.py:77,134); `Plane_np` keeps the
NumPy / pyransac3d API used at cubercnn/modeling/roi_heads/roi_heads.py:374-376 (returns +equation, fp64).
Sample ids are drawn exactly as the reference draws them -- python's `random.sample` in fit_parallel / Plane_np.fit
(plane.py:95, :171), `torch.randperm(n)[:3]` from torch's default CPU generator in the sequential Plane.fit (:41) --
so a seeded run selects the same hypotheses as the reference."""
import random

import numpy as np
import torch

from tdod import api


def _draw_ids(n_points, iters):
    return [random.sample(range(0, n_points), 3) for _ in range(iters)]


def _draw_ids_randperm(n_points, iters):
    """plane.py:41: one full torch.randperm(n_points) per iteration, first three entries kept"""
    return torch.stack([torch.randperm(n_points)[:3] for _ in range(iters)]).to(torch.int32)


def _fit(pts_soa, n_points, thresh, iters, ids=None):
    if ids is None:
        ids = torch.tensor(_draw_ids(n_points, iters), dtype=torch.int32)
    ids = ids.to(pts_soa.device, non_blocking=True)[None]
    n = torch.tensor([n_points], dtype=torch.int32, device=pts_soa.device)
    return api.ransac_plane(pts_soa, n, ids.contiguous(), iters=iters, thresh=thresh, want_mask=True)


class Plane:
    def __init__(self):
        self.inliers = []
        self.equation = []

    def fit_parallel(self, pts, thresh=0.05, minPoints=100, maxIteration=1000):
        """plane.py:79-134.  pts [P,3] CUDA fp32 -> (-equation [4], inlier indices int64)."""
        if not pts.is_cuda:
            raise ValueError("pts must be a CUDA tensor (no CPU path)")
        soa = pts.to(torch.float32).t().contiguous()[None]
        r = _fit(soa, pts.shape[0], thresh, maxIteration)
        self.equation = r["eq"][0]
        self.inliers = torch.where(r["mask"][0] > 0)[0]
        return -self.equation, self.inliers

    def fit(self, pts, thresh=0.05, minPoints=100, maxIteration=1000):
        """plane.py:22-77: the sequential form draws `torch.randperm(n)[:3]` per iteration (:41) and keeps the first
        hypothesis with a strictly larger inlier count, i.e. the first maximum -- what the kernel's argmax picks.
        The plane arithmetic (:53-66) is the same chain of fp32 torch ops as fit_parallel's."""
        if not pts.is_cuda:
            raise ValueError("pts must be a CUDA tensor (no CPU path)")
        soa = pts.to(torch.float32).t().contiguous()[None]
        r = _fit(soa, pts.shape[0], thresh, maxIteration, ids=_draw_ids_randperm(pts.shape[0], maxIteration))
        self.equation = r["eq"][0]
        self.inliers = torch.where(r["mask"][0] > 0)[0]
        return -self.equation, self.inliers


class Plane_np:
    """pyransac3d.Plane-compatible: NumPy points in, python list + index array out (roi_heads.py:374-376)."""

    def __init__(self):
        self.inliers = []
        self.equation = []

    def fit(self, pts, thresh=0.05, minPoints=100, maxIteration=1000):
        p = np.ascontiguousarray(np.asarray(pts, dtype=np.float64).T)
        soa = torch.from_numpy(p).cuda()[None]
        r = _fit(soa, p.shape[1], thresh, maxIteration)
        eq = r["eq"][0].cpu().numpy()
        self.equation = [eq[0], eq[1], eq[2], eq[3]]
        self.inliers = torch.where(r["mask"][0] > 0)[0].cpu().numpy()
        return self.equation, self.inliers
