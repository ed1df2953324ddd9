"""ProposalNetwork/utils/plane.py on the GPU: RANSAC plane fit without the 1000 x P distance matrix.

`Plane` keeps the torch API (fit / fit_parallel return the NEGATED equation, plane.py:77,134); `Plane_np` keeps the
NumPy / pyransac3d API used at cubercnn/modeling/roi_heads/roi_heads.py:374-376 (returns +equation, fp64).
Sample ids are drawn with python's `random.sample` exactly as plane.py:95 / :171 do, so a seeded run selects the
same hypotheses as the reference."""
import random

import numpy as np
import torch

from tdod import api


def _draw_ids(n_points, iters):
    return [random.sample(range(0, n_points), 3) for _ in range(iters)]


def _fit(pts_soa, n_points, thresh, iters):
    ids = torch.tensor(_draw_ids(n_points, iters), dtype=torch.int32).to(pts_soa.device, non_blocking=True)[None]
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
        """plane.py:22-77: the sequential form keeps the first hypothesis with a strictly larger inlier count,
        which is the first maximum -- the same winner fit_parallel's argmax picks."""
        return self.fit_parallel(pts, thresh, minPoints, maxIteration)


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
