"""CPU oracle for `distCUDA2` (TEST INFRASTRUCTURE, NOT PRODUCT).

Restates /root/reference/submodules/simple-knn/simple_knn.cu:132-184 semantics: for each
point the mean of the 3 smallest squared Euclidean distances to the OTHER points (self
excluded by index, so exact duplicates count with distance 0).  The CUDA algorithm there is
an exact KNN (its box pruning is conservative, simple_knn.cu:164-174), so brute force is an
exact restatement up to fp32 rounding of dx²+dy²+dz².  Pinned on the GPU box against the
reference's own kernels compiled into oracle/_ref/ (oracle/build_ref.sh).
"""
from __future__ import annotations

import numpy as np
import torch


def dist_knn3_bruteforce(points: torch.Tensor, chunk: int = 2048) -> torch.Tensor:
    """points [P,3] fp32 -> [P] fp32.  O(P²); use for P <= ~50k."""
    P = points.shape[0]
    pts = points.float()
    out = torch.empty(P, dtype=torch.float32)
    FLT_MAX = torch.finfo(torch.float32).max
    for s in range(0, P, chunk):
        q = pts[s:s + chunk]
        dx = q[:, None, 0] - pts[None, :, 0]
        dy = q[:, None, 1] - pts[None, :, 1]
        dz = q[:, None, 2] - pts[None, :, 2]
        d = dx * dx + dy * dy + dz * dz
        idx = torch.arange(s, min(P, s + chunk))
        d[torch.arange(idx.numel()), idx] = FLT_MAX
        k = min(3, P)
        best = torch.topk(d, k, dim=1, largest=False).values
        if k < 3:
            best = torch.cat([best, torch.full((best.shape[0], 3 - k), FLT_MAX)], 1)
        # with P < 4 the FLT_MAX placeholders overflow to +inf exactly as simple_knn.cu:183 does
        out[s:s + chunk] = (best[:, 0] + best[:, 1] + best[:, 2]) / 3.0
    return out


def dist_knn3_kdtree(points: torch.Tensor, workers: int = -1) -> torch.Tensor:
    """cKDTree KNN-3 (float64 distances), for large P.  Duplicates handled by index."""
    from scipy.spatial import cKDTree
    p = points.double().numpy()
    tree = cKDTree(p)
    # k=4: self + 3 neighbours; with exact duplicates self may not be column 0, so drop by index
    d, i = tree.query(p, k=min(8, p.shape[0]), workers=workers)
    self_idx = np.arange(p.shape[0])[:, None]
    d = np.where(i == self_idx, np.inf, d)
    d.sort(axis=1)
    d3 = d[:, :3] ** 2
    return torch.from_numpy(d3.mean(1).astype(np.float32))
