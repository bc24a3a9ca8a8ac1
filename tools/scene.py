"""Synthetic street-scene generator shared by tests and bench.

Deterministic (torch.Generator seed), CPU tensors; callers move them to the GPU."""
from __future__ import annotations

import math
import torch


def make_camera(width: int, height: int, yaw_deg: float = 0.0, z_shift: float = 0.0,
                x_shift: float = 0.0):
    """OpenCV pinhole, Waymo-front-like focal; returns (viewmat[4,4] world->cam, K[3,3])."""
    fx = fy = 2055.6 * width / 1920.0
    K = torch.tensor([[fx, 0.0, width / 2.0], [0.0, fy, height / 2.0], [0.0, 0.0, 1.0]])
    a = math.radians(yaw_deg)
    # camera-to-world rotation: yaw about the y (down) axis
    R_c2w = torch.tensor([[math.cos(a), 0.0, math.sin(a)], [0.0, 1.0, 0.0],
                          [-math.sin(a), 0.0, math.cos(a)]])
    t_c2w = torch.tensor([x_shift, 0.0, z_shift])
    R = R_c2w.T
    t = -(R @ t_c2w)
    V = torch.eye(4)
    V[:3, :3] = R
    V[:3, 3] = t
    return V, K


def make_gaussians(n: int, sh_degree: int = 3, seed: int = 0, scale_mean: float = 0.08,
                   box=((-40.0, 40.0), (-3.0, 6.0), (-10.0, 90.0)), device=None):
    """device=None: the CPU generator (the stream every parity test is pinned to); a CUDA device draws the same
    distributions there (fast for the 6 M-Gaussian bench scene; other values)"""
    if device is not None:
        g = torch.Generator(device=device).manual_seed(seed)
        kw = dict(generator=g, device=device)
        u = torch.rand(n, 3, **kw)
        lo = torch.tensor([b[0] for b in box], device=device)
        hi = torch.tensor([b[1] for b in box], device=device)
        means = lo + u * (hi - lo)
        scales = torch.exp(math.log(scale_mean) + 0.6 * torch.randn(n, 3, **kw))
        quats = torch.randn(n, 4, **kw)
        opacities = torch.sigmoid(2.0 * torch.randn(n, **kw))
        K = (sh_degree + 1) ** 2
        shs = torch.cat([(torch.rand(n, 1, 3, **kw) * 2 - 1) * 1.7, 0.1 * torch.randn(n, K - 1, 3, **kw)], 1)
        return dict(means=means, quats=quats, scales=scales, opacities=opacities, shs=shs)
    g = torch.Generator().manual_seed(seed)
    u = torch.rand(n, 3, generator=g)
    lo = torch.tensor([b[0] for b in box])
    hi = torch.tensor([b[1] for b in box])
    means = lo + u * (hi - lo)
    scales = torch.exp(math.log(scale_mean) + 0.6 * torch.randn(n, 3, generator=g))
    quats = torch.randn(n, 4, generator=g)
    opacities = torch.sigmoid(2.0 * torch.randn(n, generator=g))
    K = (sh_degree + 1) ** 2
    shs = torch.cat([(torch.rand(n, 1, 3, generator=g) * 2 - 1) * 1.7,
                     0.1 * torch.randn(n, K - 1, 3, generator=g)], 1) if K > 1 else \
        (torch.rand(n, 1, 3, generator=g) * 2 - 1) * 1.7
    return dict(means=means, quats=quats, scales=scales, opacities=opacities, shs=shs)


def make_lidar_points(n: int, seed: int = 0):
    g = torch.Generator().manual_seed(seed)
    u = torch.rand(n, 3, generator=g)
    lo = torch.tensor([-40.0, -3.0, 1.0])
    hi = torch.tensor([40.0, 6.0, 80.0])
    xyz = lo + u * (hi - lo)
    rgb = torch.rand(n, 3, generator=g)
    return xyz, rgb
