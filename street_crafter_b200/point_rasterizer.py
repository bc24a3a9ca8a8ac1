"""Drop-in for `diff_point_rasterization.{PointRasterizationSettings, PointRasterizer}` as
used by the LiDAR colorized-point-cloud condition render
(/root/reference/data_processor/utils/render_utils.py:129-177).

The third-party rasterizer is absent from the reference tree and unpinned, so
its semantics are RESTATED here from its call site and its in-tree sibling
`render_pointcloud_pytorch3d` (render_utils.py:21-81): every point is an isotropic splat of
world-space standard deviation `radius`, opacity as given (1.0 at the call site), colours
precomputed, composited front-to-back like the Gaussian path, and at most `max_hit`
contributing splats per pixel (points_per_pixel=10 in the sibling).

The camera is decoded from what the caller ACTUALLY passes (render_utils.py:150-163): `viewmatrix` =
world->camera TRANSPOSED and `projmatrix` = viewmatrix @ P^T, where P is the K-based OpenGL-style
projection of `convert_to_gaussian_camera` — the convention the reference tree itself carries as
`getProjectionMatrixK` (street_gaussian/utils/graphics_utils.py:79-101):
    P[0,0] = 2 fx / W   P[0,2] = 2 cx / W - 1   P[1,1] = 2 fy / H   P[1,2] = 2 cy / H - 1
    P[2,2] = (f + n) / (f - n)                  P[2,3] = -2 f n / (f - n)          P[3,2] = 1
so fx, fy, the principal point (cx, cy) and near/far (1.0 / 100.0 at render_utils.py:138-139) are
recovered from P = (viewmatrix^-1 @ projmatrix)^T.  Inria's ndc2Pix with pixel centres at integers
and this library's `fx x/z + cx` with centres at +0.5 describe the same image.  `znear` / `zfar` /
`principal` remain as optional OVERRIDES only (None = decode)."""
from __future__ import annotations

import math
from typing import NamedTuple, Optional

import torch
from torch import Tensor

from . import rendering as R


class PointRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: Tensor
    scale_modifier: float
    viewmatrix: Tensor      # [4,4] world->camera, stored TRANSPOSED (Inria convention)
    projmatrix: Tensor
    sh_degree: int
    max_hit: int
    campos: Tensor
    prefiltered: bool
    debug: bool
    znear: Optional[float] = None      # overrides; None = decoded from projmatrix
    zfar: Optional[float] = None
    principal: Optional[tuple] = None  # (cx, cy) override


def decode_camera(s: "PointRasterizationSettings"):
    """-> (fx, fy, cx, cy, near, far) from viewmatrix / projmatrix / tanfov* as the reference caller
    packs them (see the module docstring).  A projmatrix that is not of that form (e.g. identity)
    falls back to tanfov* with a centred principal point and near/far 1.0/100.0."""
    H, W = int(s.image_height), int(s.image_width)
    fx = W / (2.0 * float(s.tanfovx))
    fy = H / (2.0 * float(s.tanfovy))
    cx, cy, near, far = W / 2.0, H / 2.0, 1.0, 100.0
    pm = s.projmatrix
    if pm is not None:
        vt = torch.as_tensor(s.viewmatrix).detach().double().cpu()
        P = (torch.linalg.inv(vt) @ torch.as_tensor(pm).detach().double().cpu()).T
        A, B = float(P[2, 2]), float(P[2, 3])
        if abs(float(P[3, 2]) - 1.0) < 1e-6 and abs(float(P[3, 3])) < 1e-6 and float(P[0, 0]) > 0 and B < 0:
            fx = float(P[0, 0]) * W / 2.0
            fy = float(P[1, 1]) * H / 2.0
            cx = (float(P[0, 2]) + 1.0) * W / 2.0
            cy = (float(P[1, 2]) + 1.0) * H / 2.0
            if abs(A + 1.0) > 1e-12 and abs(A - 1.0) > 1e-12:
                # the planes were float32 constants: drop the round-off of the matrix products
                near = float("%.6g" % (-B / (A + 1.0)))
                far = float("%.6g" % (-B / (A - 1.0)))
    if s.principal is not None:
        cx, cy = (float(v) for v in s.principal)
    if s.znear is not None:
        near = float(s.znear)
    if s.zfar is not None:
        far = float(s.zfar)
    return fx, fy, cx, cy, near, far


class PointRasterizer(torch.nn.Module):
    def __init__(self, raster_settings: PointRasterizationSettings):
        super().__init__()
        self.raster_settings = raster_settings

    def forward(self, means3D: Tensor, means2D: Tensor, colors_precomp: Tensor, opacities: Tensor,
                radius: Tensor):
        """-> (image[3,H,W], depth[1,H,W], alpha[1,H,W], radii[N] int32)."""
        s = self.raster_settings
        H, W = int(s.image_height), int(s.image_width)
        N = means3D.shape[0]
        dev = means3D.device
        fx, fy, cx, cy, znear, zfar = decode_camera(s)
        K = torch.tensor([[fx, 0.0, cx], [0.0, fy, cy], [0.0, 0.0, 1.0]], dtype=torch.float32, device=dev)
        viewmat = torch.as_tensor(s.viewmatrix, dtype=torch.float32, device=dev).T.contiguous()
        scales = (radius.reshape(N, 1).float() * float(s.scale_modifier)).expand(N, 3).contiguous()
        quats = torch.zeros(N, 4, dtype=torch.float32, device=dev)
        quats[:, 0] = 1.0
        radii, means2d, depths, conics, _ = R.fully_fused_projection(
            means3D.float(), None, quats, scales, viewmat[None], K[None], W, H, near_plane=znear,
            far_plane=zfar, calc_compensations=False)
        tw, th = math.ceil(W / 16.0), math.ceil(H / 16.0)
        _, isect_ids, flatten_ids = R.isect_tiles(means2d, radii, depths, 16, tw, th, n_cameras=1)
        offsets = R.isect_offset_encode(isect_ids, 1, tw, th)
        cols = torch.cat([colors_precomp.float().reshape(N, -1)[None], depths[..., None]], dim=-1)
        bg = torch.cat([torch.as_tensor(s.bg, dtype=torch.float32, device=dev).reshape(1, -1),
                        torch.zeros(1, 1, device=dev)], dim=-1)
        rc, ra = R.rasterize_to_pixels(means2d, conics, cols, opacities.reshape(1, N).float(), W, H, 16,
                                       offsets, flatten_ids, backgrounds=bg, max_hit=int(s.max_hit))
        image = rc[0, ..., :-1].permute(2, 0, 1)
        depth = rc[0, ..., -1:].permute(2, 0, 1)
        alpha = ra[0].permute(2, 0, 1)
        return image, depth, alpha, radii[0]
