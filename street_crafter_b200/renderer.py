"""Host-side mirror of the reference's `StreetGaussianRenderer.render_kernel_gsplat`
(/root/reference/street_gaussian/models/street_gaussian_renderer.py:172-302): the same five
op calls, in the same order, with the same arguments, over plain tensors instead of the
scene-graph model (which is out of scope).  bench.py and the parity tests
drive the kernels through this function, i.e. through the very call sequence the reference
executes once to four times per camera."""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch
from torch import Tensor

from .rendering import (fully_fused_projection, fused_view, isect_offset_encode, isect_tiles,
                        rasterize_to_pixels, spherical_harmonics)


def render_kernel_gsplat(xyz3: Tensor, rgb3: Tensor, scale: Tensor, quats: Tensor, occ1: Tensor,
                         w2c: Tensor, K: Tensor, camera_center: Tensor, width: int, height: int,
                         max_sh_degree: int, znear: float = 0.001, zfar: float = 1000.0,
                         tile_size: int = 16, use_depth: bool = True, absgrad: bool = True,
                         antialiasing: bool = True, train: bool = True,
                         grad_sinks: Optional[Dict[str, Tensor]] = None, fused: bool = True,
                         aux_ids: bool = True, prepared: Optional[Dict] = None,
                         deferred: Optional[list] = None, defer_sh: bool = False) -> Dict[str, Tensor]:
    """xyz3[N,3] rgb3[N,K,3] (SH) scale[N,3] quats[N,4] occ1[N,1] w2c[1,4,4] K[1,3,3]
    camera_center[3] -> the reference's result dict (rgb[3,H,W], acc, depth, viewspace_points,
    visibility_filter, radii) plus the raw render tensors.

    fused=True (default) runs the same computation as ONE autograd node per camera
    (rendering.fused_view): the elementwise glue between the five ops (opacity * compensation,
    dirs, mask, +0.5 / clamp, cat with depth and their backward nodes) is folded into the
    kernels.  fused=False issues the five ops exactly as the reference does (:219-280).
    aux_ids=False (fused only): the sorted 64-bit isect_ids — an intermediate the reference's result
    dict does not hold — are not written out; out["aux"]["isect_ids"] is then empty.
    prepared (fused only): the state of `prepare_view` — projection + binning of this very camera already done
    (two-phase schedule of the view-sharded trainer); only SH + compositing run here."""
    if fused and use_depth and rgb3.dim() == 3 and rgb3.shape[1] in (4, 16) and tile_size == 16:
        render_colors, render_alphas, meta = fused_view(
            xyz3, quats, scale, occ1[:, 0], rgb3, w2c, K, camera_center, width, height, max_sh_degree,
            near_plane=znear, far_plane=zfar, antialiased=antialiasing, tile_size=tile_size, absgrad=absgrad,
            grad_sinks=grad_sinks, need_isect_ids=aux_ids, prepared=prepared, deferred=deferred,
            defer_sh=defer_sh)
        radii, means2d = meta["radii"], meta["means2d"]
        rendered_color = render_colors[..., :-1]
        rendered_depth = meta["depth_norm"]  # render_colors[..., -1:] / render_alphas.clamp(min=1e-10), in-kernel
        if not train:
            rendered_color = torch.clamp(rendered_color, 0.0, 1.0)
        return {
            "rgb": rendered_color[0].permute(2, 0, 1),
            "acc": render_alphas[..., 0],
            "depth": rendered_depth[..., 0],
            "viewspace_points": means2d,
            "visibility_filter": meta["visibility"][0],     # radii[0] > 0, written by the projection kernel
            "radii": meta["radii_norm"][0],                  # radii[0] / float(max(height, width)), likewise
            "render_colors": render_colors,
            "render_alphas": render_alphas,
            "aux": dict(radii=radii, means2d=means2d, depths=meta["depths"], conics=meta["conics"],
                        compensations=meta["compensations"], tiles_per_gauss=meta["tiles_per_gauss"],
                        isect_ids=meta["isect_ids"], flatten_ids=meta["flatten_ids"],
                        isect_offsets=meta["isect_offsets"], colors=meta["colors"], opacities=meta["opacities"]),
        }
    # Project Gaussians to 2D                                   (reference :219-232)
    radii, means2d, depths, conics, compensations = fully_fused_projection(
        xyz3, None, quats, scale, w2c, K, width, height, packed=False, near_plane=znear, far_plane=zfar,
        calc_compensations=antialiasing, grad_sinks=grad_sinks)
    opacities = occ1[None, :, 0]
    if compensations is not None:
        opacities = opacities * compensations
    # Identify intersecting tiles                               (reference :243-253)
    tile_width = math.ceil(width / float(tile_size))
    tile_height = math.ceil(height / float(tile_size))
    tiles_per_gauss, isect_ids, flatten_ids = isect_tiles(means2d, radii, depths, tile_size, tile_width,
                                                          tile_height, packed=False, n_cameras=1)
    isect_offsets = isect_offset_encode(isect_ids, 1, tile_width, tile_height)
    # Spherical harmonics                                       (reference :256-260)
    dirs = xyz3[None, :, :] - camera_center
    masks = radii > 0
    # the reference writes rgb3.expand(1, -1, -1, -1) (:258); unsqueeze is the same view but its
    # backward is a view too, where ExpandBackward launches a 384 MB sum over the size-1 dim
    shs = rgb3[None]
    colors = spherical_harmonics(max_sh_degree, dirs, shs, masks=masks, grad_sinks=grad_sinks)
    colors = torch.clamp_min(colors + 0.5, 0.0)
    if train and means2d.requires_grad:
        means2d.retain_grad()
    if use_depth:
        colors = torch.cat((colors, depths[..., None]), dim=-1)
    render_colors, render_alphas = rasterize_to_pixels(
        means2d, conics, colors, opacities, width, height, tile_size, isect_offsets, flatten_ids,
        backgrounds=None, packed=False, absgrad=absgrad)
    if use_depth:
        rendered_color = render_colors[..., :-1]
        rendered_depth = render_colors[..., -1:] / render_alphas.clamp(min=1e-10)
    else:
        rendered_color = render_colors
        rendered_depth = render_alphas
    rendered_acc = render_alphas
    if not train:
        rendered_color = torch.clamp(rendered_color, 0.0, 1.0)
    return {
        "rgb": rendered_color[0].permute(2, 0, 1),
        "acc": rendered_acc[..., 0],
        "depth": rendered_depth[..., 0],
        "viewspace_points": means2d,
        "visibility_filter": radii[0] > 0,
        "radii": radii[0] / float(max(height, width)),
        # raw tensors (not in the reference dict) for tests / bench
        "render_colors": render_colors,
        "render_alphas": render_alphas,
        "aux": dict(radii=radii, means2d=means2d, depths=depths, conics=conics, compensations=compensations,
                    tiles_per_gauss=tiles_per_gauss, isect_ids=isect_ids, flatten_ids=flatten_ids,
                    isect_offsets=isect_offsets, colors=colors, opacities=opacities),
    }
