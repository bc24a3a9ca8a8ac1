"""CPU oracle for the Gaussian-splatting hot path (TEST INFRASTRUCTURE, NOT PRODUCT).

PARITY UNPINNED for the gsplat ops: the reference (zzz5y/street_crafter) calls the
third-party package `gsplat` (unpinned fork HEAD, README.md:34-36) which is not vendored
in /root/reference and not installable here; the reference holds no tests or golden
vectors.  This file restates the published algorithm of gsplat 1.x's
pure-torch reference ops (`gsplat/cuda/_torch_impl.py`: _quat_scale_to_covar_preci,
_world_to_cam, _persp_proj, _fully_fused_projection, _isect_tiles, _isect_offset_encode,
_rasterize_to_pixels, _spherical_harmonics), anchored on the reference's own call sites
`street_gaussian/models/street_gaussian_renderer.py:219-280`.  The SH basis IS pinned
against the reference's own `street_gaussian/utils/sh_utils.py:57-112` (tests/golden).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg
may import this module.  The product (street_crafter_b200/) never does.

Canonical arithmetic: everything that feeds an integer result (radius, tile bounds, depth
key) is written as explicit scalar fp32 operations with a fixed association order and no
matmul / einsum / fused multiply-add, so that the CUDA kernels (compiled -fmad=false, IEEE
div/sqrt) reproduce it bit for bit.  Backward is torch autograd through these forwards.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import numpy as np
import torch
from torch import Tensor

# Frozen constants (SURVEY.md §8c "constants the oracle must freeze").
EPS2D = 0.3
EIGEN_FLOOR = 0.01
ALPHA_CAP = 0.999
ALPHA_FLOOR = 1.0 / 255.0
T_STOP = 1e-4
JAC_PAD = 0.3

SH_C0 = 0.28209479177387814
SH_C1 = 0.4886025119029199
SH_C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005,
         -1.0925484305920792, 0.5462742152960396]
SH_C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154,
         -0.4570457994644658, 1.445305721320277, -0.5900435899266435]
SH_C4 = [2.5033429417967046, -1.7701307697799304, 0.9461746957575601, -0.6690465435572892,
         0.10578554691520431, -0.6690465435572892, 0.47308734787878004, -1.7701307697799304,
         0.6258357354491761]


class _SqrtRN(torch.autograd.Function):
    """Correctly rounded (IEEE round-to-nearest) square root.  torch.sqrt on CPU goes through a
    vectorised SLEEF routine that is NOT correctly rounded (~0.6 % of fp32 results are 1 ulp off
    numpy / the FPU / CUDA's sqrt.rn), which would make the integer-feeding chain
    implementation-defined.  numpy's sqrt is the hardware instruction."""

    @staticmethod
    def forward(ctx, x):
        y = torch.from_numpy(np.sqrt(x.detach().contiguous().numpy()))
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, g):
        (y,) = ctx.saved_tensors
        return g * 0.5 / y


def sqrt_rn(x: Tensor) -> Tensor:
    return _SqrtRN.apply(x)


# --------------------------------------------------------------------------------------
# A1: fully_fused_projection   (street_gaussian_renderer.py:219-232)
# --------------------------------------------------------------------------------------
def quat_scale_to_covar(quats: Tensor, scales: Tensor):
    """quats [N,4] wxyz (un-normalised), scales [N,3] -> 6 unique covariance entries.

    Restates gsplat `_quat_scale_to_covar_preci` with explicit op order."""
    w, x, y, z = quats.unbind(-1)
    n2 = ((w * w + x * x) + y * y) + z * z
    nrm = sqrt_rn(n2).clamp_min(1e-12)
    w, x, y, z = w / nrm, x / nrm, y / nrm, z / nrm
    x2, y2, z2 = x * x, y * y, z * z
    xy, xz, yz = x * y, x * z, y * z
    wx, wy, wz = w * x, w * y, w * z
    R = [[1.0 - 2.0 * (y2 + z2), 2.0 * (xy - wz), 2.0 * (xz + wy)],
         [2.0 * (xy + wz), 1.0 - 2.0 * (x2 + z2), 2.0 * (yz - wx)],
         [2.0 * (xz - wy), 2.0 * (yz + wx), 1.0 - 2.0 * (x2 + y2)]]
    s = scales.unbind(-1)
    M = [[R[i][j] * s[j] for j in range(3)] for i in range(3)]

    def dot(i, j):
        return (M[i][0] * M[j][0] + M[i][1] * M[j][1]) + M[i][2] * M[j][2]

    # order: 00 01 02 11 12 22
    return dot(0, 0), dot(0, 1), dot(0, 2), dot(1, 1), dot(1, 2), dot(2, 2)


def fully_fused_projection(means: Tensor, covars, quats: Tensor, scales: Tensor, viewmats: Tensor,
                           Ks: Tensor, width: int, height: int, eps2d: float = EPS2D,
                           near_plane: float = 0.01, far_plane: float = 1e10,
                           radius_clip: float = 0.0, packed: bool = False,
                           calc_compensations: bool = False):
    """Returns (radii[C,N] i32, means2d[C,N,2], depths[C,N], conics[C,N,3], compensations[C,N]|None).

    Culled Gaussians get radius 0 and all float outputs 0 (upstream leaves them
    uninitialised)."""
    assert covars is None and not packed
    C = viewmats.shape[0]
    dt = means.dtype
    s00, s01, s02, s11, s12, s22 = quat_scale_to_covar(quats, scales)
    S = [[s00, s01, s02], [s01, s11, s12], [s02, s12, s22]]
    px, py, pz = means.unbind(-1)
    outs = []
    for c in range(C):
        Wm = viewmats[c]
        Wr = [[Wm[i, j] for j in range(3)] for i in range(3)]
        t = [Wm[i, 3] for i in range(3)]
        fx, fy, cx, cy = Ks[c, 0, 0], Ks[c, 1, 1], Ks[c, 0, 2], Ks[c, 1, 2]
        # world -> camera
        xc = ((Wr[0][0] * px + Wr[0][1] * py) + Wr[0][2] * pz) + t[0]
        yc = ((Wr[1][0] * px + Wr[1][1] * py) + Wr[1][2] * pz) + t[1]
        zc = ((Wr[2][0] * px + Wr[2][1] * py) + Wr[2][2] * pz) + t[2]
        A = [[(Wr[i][0] * S[0][j] + Wr[i][1] * S[1][j]) + Wr[i][2] * S[2][j] for j in range(3)]
             for i in range(3)]

        def cc(i, j):
            return (A[i][0] * Wr[j][0] + A[i][1] * Wr[j][1]) + A[i][2] * Wr[j][2]

        c00, c01, c02, c11, c12, c22 = cc(0, 0), cc(0, 1), cc(0, 2), cc(1, 1), cc(1, 2), cc(2, 2)
        # perspective projection with principal-point-aware frustum clamp
        Wf = torch.tensor(float(width), dtype=dt)
        Hf = torch.tensor(float(height), dtype=dt)
        half = torch.tensor(0.5, dtype=dt)
        pad = torch.tensor(JAC_PAD, dtype=dt)
        tan_fovx = half * Wf / fx
        tan_fovy = half * Hf / fy
        lim_x_pos = (Wf - cx) / fx + pad * tan_fovx
        lim_x_neg = cx / fx + pad * tan_fovx
        lim_y_pos = (Hf - cy) / fy + pad * tan_fovy
        lim_y_neg = cy / fy + pad * tan_fovy
        # guard the division for points on the camera plane (they are culled below)
        zs = torch.where(zc == 0, torch.ones_like(zc), zc)
        tx = zs * torch.minimum(torch.maximum(xc / zs, -lim_x_neg), lim_x_pos)
        ty = zs * torch.minimum(torch.maximum(yc / zs, -lim_y_neg), lim_y_pos)
        z2 = zs * zs
        J00 = fx / zs
        J02 = -(fx * tx) / z2
        J11 = fy / zs
        J12 = -(fy * ty) / z2
        T00 = J00 * c00 + J02 * c02
        T01 = J00 * c01 + J02 * c12
        T02 = J00 * c02 + J02 * c22
        T11 = J11 * c11 + J12 * c12
        T12 = J11 * c12 + J12 * c22
        v00 = T00 * J00 + T02 * J02
        v01 = T01 * J11 + T02 * J12
        v11 = T11 * J11 + T12 * J12
        mx = (fx * xc) / zs + cx
        my = (fy * yc) / zs + cy
        det_orig = v00 * v11 - v01 * v01
        e = torch.tensor(eps2d, dtype=dt)
        b00 = v00 + e
        b11 = v11 + e
        det = b00 * b11 - v01 * v01
        valid = (det > 0) & (zc > near_plane) & (zc < far_plane)
        dets = torch.where(det > 0, det, torch.ones_like(det))
        comp = sqrt_rn((det_orig / dets).clamp_min(0.0))
        conic_a = b11 / dets
        conic_b = -v01 / dets
        conic_c = b00 / dets
        b = half * (b00 + b11)
        v1 = b + sqrt_rn((b * b - det).clamp_min(EIGEN_FLOOR))
        radius = torch.ceil(3.0 * sqrt_rn(v1))
        valid = valid & (radius > radius_clip)
        inside = (mx + radius > 0) & (mx - radius < Wf) & (my + radius > 0) & (my - radius < Hf)
        valid = valid & inside
        # NaN-proof: invalid entries are zeroed
        zero = torch.zeros_like(mx)
        radius_i = torch.where(valid, radius, zero).to(torch.int32)
        outs.append((radius_i,
                     torch.stack([torch.where(valid, mx, zero), torch.where(valid, my, zero)], -1),
                     torch.where(valid, zc, zero),
                     torch.stack([torch.where(valid, conic_a, zero), torch.where(valid, conic_b, zero),
                                  torch.where(valid, conic_c, zero)], -1),
                     torch.where(valid, comp, zero)))
    radii = torch.stack([o[0] for o in outs])
    means2d = torch.stack([o[1] for o in outs])
    depths = torch.stack([o[2] for o in outs])
    conics = torch.stack([o[3] for o in outs])
    comps = torch.stack([o[4] for o in outs]) if calc_compensations else None
    return radii, means2d, depths, conics, comps


# --------------------------------------------------------------------------------------
# A5: spherical_harmonics   (street_gaussian_renderer.py:256-260; sh_utils.py:57-112)
# --------------------------------------------------------------------------------------
def sh_bases(deg: int, dirs: Tensor):
    """Real SH basis values Y_k(d̂) for k < (deg+1)^2; dirs are normalised here
    (gsplat normalises inside the kernel).  Sign/ordering convention identical to
    street_gaussian/utils/sh_utils.py:77-110."""
    n = torch.sqrt(((dirs[..., 0] * dirs[..., 0] + dirs[..., 1] * dirs[..., 1])
                    + dirs[..., 2] * dirs[..., 2])).clamp_min(1e-12)
    x, y, z = dirs[..., 0] / n, dirs[..., 1] / n, dirs[..., 2] / n
    Y = [torch.full_like(x, SH_C0)]
    if deg > 0:
        Y += [-SH_C1 * y, SH_C1 * z, -SH_C1 * x]
    if deg > 1:
        xx, yy, zz = x * x, y * y, z * z
        xy, yz, xz = x * y, y * z, x * z
        Y += [SH_C2[0] * xy, SH_C2[1] * yz, SH_C2[2] * (2.0 * zz - xx - yy), SH_C2[3] * xz,
              SH_C2[4] * (xx - yy)]
    if deg > 2:
        Y += [SH_C3[0] * y * (3 * xx - yy), SH_C3[1] * xy * z, SH_C3[2] * y * (4 * zz - xx - yy),
              SH_C3[3] * z * (2 * zz - 3 * xx - 3 * yy), SH_C3[4] * x * (4 * zz - xx - yy),
              SH_C3[5] * z * (xx - yy), SH_C3[6] * x * (xx - 3 * yy)]
    if deg > 3:
        Y += [SH_C4[0] * xy * (xx - yy), SH_C4[1] * yz * (3 * xx - yy), SH_C4[2] * xy * (7 * zz - 1),
              SH_C4[3] * yz * (7 * zz - 3), SH_C4[4] * (zz * (35 * zz - 30) + 3),
              SH_C4[5] * xz * (7 * zz - 3), SH_C4[6] * (xx - yy) * (7 * zz - 1),
              SH_C4[7] * xz * (xx - 3 * yy),
              SH_C4[8] * (xx * (xx - 3 * yy) - yy * (3 * xx - yy))]
    return torch.stack(Y, -1)  # [..., K]


def spherical_harmonics(degrees_to_use: int, dirs: Tensor, coeffs: Tensor,
                        masks: Optional[Tensor] = None) -> Tensor:
    """dirs [...,3] (un-normalised), coeffs [...,K,3], masks [...] bool -> colours [...,3].
    Masked-out entries are 0 (value and gradient)."""
    nb = (degrees_to_use + 1) ** 2
    assert coeffs.shape[-2] >= nb
    Y = sh_bases(degrees_to_use, dirs)  # [..., nb]
    col = (Y[..., :, None] * coeffs[..., :nb, :]).sum(-2)
    if masks is not None:
        col = torch.where(masks[..., None], col, torch.zeros_like(col))
    return col


# --------------------------------------------------------------------------------------
# A3/A4: isect_tiles / isect_offset_encode   (street_gaussian_renderer.py:243-253)
# --------------------------------------------------------------------------------------
def tile_bounds(means2d: Tensor, radii: Tensor, tile_size: int, tile_width: int, tile_height: int):
    ts = float(tile_size)
    r = radii.to(means2d.dtype) / ts
    tx = means2d[..., 0] / ts
    ty = means2d[..., 1] / ts
    x0 = torch.floor(tx - r).clamp(0, tile_width).to(torch.int64)
    x1 = torch.ceil(tx + r).clamp(0, tile_width).to(torch.int64)
    y0 = torch.floor(ty - r).clamp(0, tile_height).to(torch.int64)
    y1 = torch.ceil(ty + r).clamp(0, tile_height).to(torch.int64)
    vis = radii > 0
    z = torch.zeros_like(x0)
    return (torch.where(vis, x0, z), torch.where(vis, x1, z), torch.where(vis, y0, z),
            torch.where(vis, y1, z))


def isect_tiles(means2d: Tensor, radii: Tensor, depths: Tensor, tile_size: int, tile_width: int,
                tile_height: int, sort: bool = True, packed: bool = False, n_cameras: int = 1):
    """-> (tiles_per_gauss[C,N] i32, isect_ids[I] i64, flatten_ids[I] i32).

    key = cam << (32+tile_n_bits) | tile << 32 | float_bits(depth); value = cam*N + gauss;
    emission order: camera, Gaussian index, tile row-major; sort is stable."""
    assert not packed
    C, N = radii.shape
    x0, x1, y0, y1 = tile_bounds(means2d.detach(), radii, tile_size, tile_width, tile_height)
    w = x1 - x0
    tpg = (w * (y1 - y0)).reshape(-1)  # [C*N]
    n_tiles = tile_width * tile_height
    tile_n_bits = int(math.floor(math.log2(n_tiles))) + 1
    I = int(tpg.sum())
    flat = torch.repeat_interleave(torch.arange(C * N, dtype=torch.int64), tpg)
    start = torch.cumsum(tpg, 0) - tpg
    k = torch.arange(I, dtype=torch.int64) - start[flat]
    wf = w.reshape(-1)[flat]
    ti = y0.reshape(-1)[flat] + k // wf.clamp_min(1)
    tj = x0.reshape(-1)[flat] + k % wf.clamp_min(1)
    tile_id = ti * tile_width + tj
    cam = flat // N
    dbits = depths.detach().reshape(-1).to(torch.float32).contiguous().view(torch.int32)[flat]
    dbits = dbits.to(torch.int64) & 0xFFFFFFFF
    ids = (cam << (32 + tile_n_bits)) | (tile_id << 32) | dbits
    flatten = flat.to(torch.int32)
    if sort:
        ids, order = torch.sort(ids, stable=True)
        flatten = flatten[order]
    return tpg.reshape(C, N).to(torch.int32), ids, flatten


def isect_offset_encode(isect_ids: Tensor, n_cameras: int, tile_width: int, tile_height: int) -> Tensor:
    n_tiles = tile_width * tile_height
    tile_n_bits = int(math.floor(math.log2(n_tiles))) + 1
    cam = isect_ids >> (32 + tile_n_bits)
    tile = (isect_ids >> 32) & ((1 << tile_n_bits) - 1)
    slot = cam * n_tiles + tile
    queries = torch.arange(n_cameras * n_tiles, dtype=torch.int64)
    off = torch.searchsorted(slot.contiguous(), queries, right=False)
    return off.to(torch.int32).reshape(n_cameras, tile_height, tile_width)


# --------------------------------------------------------------------------------------
# A7/A8: rasterize_to_pixels   (street_gaussian_renderer.py:267-280)
# --------------------------------------------------------------------------------------
def rasterize_to_pixels(means2d: Tensor, conics: Tensor, colors: Tensor, opacities: Tensor,
                        image_width: int, image_height: int, tile_size: int, isect_offsets: Tensor,
                        flatten_ids: Tensor, backgrounds: Optional[Tensor] = None,
                        packed: bool = False, absgrad: bool = False, max_hit: int = 0,
                        return_stats: bool = False, return_margin: bool = False):
    """-> (render_colors[C,H,W,CH], render_alphas[C,H,W,1]).

    return_margin: also returns margin[C,H,W] = the smallest RELATIVE distance of any visited
    (pixel, Gaussian) pair to one of the algorithm's hard discontinuities (alpha = 1/255 accept test,
    T = 1e-4 stop test): the tests require every pixel that misses the image tolerance to sit on one.

    Per pixel centre (j+.5, i+.5), front to back over the tile's sorted list:
      sigma = .5(a dx² + c dy²) + b dx dy;  alpha = min(.999, o·exp(-sigma));
      skip if sigma < 0 or alpha < 1/255; stop BEFORE a Gaussian that would take T to <= 1e-4.
    max_hit > 0 (LiDAR point-splat mode): stop after max_hit contributing splats.
    `absgrad`: gradient bookkeeping only exists in the CUDA op; here |grad| of means2d is
    obtained by the test through per-pixel autograd (see tests)."""
    assert not packed
    C, N = opacities.shape
    CH = colors.shape[-1]
    dt = means2d.dtype
    tile_h, tile_w = isect_offsets.shape[1:]
    n_isects = flatten_ids.shape[0]
    offs = isect_offsets.reshape(-1).tolist() + [n_isects]
    out_c = torch.zeros(C, tile_h * tile_size, tile_w * tile_size, CH, dtype=dt)
    out_a = torch.zeros(C, tile_h * tile_size, tile_w * tile_size, 1, dtype=dt)
    m2 = means2d.reshape(-1, 2)
    cn = conics.reshape(-1, 3)
    col = colors.reshape(-1, CH)
    op = opacities.reshape(-1)
    ii, jj = torch.meshgrid(torch.arange(tile_size), torch.arange(tile_size), indexing="ij")
    visited = 0
    contributing = 0
    tiles_c, tiles_a, tiles_pos = [], [], []
    margin = torch.full((C, tile_h * tile_size, tile_w * tile_size), float("inf"), dtype=dt) if return_margin else None
    for c in range(C):
        for t in range(tile_h * tile_w):
            s, e = offs[c * tile_h * tile_w + t], offs[c * tile_h * tile_w + t + 1]
            if e <= s:
                continue
            ty, tx = divmod(t, tile_w)
            g = flatten_ids[s:e].long()
            px = (tx * tile_size + jj.reshape(-1)).to(dt) + 0.5  # [P]
            py = (ty * tile_size + ii.reshape(-1)).to(dt) + 0.5
            dx = m2[g, 0][None, :] - px[:, None]  # [P,G]
            dy = m2[g, 1][None, :] - py[:, None]
            a, b, cc = cn[g, 0][None], cn[g, 1][None], cn[g, 2][None]
            sigma = 0.5 * (a * dx * dx + cc * dy * dy) + b * dx * dy
            alpha = torch.clamp_max(op[g][None] * torch.exp(-sigma), ALPHA_CAP)
            raw_alpha = alpha
            keep = (sigma >= 0) & (alpha >= ALPHA_FLOOR)
            if max_hit > 0:
                keep = keep & (torch.cumsum(keep.to(torch.int32), 1) <= max_hit)
            alpha = torch.where(keep, alpha, torch.zeros_like(alpha))
            Tn = torch.cumprod(1.0 - alpha, dim=1)
            live = Tn > T_STOP
            Tp = torch.cat([torch.ones_like(Tn[:, :1]), Tn[:, :-1]], 1)
            wgt = torch.where(live, alpha * Tp, torch.zeros_like(alpha))  # [P,G]
            pix = wgt @ col[g]  # [P,CH]
            Tfin = torch.prod(torch.where(live, 1.0 - alpha, torch.ones_like(alpha)), dim=1)
            if return_margin:
                with torch.no_grad():
                    visited_m = Tp > T_STOP  # pairs the front-to-back walk reaches
                    big = torch.full_like(alpha, float("inf"))
                    m_a = torch.where(visited_m, (raw_alpha - ALPHA_FLOOR).abs() / ALPHA_FLOOR, big)
                    m_t = torch.where(visited_m, (Tn - T_STOP).abs() / T_STOP, big)
                    mm = torch.minimum(m_a, m_t).min(dim=1).values
                    margin[c, ty * tile_size:(ty + 1) * tile_size, tx * tile_size:(tx + 1) * tile_size] = \
                        mm.reshape(tile_size, tile_size)
            if return_stats:
                inb = ((px < image_width) & (py < image_height))[:, None]
                nlive = live.sum(1, keepdim=True)
                visited += int((torch.clamp_max(nlive + 1, live.shape[1]) * inb).sum())
                contributing += int((keep & live & inb).sum())
            tiles_c.append(pix.reshape(tile_size, tile_size, CH))
            tiles_a.append((1.0 - Tfin).reshape(tile_size, tile_size, 1))
            tiles_pos.append((c, ty, tx))
    # assemble (index_put keeps autograd happy)
    for (c, ty, tx), pc_, pa_ in zip(tiles_pos, tiles_c, tiles_a):
        out_c[c, ty * tile_size:(ty + 1) * tile_size, tx * tile_size:(tx + 1) * tile_size] = pc_
        out_a[c, ty * tile_size:(ty + 1) * tile_size, tx * tile_size:(tx + 1) * tile_size] = pa_
    out_c = out_c[:, :image_height, :image_width]
    out_a = out_a[:, :image_height, :image_width]
    if backgrounds is not None:
        out_c = out_c + backgrounds[:, None, None, :] * (1.0 - out_a)
    if return_margin:
        margin = margin[:, :image_height, :image_width]
        if return_stats:
            return out_c, out_a, {"visited": visited, "contributing": contributing, "margin": margin}
        return out_c, out_a, margin
    if return_stats:
        return out_c, out_a, {"visited": visited, "contributing": contributing}
    return out_c, out_a


# --------------------------------------------------------------------------------------
# The reference renderer's call sequence (street_gaussian_renderer.py:214-300), on the oracle.
# --------------------------------------------------------------------------------------
def render_reference_sequence(means, quats, scales, opacities, shs, viewmat, K, width, height,
                              sh_degree, near=0.001, far=1000.0, tile_size=16, antialiasing=True,
                              cam_center=None, return_aux=False):
    """means[N,3] quats[N,4] scales[N,3] opacities[N] shs[N,K,3] viewmat[4,4] K[3,3]
    -> render_colors[1,H,W,4] (RGB + depth-weighted), render_alphas[1,H,W,1]."""
    radii, means2d, depths, conics, comps = fully_fused_projection(
        means, None, quats, scales, viewmat[None], K[None], width, height,
        near_plane=near, far_plane=far, calc_compensations=antialiasing)
    op = opacities[None]
    if comps is not None:
        op = op * comps
    tw = math.ceil(width / float(tile_size))
    th = math.ceil(height / float(tile_size))
    tpg, isect_ids, flatten_ids = isect_tiles(means2d, radii, depths, tile_size, tw, th, n_cameras=1)
    isect_offsets = isect_offset_encode(isect_ids, 1, tw, th)
    if cam_center is None:
        Rm = viewmat[:3, :3]
        cam_center = -(Rm.T @ viewmat[:3, 3])
    dirs = means[None] - cam_center
    colors = spherical_harmonics(sh_degree, dirs, shs[None], masks=radii > 0)
    colors = torch.clamp_min(colors + 0.5, 0.0)
    colors = torch.cat((colors, depths[..., None]), dim=-1)
    if return_aux:
        rc, ra, margin = rasterize_to_pixels(means2d, conics, colors, op, width, height, tile_size,
                                             isect_offsets, flatten_ids, return_margin=True)
    else:
        rc, ra = rasterize_to_pixels(means2d, conics, colors, op, width, height, tile_size,
                                     isect_offsets, flatten_ids)
    if return_aux:
        return rc, ra, dict(margin=margin, radii=radii, means2d=means2d, depths=depths, conics=conics,
                            compensations=comps, tiles_per_gauss=tpg, isect_ids=isect_ids,
                            flatten_ids=flatten_ids, isect_offsets=isect_offsets, colors=colors,
                            opacities=op)
    return rc, ra
