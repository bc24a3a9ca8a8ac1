"""Drop-in for the `gsplat.rendering` names the reference renderer imports
(/root/reference/street_gaussian/models/street_gaussian_renderer.py:204):

    rasterization, fully_fused_projection, isect_tiles, isect_offset_encode,
    rasterize_to_pixels, spherical_harmonics

Same positional order, keyword names, shapes, dtypes and autograd conventions as the
reference's call sites (street_gaussian_renderer.py:219-280): `radii[C,N]` int32,
`packed=False` only, `means2d.absgrad` set as a side effect of the rasterizer's backward when
`absgrad=True` (consumed at street_gaussian/models/street_gaussian_model.py:505-506).

Host code is plumbing only: tensors are allocated with torch, every computation is a
hand-written sm_100a kernel reached through the C ABI in include/scb200.h.  There is no CPU
path: CPU tensors raise."""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Tuple

import torch
from torch import Tensor

from ._lib import check, lib


# ------------------------------------------------------------------------------------------
# helpers
# ------------------------------------------------------------------------------------------
def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t: Optional[Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _f32c(t: Tensor, name: str, align: int = 16) -> Tensor:
    """contiguous fp32 CUDA tensor whose base pointer is `align`-byte aligned."""
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor: street_crafter_b200 has no CPU path")
    if t.dtype != torch.float32:
        raise AssertionError(f"{name} must be float32, got {t.dtype}")
    t = t.contiguous()
    if t.data_ptr() % align:
        t = t.clone(memory_format=torch.contiguous_format)
    return t


def _grad_or_zeros(g: Optional[Tensor], like_shape, device) -> Tensor:
    if g is None:
        return torch.zeros(like_shape, dtype=torch.float32, device=device)
    return _f32c(g, "grad")


# ------------------------------------------------------------------------------------------
# A1  fully_fused_projection
# ------------------------------------------------------------------------------------------
class _FullyFusedProjection(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means, quats, scales, viewmats, Ks, width, height, eps2d, near_plane, far_plane,
                radius_clip, calc_compensations, sinks=None):
        means = _f32c(means, "means")
        quats = _f32c(quats, "quats")
        scales = _f32c(scales, "scales")
        viewmats = _f32c(viewmats, "viewmats")
        Ks = _f32c(Ks, "Ks")
        N, C = means.shape[0], viewmats.shape[0]
        dev = means.device
        radii = torch.empty((C, N), dtype=torch.int32, device=dev)
        means2d = torch.empty((C, N, 2), dtype=torch.float32, device=dev)
        depths = torch.empty((C, N), dtype=torch.float32, device=dev)
        conics = torch.empty((C, N, 3), dtype=torch.float32, device=dev)
        comps = torch.empty((C, N), dtype=torch.float32, device=dev) if calc_compensations else None
        with torch.cuda.device(dev):
            check(lib.sc_projection_fwd(N, C, _p(means), _p(quats), _p(scales), _p(viewmats), _p(Ks),
                                        int(width), int(height), float(eps2d), float(near_plane),
                                        float(far_plane), float(radius_clip), _p(radii), _p(means2d),
                                        _p(depths), _p(conics), _p(comps), _stream()),
                  "sc_projection_fwd")
        ctx.save_for_backward(means, quats, scales, viewmats, Ks, radii)
        ctx.set_materialize_grads(False)
        ctx.dims = (int(width), int(height), float(eps2d), bool(calc_compensations))
        ctx.sinks = sinks
        ctx.mark_non_differentiable(radii)
        if comps is None:
            return radii, means2d, depths, conics
        return radii, means2d, depths, conics, comps

    @staticmethod
    def backward(ctx, v_radii, v_means2d, v_depths, v_conics, v_comps=None):
        means, quats, scales, viewmats, Ks, radii = ctx.saved_tensors
        width, height, eps2d, has_comp = ctx.dims
        N, C = means.shape[0], viewmats.shape[0]
        dev = means.device
        v_means2d = _grad_or_zeros(v_means2d, (C, N, 2), dev)
        v_depths = _grad_or_zeros(v_depths, (C, N), dev)
        v_conics = _grad_or_zeros(v_conics, (C, N, 3), dev)
        if has_comp and v_comps is not None:
            v_comps = _f32c(v_comps, "v_compensations")
        else:
            v_comps = None
        need = ctx.needs_input_grad
        sinks = ctx.sinks
        if sinks is not None:
            # gradient-arena mode: += straight into the caller's buffers, nothing flows to autograd
            assert C == 1, "gradient sinks need C == 1"
            v_means, v_quats, v_scales = (sinks.get(k) for k in ("means", "quats", "scales"))
        else:
            v_means = torch.empty_like(means) if need[0] else None
            v_quats = torch.empty_like(quats) if need[1] else None
            v_scales = torch.empty_like(scales) if need[2] else None
        if need[3]:
            raise NotImplementedError("viewmats gradients are not produced (cameras are not optimised "
                                      "in the reference, street_gaussian/config/config.py:126)")
        with torch.cuda.device(dev):
            check(lib.sc_projection_bwd(N, C, _p(means), _p(quats), _p(scales), _p(viewmats), _p(Ks), width,
                                        height, eps2d, _p(radii), _p(v_means2d), _p(v_depths), _p(v_conics),
                                        _p(v_comps), _p(v_means), _p(v_quats), _p(v_scales),
                                        0 if sinks is None else 1, _stream()),
                  "sc_projection_bwd")
        if sinks is not None:
            return (None,) * 13
        return (v_means, v_quats, v_scales, None, None, None, None, None, None, None, None, None, None)


def fully_fused_projection(means: Tensor, covars: Optional[Tensor], quats: Optional[Tensor],
                           scales: Optional[Tensor], viewmats: Tensor, Ks: Tensor, width: int, height: int,
                           eps2d: float = 0.3, near_plane: float = 0.01, far_plane: float = 1e10,
                           radius_clip: float = 0.0, packed: bool = False, sparse_grad: bool = False,
                           calc_compensations: bool = False, camera_model: str = "pinhole",
                           grad_sinks: Optional[Dict[str, Tensor]] = None,
                           ) -> Tuple[Tensor, Tensor, Tensor, Tensor, Optional[Tensor]]:
    """-> (radii[C,N] i32, means2d[C,N,2], depths[C,N], conics[C,N,3], compensations[C,N]|None).

    `grad_sinks` (extension, C == 1): {"means","quats","scales"} -> contiguous fp32 buffers the
    backward ADDS its gradients into (the flat gradient arena of parallel.py) instead of
    returning them to autograd — no AccumulateGrad add kernels, culled Gaussians untouched."""
    N = means.shape[0]
    C = viewmats.shape[0]
    assert means.shape == (N, 3), means.shape
    assert viewmats.shape == (C, 4, 4), viewmats.shape
    assert Ks.shape == (C, 3, 3), Ks.shape
    if covars is not None:
        raise NotImplementedError("covars input is not used by the reference (it passes quats/scales)")
    if packed:
        raise NotImplementedError("packed=True is not used by the reference (packed=False at "
                                  "street_gaussian_renderer.py:228)")
    assert camera_model == "pinhole"
    assert quats is not None and scales is not None, "quats and scales are required"
    assert quats.shape == (N, 4), quats.shape
    assert scales.shape == (N, 3), scales.shape
    out = _FullyFusedProjection.apply(means, quats, scales, viewmats, Ks, width, height, eps2d, near_plane,
                                      far_plane, radius_clip, calc_compensations, grad_sinks)
    if calc_compensations:
        return out
    return out + (None,)


# ------------------------------------------------------------------------------------------
# A5  spherical_harmonics
# ------------------------------------------------------------------------------------------
class _SphericalHarmonics(torch.autograd.Function):
    @staticmethod
    def forward(ctx, degree, dirs, coeffs, masks, sinks=None):
        dirs = _f32c(dirs, "dirs", 4)
        coeffs = _f32c(coeffs, "coeffs")
        K = coeffs.shape[-2]
        M = dirs.numel() // 3
        m8 = None
        if masks is not None:
            m8 = masks.contiguous().view(torch.uint8) if masks.dtype == torch.bool else \
                masks.to(torch.uint8).contiguous()
        colors = torch.empty(dirs.shape, dtype=torch.float32, device=dirs.device)
        with torch.cuda.device(dirs.device):
            check(lib.sc_sh_fwd(M, K, int(degree), _p(dirs), _p(coeffs), _p(m8), _p(colors), _stream()),
                  "sc_sh_fwd")
        ctx.save_for_backward(dirs, coeffs, m8 if m8 is not None else torch.empty(0, device=dirs.device))
        ctx.degree = int(degree)
        ctx.has_mask = m8 is not None
        ctx.sinks = sinks
        return colors

    @staticmethod
    def backward(ctx, v_colors):
        dirs, coeffs, m8 = ctx.saved_tensors
        if not ctx.has_mask:
            m8 = None
        K = coeffs.shape[-2]
        M = dirs.numel() // 3
        v_colors = _f32c(v_colors, "v_colors", 4)
        sinks = ctx.sinks
        if sinks is not None:
            # gradient-arena mode: v_coeffs += into the sh slice, v_dirs += into the means slice
            # (dirs = means - camera centre), masked Gaussians untouched
            with torch.cuda.device(dirs.device):
                check(lib.sc_sh_bwd(M, K, ctx.degree, _p(dirs), _p(coeffs), _p(m8), _p(v_colors),
                                    _p(sinks["shs"]), _p(sinks.get("means")), 1, _stream()), "sc_sh_bwd")
            return None, None, None, None, None
        v_coeffs = torch.empty_like(coeffs)
        v_dirs = torch.empty_like(dirs) if ctx.needs_input_grad[1] else None
        with torch.cuda.device(dirs.device):
            check(lib.sc_sh_bwd(M, K, ctx.degree, _p(dirs), _p(coeffs), _p(m8), _p(v_colors), _p(v_coeffs),
                                _p(v_dirs), 0, _stream()), "sc_sh_bwd")
        return None, v_dirs, (v_coeffs if ctx.needs_input_grad[2] else None), None, None


def spherical_harmonics(degrees_to_use: int, dirs: Tensor, coeffs: Tensor,
                        masks: Optional[Tensor] = None,
                        grad_sinks: Optional[Dict[str, Tensor]] = None) -> Tensor:
    """dirs [...,3] (normalised inside the kernel), coeffs [...,K,3], masks [...] -> [...,3].

    `grad_sinks` (extension): {"shs": buffer of coeffs' shape, "means": buffer of dirs' shape}
    the backward ADDS into (see fully_fused_projection)."""
    assert (degrees_to_use + 1) ** 2 <= coeffs.shape[-2], coeffs.shape
    assert dirs.shape[:-1] == coeffs.shape[:-2], (dirs.shape, coeffs.shape)
    assert dirs.shape[-1] == 3 and coeffs.shape[-1] == 3
    if masks is not None:
        assert masks.shape == dirs.shape[:-1], masks.shape
    if grad_sinks is not None:
        assert grad_sinks["shs"].numel() == coeffs.numel() and grad_sinks["shs"].is_contiguous()
    return _SphericalHarmonics.apply(degrees_to_use, dirs, coeffs, masks, grad_sinks)


# ------------------------------------------------------------------------------------------
# A3/A4  isect_tiles / isect_offset_encode
# ------------------------------------------------------------------------------------------
def _tile_n_bits(tile_width: int, tile_height: int) -> int:
    return int(math.floor(math.log2(tile_width * tile_height))) + 1


# pinned landing buffers for the {n_visible, n_isects, n_items} read-back, one per (device, stream)
_PINNED_TOTALS: Dict[Tuple[int, int], Tensor] = {}


def _isect_begin(means2d: Tensor, radii: Tensor, depths: Tensor, tile_size: int, tile_width: int,
                 tile_height: int, tpg: Tensor, st: int, capacity: Optional[int] = None,
                 counts_host: Optional[Tensor] = None):
    """phase 1 of the fused binning pipeline: compaction of the emitting entries + their stable
    depth order + scan (one C call), then an ASYNC copy of {n_visible, n_isects} to pinned host
    memory and an event.  The caller may enqueue independent GPU work before _isect_finish()
    waits on the event, so the GPU stays busy across the host read-back.  One camera and a tile
    grid of at most 128 x 128 take the fast path (sc_bin_*), anything else the generic one."""
    C, N = radii.shape
    CN = C * N
    dev = means2d.device
    fast = bool(lib.sc_bin_supported(C, tile_width, tile_height))
    ws_bytes = int(lib.sc_isect_workspace_bytes(CN, tile_width, tile_height))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    if fast:
        check(lib.sc_bin_prepare(CN, _p(means2d), _p(radii), _p(depths), tile_size, tile_width, tile_height,
                                 _p(tpg), _p(ws), ws_bytes, st), "sc_bin_prepare")
    else:
        check(lib.sc_isect_prepare(CN, _p(means2d), _p(radii), _p(depths), tile_size, tile_width, tile_height,
                                   _p(tpg), _p(ws), ws_bytes, st), "sc_isect_prepare")
    if capacity is not None:
        # sync-free mode (fast path only): the counts stay on the device; a copy travels to the caller's pinned
        # buffer for the capacity check it makes LATER (nothing waits for it here)
        if not fast:
            raise RuntimeError("sync-free binning needs the fast path (one camera, tile grid <= 128 x 128)")
        counts = ws[:24].view(torch.int64).clone()
        if counts_host is not None:
            counts_host.copy_(counts, non_blocking=True)
        return (means2d, radii, depths, tile_size, tile_width, tile_height, tpg, st, ws, counts, int(capacity), fast)
    key = (dev.index if dev.index is not None else torch.cuda.current_device(), st)
    host = _PINNED_TOTALS.get(key)
    if host is None:
        host = _PINNED_TOTALS[key] = torch.empty(3, dtype=torch.int64).pin_memory()
    host.copy_(ws[:24].view(torch.int64), non_blocking=True)
    ev = torch.cuda.Event()
    ev.record()
    return (means2d, radii, depths, tile_size, tile_width, tile_height, tpg, st, ws, host, ev, fast)


def _isect_finish(state, need_ids: bool = True):
    """phase 2: wait for the counts (the one host sync of the pipeline), then emit + tile sort (one C
    call).  -> (tiles_per_gauss, isect_ids | None, flatten_ids, isect_offsets | None): the fast path
    also produces the tile offsets (and skips the 64-bit ids when `need_ids` is False)."""
    means2d, radii, depths, tile_size, tile_width, tile_height, tpg, st, ws, host, ev, fast = state
    C, N = radii.shape
    dev = means2d.device
    if isinstance(ev, int):   # sync-free mode: (…, counts on the device, capacity, fast)
        counts, capacity = host, ev
        flat = torch.empty(capacity, dtype=torch.int32, device=dev)
        ids = torch.empty(capacity, dtype=torch.int64, device=dev) if need_ids else None
        offsets = torch.empty((1, tile_height, tile_width), dtype=torch.int32, device=dev)
        sws_bytes = int(lib.sc_bin_sort_workspace_bytes(capacity))
        sws = torch.empty(sws_bytes, dtype=torch.uint8, device=dev)
        check(lib.sc_bin_finish_dev(N, _p(depths), tile_width, tile_height, capacity, _p(ws), _p(flat), _p(ids),
                                    _p(offsets), _p(sws), sws_bytes, st), "sc_bin_finish_dev")
        return tpg, ids, flat, offsets
    bits = _tile_n_bits(tile_width, tile_height)
    cam_bits = 0 if C == 1 else int(math.floor(math.log2(C))) + 1
    ev.synchronize()
    n_visible, n_isects, n_items = host.tolist()
    if fast:
        flat = torch.empty(n_isects, dtype=torch.int32, device=dev)
        ids = torch.empty(n_isects, dtype=torch.int64, device=dev) if need_ids else None
        offsets = torch.empty((1, tile_height, tile_width), dtype=torch.int32, device=dev)
        sws_bytes = int(lib.sc_bin_sort_workspace_bytes(n_isects))
        sws = torch.empty(sws_bytes, dtype=torch.uint8, device=dev)
        check(lib.sc_bin_finish(N, _p(depths), tile_width, tile_height, n_visible, n_isects, n_items, _p(ws), _p(flat),
                                _p(ids), _p(offsets), _p(sws), sws_bytes, st), "sc_bin_finish")
        return tpg, ids, flat, offsets
    ids = torch.empty((2, n_isects), dtype=torch.int64, device=dev)
    flat = torch.empty((2, n_isects), dtype=torch.int32, device=dev)
    if n_isects == 0:
        return tpg, ids[0], flat[0], None
    sws_bytes = int(lib.sc_sort_workspace_bytes(n_isects))
    sws = torch.empty(sws_bytes, dtype=torch.uint8, device=dev)
    check(lib.sc_isect_emit_sorted(N, C, _p(means2d), _p(radii), _p(depths), tile_size, tile_width,
                                   tile_height, bits, cam_bits, n_visible, n_isects, _p(ws),
                                   _p(ids[0]), _p(flat[0]), _p(ids[1]), _p(flat[1]), _p(sws),
                                   sws_bytes, st), "sc_isect_emit_sorted")
    return tpg, ids[0], flat[0], None


@torch.no_grad()
def isect_tiles(means2d: Tensor, radii: Tensor, depths: Tensor, tile_size: int, tile_width: int,
                tile_height: int, sort: bool = True, packed: bool = False, n_cameras: Optional[int] = None,
                camera_ids: Optional[Tensor] = None, gaussian_ids: Optional[Tensor] = None,
                ) -> Tuple[Tensor, Tensor, Tensor]:
    """-> (tiles_per_gauss[C,N] i32, isect_ids[I] i64 (sorted), flatten_ids[I] i32).

    One host read-back (the total number of intersections sizes the outputs), exactly as the
    reference's op does."""
    if packed:
        raise NotImplementedError("packed=True is not used by the reference")
    C, N = radii.shape
    assert means2d.shape == (C, N, 2) and depths.shape == (C, N), (means2d.shape, depths.shape)
    if n_cameras is not None:
        assert n_cameras == C
    dev = means2d.device
    means2d = _f32c(means2d.detach(), "means2d", 8)
    depths = _f32c(depths.detach(), "depths", 4)
    radii = radii.contiguous()
    assert radii.dtype == torch.int32
    CN = C * N
    bits = _tile_n_bits(tile_width, tile_height)
    cam_bits = 0 if C == 1 else int(math.floor(math.log2(C))) + 1
    assert 32 + bits + cam_bits <= 64
    st = _stream()
    end_bit = 32 + bits + cam_bits
    with torch.cuda.device(dev):
        tpg = torch.empty((C, N), dtype=torch.int32, device=dev)
        if sort:
            state = _isect_begin(means2d, radii, depths, tile_size, tile_width, tile_height, tpg, st)
            return _isect_finish(state)[:3]
        # unsorted (gsplat's sort=False): entry order, through the stand-alone entry points
        cum = torch.empty(CN + 1, dtype=torch.int64, device=dev)
        ws_bytes = int(lib.sc_scan_workspace_bytes(CN))
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        check(lib.sc_isect_count(CN, _p(means2d), _p(radii), None, tile_size, tile_width, tile_height,
                                 _p(tpg), None, None, st), "sc_isect_count")
        check(lib.sc_exclusive_scan_i32(CN, _p(tpg), _p(cum), cum.data_ptr() + 8 * CN, _p(ws), ws_bytes, st),
              "sc_exclusive_scan_i32")
        n_isects = int(cum[CN].item())
        ids_a = torch.empty(n_isects, dtype=torch.int64, device=dev)
        flat_a = torch.empty(n_isects, dtype=torch.int32, device=dev)
        if n_isects:
            check(lib.sc_isect_emit(N, C, _p(means2d), _p(radii), _p(depths), _p(cum), None, tile_size,
                                    tile_width, tile_height, bits, n_isects, _p(ids_a), _p(flat_a), st),
                  "sc_isect_emit")
    return tpg, ids_a, flat_a


@torch.no_grad()
def isect_offset_encode(isect_ids: Tensor, n_cameras: int, tile_width: int, tile_height: int) -> Tensor:
    """-> offsets[C, tile_height, tile_width] i32 (first sorted index of every tile)."""
    assert isect_ids.dtype == torch.int64
    if not isect_ids.is_cuda:
        raise RuntimeError("isect_ids must be a CUDA tensor: street_crafter_b200 has no CPU path")
    isect_ids = isect_ids.contiguous()
    if isect_ids.data_ptr() % 16:
        isect_ids = isect_ids.clone()  # the kernel reads two keys per 16-byte load
    dev = isect_ids.device
    offsets = torch.empty((n_cameras, tile_height, tile_width), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        check(lib.sc_isect_offset_encode(isect_ids.numel(), _p(isect_ids), n_cameras, tile_width, tile_height,
                                         _tile_n_bits(tile_width, tile_height), _p(offsets), _stream()),
              "sc_isect_offset_encode")
    return offsets


# ------------------------------------------------------------------------------------------
# A7/A8  rasterize_to_pixels
# ------------------------------------------------------------------------------------------
# measurement hook (bench.py): when "buf" holds a 2-element int64 CUDA tensor the forward
# kernel adds its visited / contributing pair counts to it (a separate template instance).
_STATS = {"buf": None}


class _RasterizeToPixels(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means2d, conics, colors, opacities, backgrounds, width, height, tile_size,
                isect_offsets, flatten_ids, absgrad, max_hit):
        means2d_in = means2d
        means2d = _f32c(means2d, "means2d", 8)
        conics = _f32c(conics, "conics", 4)
        colors = _f32c(colors, "colors", 4)
        opacities = _f32c(opacities, "opacities", 4)
        if backgrounds is not None:
            backgrounds = _f32c(backgrounds, "backgrounds", 4)
        C, N = opacities.shape
        CH = colors.shape[-1]
        dev = means2d.device
        isect_offsets = isect_offsets.contiguous()
        flatten_ids = flatten_ids.contiguous()
        assert isect_offsets.dtype == torch.int32 and flatten_ids.dtype == torch.int32
        n_isects = flatten_ids.numel()
        render_colors = torch.empty((C, height, width, CH), dtype=torch.float32, device=dev)
        render_alphas = torch.empty((C, height, width, 1), dtype=torch.float32, device=dev)
        last_ids = torch.empty((C, height, width), dtype=torch.int32, device=dev)
        final_T = torch.empty((C, height, width), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            check(lib.sc_rasterize_fwd(C, N, n_isects, CH, width, height, tile_size, _p(means2d), _p(conics),
                                       _p(colors), _p(opacities), _p(backgrounds), _p(isect_offsets),
                                       _p(flatten_ids), int(max_hit), _p(render_colors), _p(render_alphas),
                                       _p(last_ids), _p(final_T), _p(_STATS["buf"]), _stream()),
                  "sc_rasterize_fwd")
        ctx.save_for_backward(means2d, conics, colors, opacities,
                              backgrounds if backgrounds is not None else torch.empty(0, device=dev),
                              isect_offsets, flatten_ids, last_ids, final_T)
        ctx.set_materialize_grads(False)
        ctx.cfg = (width, height, tile_size, bool(absgrad), backgrounds is not None)
        # the caller's tensor object: `.absgrad` is attached to it in backward
        # (street_gaussian/models/street_gaussian_model.py:505-506 reads it)
        ctx.means2d_ref = means2d_in
        return render_colors, render_alphas

    @staticmethod
    def backward(ctx, v_render_colors, v_render_alphas):
        (means2d, conics, colors, opacities, backgrounds, isect_offsets, flatten_ids, last_ids,
         final_T) = ctx.saved_tensors
        width, height, tile_size, absgrad, has_bg = ctx.cfg
        if not has_bg:
            backgrounds = None
        C, N = opacities.shape
        CH = colors.shape[-1]
        dev = means2d.device
        v_render_colors = _grad_or_zeros(v_render_colors, (C, height, width, CH), dev)
        if v_render_alphas is not None:  # None: the kernel takes a null pointer (no gradient into alpha)
            v_render_alphas = _f32c(v_render_alphas, "grad")
        # one zero-filled arena for all five gradient outputs (a single memset launch)
        n2, n3, nc, n1 = C * N * 2, C * N * 3, C * N * CH, C * N
        sizes = [n2, n2 if absgrad else 0, n3, nc, n1]
        arena = torch.zeros(sum(sizes), dtype=torch.float32, device=dev)
        parts = torch.split(arena, sizes)
        v_means2d = parts[0].view(C, N, 2)
        v_abs = parts[1].view(C, N, 2) if absgrad else None
        v_conics = parts[2].view(C, N, 3)
        v_colors = parts[3].view(C, N, CH)
        v_opacities = parts[4].view(C, N)
        with torch.cuda.device(dev):
            check(lib.sc_rasterize_bwd(C, N, flatten_ids.numel(), CH, width, height, tile_size, _p(means2d),
                                       _p(conics), _p(colors), _p(opacities), _p(backgrounds),
                                       _p(isect_offsets), _p(flatten_ids), _p(final_T), _p(last_ids),
                                       _p(v_render_colors), _p(v_render_alphas), _p(v_means2d), _p(v_abs),
                                       _p(v_conics), _p(v_colors), _p(v_opacities), _stream()),
                  "sc_rasterize_bwd")
        if absgrad:
            ctx.means2d_ref.absgrad = v_abs
        v_bg = None
        if has_bg and ctx.needs_input_grad[4]:
            v_bg = (v_render_colors * final_T[..., None]).sum(dim=(1, 2))
        return (v_means2d, v_conics, v_colors, v_opacities, v_bg, None, None, None, None, None, None, None)


def rasterize_to_pixels(means2d: Tensor, conics: Tensor, colors: Tensor, opacities: Tensor, image_width: int,
                        image_height: int, tile_size: int, isect_offsets: Tensor, flatten_ids: Tensor,
                        backgrounds: Optional[Tensor] = None, masks: Optional[Tensor] = None,
                        packed: bool = False, absgrad: bool = False, max_hit: int = 0,
                        ) -> Tuple[Tensor, Tensor]:
    """-> (render_colors[C,H,W,CH], render_alphas[C,H,W,1]).  `max_hit` (extension, default off)
    caps the number of contributing splats per pixel for the LiDAR point-splat mode."""
    if packed:
        raise NotImplementedError("packed=True is not used by the reference")
    if masks is not None:
        raise NotImplementedError("tile masks are not used by the reference")
    C, N = opacities.shape
    assert means2d.shape == (C, N, 2), means2d.shape
    assert conics.shape == (C, N, 3), conics.shape
    assert colors.shape[:2] == (C, N), colors.shape
    if backgrounds is not None:
        assert backgrounds.shape == (C, colors.shape[-1]), backgrounds.shape
    tile_h, tile_w = isect_offsets.shape[1:]
    assert tile_h * tile_size >= image_height and tile_w * tile_size >= image_width, \
        "tile grid does not cover the image"
    CH = colors.shape[-1]
    if CH > 5:
        # wide feature maps are composited 4 channels at a time (not on the reference's path)
        outs, alphas = [], None
        for s in range(0, CH, 4):
            bg = None if backgrounds is None else backgrounds[:, s:s + 4]
            rc, ra = _RasterizeToPixels.apply(means2d, conics, colors[..., s:s + 4], opacities, bg, image_width,
                                              image_height, tile_size, isect_offsets, flatten_ids,
                                              absgrad, max_hit)
            outs.append(rc)
            alphas = ra if alphas is None else alphas
        return torch.cat(outs, -1), alphas
    return _RasterizeToPixels.apply(means2d, conics, colors, opacities, backgrounds, image_width, image_height,
                                    tile_size, isect_offsets, flatten_ids, absgrad, max_hit)


def _project_begin(means, quats, scales, opacities, viewmats, Ks, width, height, eps2d, near_plane, far_plane,
                   radius_clip, antialiased, tile_size, st, capacity=None, counts_host=None):
    """projection (+ opacity * compensation, result packing) and the first half of the binning for ONE camera"""
    N = means.shape[0]
    dev = means.device
    radii = torch.empty((1, N), dtype=torch.int32, device=dev)
    means2d = torch.empty((1, N, 2), dtype=torch.float32, device=dev)
    depths = torch.empty((1, N), dtype=torch.float32, device=dev)
    conics = torch.empty((1, N, 3), dtype=torch.float32, device=dev)
    comps = torch.empty((1, N), dtype=torch.float32, device=dev) if antialiased else None
    op_c = torch.empty((1, N), dtype=torch.float32, device=dev)
    # the renderer's result packing (row A11), written by the producing kernels
    vis = torch.empty((1, N), dtype=torch.bool, device=dev)
    radii_norm = torch.empty((1, N), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(lib.sc_projection_fwd_opacity(N, 1, _p(means), _p(quats), _p(scales), _p(opacities),
                                            _p(viewmats), _p(Ks), int(width), int(height), float(eps2d),
                                            float(near_plane), float(far_plane), float(radius_clip),
                                            _p(radii), _p(means2d), _p(depths), _p(conics), _p(comps),
                                            _p(op_c), _p(vis), _p(radii_norm), st),
              "sc_projection_fwd_opacity")
        tile_width = math.ceil(width / float(tile_size))
        tile_height = math.ceil(height / float(tile_size))
        tpg = torch.empty((1, N), dtype=torch.int32, device=dev)
        state = _isect_begin(means2d, radii, depths, tile_size, tile_width, tile_height, tpg, st, capacity, counts_host)
    pb = dict(radii=radii, means2d=means2d, depths=depths, conics=conics, comps=comps, op_c=op_c, vis=vis,
              radii_norm=radii_norm, state=state, tile_width=tile_width, tile_height=tile_height,
              key=(N, int(width), int(height), int(tile_size), bool(antialiased)))
    if capacity is not None:
        pb["counts"], pb["capacity"] = state[9], int(capacity)
    return pb


def _project_finish(pb, need_ids):
    dev = pb["radii"].device
    with torch.cuda.device(dev):
        tpg, isect_ids, flatten_ids, isect_offsets = _isect_finish(pb.pop("state"), need_ids)
    if isect_offsets is None:
        isect_offsets = isect_offset_encode(isect_ids, 1, pb["tile_width"], pb["tile_height"])
    if isect_ids is None:   # not materialised (need_ids=False): the compositor only reads flatten_ids
        isect_ids = torch.empty(0, dtype=torch.int64, device=dev)
    pb.update(tpg=tpg, isect_ids=isect_ids, flatten_ids=flatten_ids, isect_offsets=isect_offsets)
    return pb


@torch.no_grad()
def prepare_view_begin(means: Tensor, quats: Tensor, scales: Tensor, opacities: Tensor, viewmats: Tensor, Ks: Tensor,
                       width: int, height: int, eps2d: float = 0.3, near_plane: float = 0.01, far_plane: float = 1e10,
                       radius_clip: float = 0.0, antialiased: bool = True, tile_size: int = 16,
                       shs: Optional[Tensor] = None, campos: Optional[Tensor] = None, sh_degree: Optional[int] = None,
                       capacity: Optional[int] = None, counts_host: Optional[Tensor] = None):
    """First half of `prepare_view`: enqueues the projection, the first half of the binning and the asynchronous
    read-back of its counts on the CURRENT stream and returns without waiting.  With `shs` / `campos` / `sh_degree`
    the SH colours of the view are evaluated here too (they only need the projection's radii and depths), so that
    `fused_view(..., prepared=state)` starts with the compositing.  `prepare_view_finish(state)` waits for the
    counts and enqueues the second half of the binning (on the stream current THEN; use the same one).
    `capacity` (an upper bound of the view's intersection count, e.g. 1.25 x the last step's) selects the SYNC-FREE
    binning: nothing is read back, `prepare_view_finish` does not wait, a list beyond the capacity is DROPPED (the
    view renders empty); the counts
    land in `counts_host` (pinned int64[3]: visible, intersections, items) for `capacity_exceeded`.  Such a pass has
    no host dependency: it can be captured in a CUDA graph (pipeline.GraphedStep)."""
    means, quats, scales = _f32c(means.detach(), "means"), _f32c(quats.detach(), "quats"), _f32c(scales.detach(), "scales")
    opacities = _f32c(opacities.detach(), "opacities", 4)
    st = _stream()
    pb = _project_begin(means, quats, scales, opacities, _f32c(viewmats, "viewmats"), _f32c(Ks, "Ks"), width, height, eps2d,
                        near_plane, far_plane, radius_clip, antialiased, tile_size, st, capacity, counts_host)
    if shs is not None:
        shs = _f32c(shs.detach(), "shs")
        campos = _f32c(campos, "camera centre", 4)
        N, Kc = shs.shape[0], shs.shape[1]
        colors4 = torch.empty((1, N, 4), dtype=torch.float32, device=means.device)
        with torch.cuda.device(means.device):
            check(lib.sc_sh_fwd_view(N, Kc, int(sh_degree), _p(means), _p(campos), _p(shs), _p(pb["radii"]),
                                     _p(pb["depths"]), _p(colors4), st), "sc_sh_fwd_view")
        pb["colors4"] = colors4
        pb["sh_key"] = (Kc, int(sh_degree))
    return pb


@torch.no_grad()
def prepare_view_finish(pb: Dict, need_isect_ids: bool = False) -> Dict:
    return _project_finish(pb, need_isect_ids)


def capacity_exceeded(counts_host: Tensor, capacity: int) -> bool:
    """after the stream has passed a sync-free pass: did the view's intersection list fit its capacity?"""
    return int(counts_host[1]) > int(capacity)


def prepared_tensors(pb: Dict) -> List[Tensor]:
    """the device tensors of a prepared state (for Tensor.record_stream when it crosses streams)"""
    return [v for v in pb.values() if isinstance(v, Tensor) and v.is_cuda]


@torch.no_grad()
def prepare_view(means: Tensor, quats: Tensor, scales: Tensor, opacities: Tensor, viewmats: Tensor, Ks: Tensor,
                 width: int, height: int, eps2d: float = 0.3, near_plane: float = 0.01, far_plane: float = 1e10,
                 radius_clip: float = 0.0, antialiased: bool = True, tile_size: int = 16, need_isect_ids: bool = False,
                 shs: Optional[Tensor] = None, campos: Optional[Tensor] = None, sh_degree: Optional[int] = None):
    """Phase 1 of a camera pass — projection and the complete binning, everything that reads only the GEOMETRIC
    parameters — as a state object for `fused_view(..., prepared=state)`.  A multi-view step can run phase 1 of all
    its views first: the view-sharded trainer does, so that the all-gather of the SH coefficients (81 % of the
    parameter bytes) travels under it (parallel.ShardedAdam).  Same kernels, same results."""
    pb = prepare_view_begin(means, quats, scales, opacities, viewmats, Ks, width, height, eps2d, near_plane, far_plane,
                            radius_clip, antialiased, tile_size, shs, campos, sh_degree)
    return prepare_view_finish(pb, need_isect_ids)


# ------------------------------------------------------------------------------------------
# one camera pass as ONE autograd node (rows A1..A8b with A2/A5-prologue/A6 folded into kernels)
# ------------------------------------------------------------------------------------------
class _FusedView(torch.autograd.Function):
    """projection (+ opacity * compensation) -> binning -> SH (dirs, mask, +0.5, clamp, depth
    channel in-kernel) -> compositing for ONE camera, with a single backward that runs the
    three backward kernels back to back.  Removes the ~20 elementwise / cat / fill launches and
    autograd nodes the op-by-op sequence of the reference renderer needs per camera
    (street_gaussian_renderer.py:233-266); results are the same tensors."""

    @staticmethod
    def forward(ctx, means, quats, scales, opacities, shs, viewmats, Ks, campos, width, height, sh_degree,
                eps2d, near_plane, far_plane, radius_clip, antialiased, tile_size, absgrad, sinks, need_ids=True,
                prepared=None, deferred=None, defer_sh=False):
        means = _f32c(means, "means")
        quats = _f32c(quats, "quats")
        scales = _f32c(scales, "scales")
        opacities = _f32c(opacities, "opacities", 4)
        shs = _f32c(shs, "shs")
        viewmats = _f32c(viewmats, "viewmats")
        Ks = _f32c(Ks, "Ks")
        campos = _f32c(campos, "camera centre", 4)
        N, Kc = shs.shape[0], shs.shape[1]
        dev = means.device
        st = _stream()
        colors4 = None
        if prepared is None or "colors4" not in prepared:
            colors4 = torch.empty((1, N, 4), dtype=torch.float32, device=dev)
        if prepared is None:
            pb = _project_begin(means, quats, scales, opacities, viewmats, Ks, width, height, eps2d, near_plane, far_plane,
                                radius_clip, antialiased, tile_size, st)
            with torch.cuda.device(dev):
                # independent of the binning: runs on the GPU while the host waits for the counts
                check(lib.sc_sh_fwd_view(N, Kc, int(sh_degree), _p(means), _p(campos), _p(shs), _p(pb["radii"]),
                                         _p(pb["depths"]), _p(colors4), st), "sc_sh_fwd_view")
            pb = _project_finish(pb, need_ids)
        else:
            # two-phase schedule (prepare_view): projection + binning already ran — e.g. for every view of the
            # step while the SH coefficients were still arriving from the other ranks
            pb = prepared
            assert pb["key"] == (N, int(width), int(height), int(tile_size), bool(antialiased)), "prepared for another view"
            if colors4 is None:
                assert pb["sh_key"] == (Kc, int(sh_degree)), "prepared with other SH settings"
                colors4 = pb["colors4"]
            else:
                with torch.cuda.device(dev):
                    check(lib.sc_sh_fwd_view(N, Kc, int(sh_degree), _p(means), _p(campos), _p(shs), _p(pb["radii"]),
                                             _p(pb["depths"]), _p(colors4), st), "sc_sh_fwd_view")
        radii, means2d, depths, conics, comps, op_c, vis, radii_norm = (pb[k] for k in (
            "radii", "means2d", "depths", "conics", "comps", "op_c", "vis", "radii_norm"))
        tpg, isect_ids, flatten_ids, isect_offsets = (pb[k] for k in ("tpg", "isect_ids", "flatten_ids", "isect_offsets"))
        with torch.cuda.device(dev):
            render_colors = torch.empty((1, height, width, 4), dtype=torch.float32, device=dev)
            render_alphas = torch.empty((1, height, width, 1), dtype=torch.float32, device=dev)
            last_ids = torch.empty((1, height, width), dtype=torch.int32, device=dev)
            final_T = torch.empty((1, height, width), dtype=torch.float32, device=dev)
            depth_norm = torch.empty((1, height, width, 1), dtype=torch.float32, device=dev)
            if "capacity" in pb:   # sync-free binning: the list length is on the device
                check(lib.sc_rasterize_fwd_depth_dev(1, N, flatten_ids.numel(), pb["counts"].data_ptr() + 8, width, height,
                                                     tile_size, _p(means2d), _p(conics), _p(colors4), _p(op_c),
                                                     _p(isect_offsets), _p(flatten_ids), _p(render_colors),
                                                     _p(render_alphas), _p(depth_norm), _p(last_ids), _p(final_T),
                                                     _p(_STATS["buf"]), st), "sc_rasterize_fwd_depth_dev")
            else:
                check(lib.sc_rasterize_fwd_depth(1, N, flatten_ids.numel(), width, height, tile_size, _p(means2d),
                                                 _p(conics), _p(colors4), _p(op_c), _p(isect_offsets),
                                                 _p(flatten_ids), _p(render_colors), _p(render_alphas),
                                                 _p(depth_norm), _p(last_ids), _p(final_T), _p(_STATS["buf"]), st),
                      "sc_rasterize_fwd_depth")
        ctx.save_for_backward(means, quats, scales, opacities, shs, viewmats, Ks, campos, radii, means2d, conics,
                              colors4, op_c, isect_offsets, flatten_ids, last_ids, final_T, render_alphas,
                              depth_norm)
        ctx.cfg = (int(width), int(height), int(tile_size), float(eps2d), bool(antialiased), bool(absgrad),
                   int(sh_degree))
        ctx.sinks = sinks
        ctx.deferred = deferred
        ctx.defer_sh = bool(defer_sh)
        ctx.means2d_ref = None  # weakref to the returned means2d: receives .grad / .absgrad in backward
        ctx.set_materialize_grads(False)  # no zero-filled "gradients" for the auxiliary outputs
        aux = (radii, means2d, depths, conics, op_c, colors4, tpg, isect_ids, flatten_ids, isect_offsets, vis,
               radii_norm)
        ctx.mark_non_differentiable(*aux)
        if comps is not None:
            ctx.mark_non_differentiable(comps)
        return (render_colors, render_alphas, depth_norm) + aux + ((comps,) if comps is not None else ())

    @staticmethod
    def backward(ctx, v_render_colors, v_render_alphas, v_depth_norm, *_unused):
        (means, quats, scales, opacities, shs, viewmats, Ks, campos, radii, means2d, conics, colors4, op_c,
         isect_offsets, flatten_ids, last_ids, final_T, render_alphas, depth_norm) = ctx.saved_tensors
        width, height, tile_size, eps2d, antialiased, absgrad, sh_degree = ctx.cfg
        N, Kc = shs.shape[0], shs.shape[1]
        dev = means.device
        st = _stream()
        v_render_colors = _grad_or_zeros(v_render_colors, (1, height, width, 4), dev)
        if v_render_alphas is not None:  # None: no gradient into the alpha map, the kernel takes a null pointer
            v_render_alphas = _f32c(v_render_alphas, "grad")
        if v_depth_norm is not None:
            # depth_norm = rc[..., 3] / clamp_min(alpha, 1e-10) (only when a depth loss is attached):
            # d/d rc_d = 1 / ac,  d/d alpha = -depth_norm / ac where alpha >= 1e-10
            ac = render_alphas.clamp(min=1e-10)
            v_render_colors = v_render_colors.clone()
            v_render_colors[..., 3:] += v_depth_norm / ac
            v_a = -v_depth_norm * depth_norm / ac * (render_alphas >= 1e-10)
            v_render_alphas = v_a if v_render_alphas is None else v_render_alphas + v_a
        sizes = [4 * N, 2 * N, 2 * N if absgrad else 0, 3 * N, N]  # colours first: 16-byte aligned rows
        arena = torch.zeros(sum(sizes), dtype=torch.float32, device=dev)  # one memset for all five
        v_colors4, v_means2d, v_abs, v_conics, v_opc = torch.split(arena, sizes)
        sinks = ctx.sinks
        if sinks is not None:
            g_means, g_quats, g_scales, g_op, g_shs = (sinks[k] for k in ("means", "quats", "scales",
                                                                           "opacities", "shs"))
        else:
            g_means, g_quats, g_scales = torch.zeros_like(means), torch.zeros_like(quats), torch.zeros_like(scales)
            g_op, g_shs = torch.zeros_like(opacities), torch.zeros_like(shs)
        with torch.cuda.device(dev):
            check(lib.sc_rasterize_bwd(1, N, flatten_ids.numel(), 4, width, height, tile_size, _p(means2d),
                                       _p(conics), _p(colors4), _p(op_c), None, _p(isect_offsets),
                                       _p(flatten_ids), _p(final_T), _p(last_ids), _p(v_render_colors),
                                       _p(v_render_alphas), _p(v_means2d), _p(v_abs) if absgrad else None,
                                       _p(v_conics), _p(v_colors4), _p(v_opc), st), "sc_rasterize_bwd")
            def sh_bwd(keep=(arena, means, campos, shs, radii, colors4)):
                with torch.cuda.device(dev):
                    check(lib.sc_sh_bwd_view(N, Kc, sh_degree, _p(means), _p(campos), _p(shs), _p(radii), _p(colors4),
                                             _p(v_colors4), _p(g_shs), _p(g_means), 1, _stream()), "sc_sh_bwd_view")

            def projection_bwd(keep=(arena, means, quats, scales, opacities, viewmats, Ks, radii)):
                with torch.cuda.device(dev):
                    check(lib.sc_projection_bwd_opacity(N, 1, _p(means), _p(quats), _p(scales), _p(opacities),
                                                        _p(viewmats), _p(Ks), width, height, eps2d, _p(radii),
                                                        _p(v_means2d), v_colors4.data_ptr() + 12, 4, _p(v_conics),
                                                        _p(v_opc), 1 if antialiased else 0, _p(g_means), _p(g_quats),
                                                        _p(g_scales), _p(g_op), 1, _stream()), "sc_projection_bwd_opacity")

            def tail(keep=(arena, means, quats, scales, opacities, viewmats, Ks, radii, campos, shs, colors4)):
                # may run on ANOTHER stream than this backward (pipeline.ViewPipeline): tell the allocator
                cur = torch.cuda.current_stream(dev)
                for t in keep:
                    t.record_stream(cur)
                sh_bwd()
                projection_bwd()
            deferred_ok = ctx.deferred is not None and sinks is not None
            if deferred_ok and ctx.defer_sh:
                # gradient-arena mode, view pipeline: the two HBM-bound backward kernels run later, on the stream the
                # caller chooses, beside the next view's compositing
                ctx.deferred.append(tail)
            else:
                sh_bwd()
                if deferred_ok:
                    # gradient-arena mode only: the caller runs the projection backward of all its views LATER (so
                    # that the exchange of the SH gradients, final once the last view's SH backward has run, travels
                    # under them)
                    ctx.deferred.append(projection_bwd)
                else:
                    projection_bwd()
        m2 = ctx.means2d_ref() if ctx.means2d_ref is not None else None
        if m2 is not None:
            # what the reference reads after backward (street_gaussian_model.py:504-508)
            m2.grad = v_means2d.view(1, N, 2)
            if absgrad:
                m2.absgrad = v_abs.view(1, N, 2)
        if sinks is not None:
            return (None,) * 23
        need = ctx.needs_input_grad
        return (g_means if need[0] else None, g_quats if need[1] else None, g_scales if need[2] else None,
                g_op if need[3] else None, g_shs if need[4] else None) + (None,) * 18


def fused_view(means: Tensor, quats: Tensor, scales: Tensor, opacities: Tensor, shs: Tensor, viewmats: Tensor,
               Ks: Tensor, campos: Tensor, width: int, height: int, sh_degree: int, eps2d: float = 0.3,
               near_plane: float = 0.01, far_plane: float = 1e10, radius_clip: float = 0.0,
               antialiased: bool = True, tile_size: int = 16, absgrad: bool = False,
               grad_sinks: Optional[Dict[str, Tensor]] = None, need_isect_ids: bool = True,
               prepared: Optional[Dict] = None, deferred: Optional[list] = None,
               defer_sh: bool = False) -> Tuple[Tensor, Tensor, Dict]:
    """One camera (viewmats[1,4,4], Ks[1,3,3], campos[3] on the device), SH colours with K = 4 or
    16 coefficients, RGB + depth channel: -> (render_colors[1,H,W,4], render_alphas[1,H,W,1], meta).
    `meta["means2d"].grad` / `.absgrad` are filled by the backward pass.  `need_isect_ids=False`
    leaves meta["isect_ids"] empty: the 64-bit keys are an intermediate of the binning that only the
    op-by-op API hands out (the reference renderer's result dict does not hold them).  `deferred` (a list, with
    `grad_sinks` only): the backward appends its projection-backward launch to it instead of running it; the caller
    calls the entries once the backward passes of all its views are through (parallel.ShardedAdam).  `defer_sh`: the
    deferred entry also holds the SH backward (pipeline.ViewPipeline runs both on its memory stream)."""
    import weakref
    assert viewmats.shape == (1, 4, 4) and Ks.shape == (1, 3, 3) and campos.shape == (3,)
    assert shs.dim() == 3 and shs.shape[1] in (4, 16) and (sh_degree + 1) ** 2 <= shs.shape[1]
    assert opacities.shape == (means.shape[0],)
    if grad_sinks is not None:
        assert all(k in grad_sinks for k in ("means", "quats", "scales", "opacities", "shs"))
    out = _FusedView.apply(means, quats, scales, opacities, shs, viewmats, Ks, campos, width, height, sh_degree,
                           eps2d, near_plane, far_plane, radius_clip, antialiased, tile_size, absgrad, grad_sinks,
                           need_isect_ids, prepared, deferred, defer_sh)
    (rc, ra, depth_norm, radii, means2d, depths, conics, op_c, colors4, tpg, isect_ids, flatten_ids, isect_offsets,
     vis, radii_norm) = out[:15]
    comps = out[15] if len(out) > 15 else None
    if rc.grad_fn is not None:
        rc.grad_fn.means2d_ref = weakref.ref(means2d)
    meta = {"camera_ids": None, "gaussian_ids": None, "radii": radii, "means2d": means2d, "depths": depths,
            "conics": conics, "opacities": op_c, "compensations": comps, "colors": colors4,
            "tile_width": math.ceil(width / float(tile_size)), "tile_height": math.ceil(height / float(tile_size)),
            "tiles_per_gauss": tpg, "isect_ids": isect_ids, "flatten_ids": flatten_ids,
            "isect_offsets": isect_offsets, "width": width, "height": height, "tile_size": tile_size,
            "n_cameras": 1,
            # result packing of the reference renderer (street_gaussian_renderer.py:283, :293-294),
            # produced by the kernels: depth / clamp(alpha), radii > 0, radii / max(H, W)
            "depth_norm": depth_norm, "visibility": vis, "radii_norm": radii_norm}
    return rc, ra, meta


# ------------------------------------------------------------------------------------------
# rasterization(): the fused high-level entry (imported by the reference, see header)
# ------------------------------------------------------------------------------------------
def rasterization(means: Tensor, quats: Tensor, scales: Tensor, opacities: Tensor, colors: Tensor,
                  viewmats: Tensor, Ks: Tensor, width: int, height: int, near_plane: float = 0.01,
                  far_plane: float = 1e10, radius_clip: float = 0.0, eps2d: float = 0.3,
                  sh_degree: Optional[int] = None, packed: bool = False, tile_size: int = 16,
                  backgrounds: Optional[Tensor] = None, render_mode: str = "RGB", sparse_grad: bool = False,
                  absgrad: bool = False, rasterize_mode: str = "classic", channel_chunk: int = 32,
                  distributed: bool = False, camera_model: str = "pinhole", covars: Optional[Tensor] = None,
                  max_hit: int = 0) -> Tuple[Tensor, Tensor, Dict]:
    """means[N,3] quats[N,4] scales[N,3] opacities[N] colors[N,D]|[N,K,3] (SH when sh_degree is
    given) viewmats[C,4,4] Ks[C,3,3] -> (render_colors[C,H,W,X], render_alphas[C,H,W,1], meta).

    Equivalent to the reference renderer's inlined sequence
    (street_gaussian_renderer.py:219-280) with packed=False; `render_mode` in
    {RGB, D, ED, RGB+D, RGB+ED}; `rasterize_mode` in {classic, antialiased}."""
    assert render_mode in ("RGB", "D", "ED", "RGB+D", "RGB+ED"), render_mode
    assert rasterize_mode in ("classic", "antialiased"), rasterize_mode
    if packed or sparse_grad or distributed or covars is not None:
        raise NotImplementedError("packed / sparse_grad / distributed / covars are not used by the reference")
    N = means.shape[0]
    C = viewmats.shape[0]
    assert opacities.shape == (N,), opacities.shape
    if (C == 1 and sh_degree is not None and colors.dim() == 3 and colors.shape[1] in (4, 16)
            and render_mode in ("RGB+D", "RGB+ED") and backgrounds is None and max_hit == 0
            and camera_model == "pinhole"):
        # the reference renderer's configuration: one autograd node per camera pass
        campos = torch.inverse(viewmats)[0, :3, 3]
        render_colors, render_alphas, meta = fused_view(
            means, quats, scales, opacities, colors, viewmats, Ks, campos, width, height, sh_degree,
            eps2d=eps2d, near_plane=near_plane, far_plane=far_plane, radius_clip=radius_clip,
            antialiased=(rasterize_mode == "antialiased"), tile_size=tile_size, absgrad=absgrad)
        if render_mode == "RGB+ED":
            render_colors = torch.cat([render_colors[..., :-1],
                                       render_colors[..., -1:] / render_alphas.clamp(min=1e-10)], dim=-1)
        return render_colors, render_alphas, meta
    radii, means2d, depths, conics, comps = fully_fused_projection(
        means, None, quats, scales, viewmats, Ks, width, height, eps2d=eps2d, near_plane=near_plane,
        far_plane=far_plane, radius_clip=radius_clip, calc_compensations=(rasterize_mode == "antialiased"),
        camera_model=camera_model)
    op = opacities[None].expand(C, N)
    if comps is not None:
        op = op * comps
    tile_width = math.ceil(width / float(tile_size))
    tile_height = math.ceil(height / float(tile_size))
    tiles_per_gauss, isect_ids, flatten_ids = isect_tiles(means2d, radii, depths, tile_size, tile_width,
                                                          tile_height, n_cameras=C)
    isect_offsets = isect_offset_encode(isect_ids, C, tile_width, tile_height)
    if sh_degree is None:
        cols = colors[None].expand(C, -1, -1) if colors.dim() == 2 else colors
    else:
        campos = torch.inverse(viewmats)[:, :3, 3]
        dirs = means[None, :, :] - campos[:, None, :]
        shs = colors[None].expand(C, -1, -1, -1) if colors.dim() == 3 else colors
        cols = spherical_harmonics(sh_degree, dirs, shs, masks=radii > 0)
        cols = torch.clamp_min(cols + 0.5, 0.0)
    if render_mode in ("RGB+D", "RGB+ED"):
        cols = torch.cat((cols, depths[..., None]), dim=-1)
        if backgrounds is not None:
            backgrounds = torch.cat([backgrounds, torch.zeros(C, 1, device=backgrounds.device)], dim=-1)
    elif render_mode in ("D", "ED"):
        cols = depths[..., None]
        if backgrounds is not None:
            backgrounds = torch.zeros(C, 1, device=backgrounds.device)
    render_colors, render_alphas = rasterize_to_pixels(
        means2d, conics, cols, op, width, height, tile_size, isect_offsets, flatten_ids,
        backgrounds=backgrounds, absgrad=absgrad, max_hit=max_hit)
    if render_mode in ("ED", "RGB+ED"):
        render_colors = torch.cat([render_colors[..., :-1],
                                   render_colors[..., -1:] / render_alphas.clamp(min=1e-10)], dim=-1)
    meta = {"camera_ids": None, "gaussian_ids": None, "radii": radii, "means2d": means2d, "depths": depths,
            "conics": conics, "opacities": op, "tile_width": tile_width, "tile_height": tile_height,
            "tiles_per_gauss": tiles_per_gauss, "isect_ids": isect_ids, "flatten_ids": flatten_ids,
            "isect_offsets": isect_offsets, "width": width, "height": height, "tile_size": tile_size,
            "n_cameras": C}
    return render_colors, render_alphas, meta
