"""Fused photometric loss: the reference's masked L1 + SSIM terms
(/root/reference/street_gaussian/utils/loss_utils.py:21-38 `l1_loss`, :96-131 `ssim`) and their
combination at /root/reference/train.py:183-186

    loss = (1 - lambda_dssim) * lambda_l1 * Ll1 + lambda_dssim * (1 - ssim_value)        (no LPIPS)

as ONE forward and ONE backward kernel (csrc/loss.cu) that read the compositor's output where it
lies and write the gradient in the layout the rasterizer's backward consumes.  `l1_loss` and
`ssim` keep the reference's names and argument order.  No CPU path."""
from __future__ import annotations

from typing import Optional, Tuple

import torch
from torch import Tensor

from ._lib import check, lib


def _strides_chw(image: Tensor) -> Optional[Tuple[int, int]]:
    """(channel_stride, pixel_stride) if element (c,y,x) lives at c*cs + (y*W + x)*ps, else None"""
    C, H, W = image.shape
    sc, sy, sx = image.stride()
    if sx > 0 and sc > 0 and (H == 1 or sy == W * sx):
        return sc, sx
    return None


class _Photometric(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, gt, mask, lambda_dssim, lambda_l1, hwc=False):
        if not image.is_cuda:
            raise RuntimeError("image must be a CUDA tensor: street_crafter_b200 has no CPU path")
        assert image.dtype == torch.float32
        ctx.base_shape = None
        if hwc:
            # the compositor's own output [1,H,W,CH] / [H,W,CH] (CH >= 3): the autograd input is that
            # tensor itself, so the gradient goes back without view-backward nodes
            ctx.base_shape = tuple(image.shape)
            image = image.reshape(image.shape[-3], image.shape[-2], image.shape[-1]).permute(2, 0, 1)[:3]
        assert image.dim() == 3 and image.shape[0] == 3, image.shape
        C, H, W = image.shape
        st = _strides_chw(image)
        if st is None:
            image = image.contiguous()
            st = (H * W, 1)
        gt = gt.to(torch.float32).contiguous()
        assert gt.shape == (3, H, W), gt.shape
        m8 = None
        if mask is not None:
            assert mask.numel() == H * W, mask.shape
            m8 = mask.reshape(H, W).contiguous()
            m8 = m8.view(torch.uint8) if m8.dtype == torch.bool else (m8 != 0).view(torch.uint8)
        dev = image.device
        ws_bytes = int(lib.sc_photometric_workspace_bytes(H, W))
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        out = torch.empty(3, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            check(lib.sc_photometric_fwd(H, W, image.data_ptr(), st[0], st[1], gt.data_ptr(),
                                         None if m8 is None else m8.data_ptr(), float(lambda_dssim),
                                         float(lambda_l1), ws.data_ptr(), ws_bytes, out.data_ptr(),
                                         torch.cuda.current_stream().cuda_stream), "sc_photometric_fwd")
        ctx.save_for_backward(image, gt, m8 if m8 is not None else torch.empty(0, device=dev), ws)
        ctx.cfg = (H, W, st, m8 is not None, float(lambda_dssim), float(lambda_l1))
        ctx.set_materialize_grads(False)
        loss, l1, ssim_v = out[0], out[1], out[2]
        ctx.mark_non_differentiable(l1, ssim_v)
        return loss, l1, ssim_v

    @staticmethod
    def backward(ctx, v_loss, _v_l1, _v_ssim):
        image, gt, m8, ws = ctx.saved_tensors
        H, W, st, has_mask, lambda_dssim, lambda_l1 = ctx.cfg
        if v_loss is None:
            return None, None, None, None, None, None
        dev = image.device
        up = v_loss.to(torch.float32).contiguous()
        # gradient in the layout of `image`: a strided view of the compositor's [H,W,CH] output
        # gets a [H,W,CH] buffer (extra channels zero), a contiguous image a contiguous one
        zero_extra = 0
        if st[1] > 1:
            # channels beyond RGB (the depth channel) are zeroed by the kernel, not by a fill
            interleaved = st[0] == 1 and st[1] - 3 <= 8
            g_full = (torch.empty if interleaved else torch.zeros)((H, W, st[1]), dtype=torch.float32, device=dev)
            g = g_full.permute(2, 0, 1)[:3] if st[0] == 1 else None
            zero_extra = st[1] - 3 if interleaved else 0
        else:
            g_full = torch.empty((3, H, W), dtype=torch.float32, device=dev)
            g = g_full if st[0] == H * W else None
        if g is None:  # unusual strides: dense [3,H,W] gradient
            g_full = torch.empty((3, H, W), dtype=torch.float32, device=dev)
            g = g_full
            zero_extra = 0
        vs = _strides_chw(g)
        with torch.cuda.device(dev):
            check(lib.sc_photometric_bwd(H, W, image.data_ptr(), st[0], st[1], gt.data_ptr(),
                                         m8.data_ptr() if has_mask else None, lambda_dssim, lambda_l1,
                                         ws.data_ptr(), up.data_ptr(), g.data_ptr(), vs[0], vs[1], zero_extra,
                                         torch.cuda.current_stream().cuda_stream), "sc_photometric_bwd")
        if ctx.base_shape is not None:
            assert g_full.dim() == 3 and g_full.shape[-1] == ctx.base_shape[-1]
            return g_full.view(ctx.base_shape), None, None, None, None, None
        return g, None, None, None, None, None


def photometric_loss_render(render_colors: Tensor, gt: Tensor, mask: Optional[Tensor] = None,
                            lambda_dssim: float = 0.2, lambda_l1: float = 1.0) -> Tuple[Tensor, Tensor, Tensor]:
    """Same loss on the compositor's own output `render_colors` [1,H,W,CH] or [H,W,CH] (RGB in the
    first three channels; further channels, e.g. depth, get zero gradient): the gradient is
    produced directly in that layout."""
    assert render_colors.shape[-1] >= 3 and render_colors.dim() in (3, 4)
    if render_colors.dim() == 4:
        assert render_colors.shape[0] == 1
    if not render_colors.is_contiguous():
        render_colors = render_colors.contiguous()
    return _Photometric.apply(render_colors, gt, mask, lambda_dssim, lambda_l1, True)


def photometric_loss(image: Tensor, gt: Tensor, mask: Optional[Tensor] = None, lambda_dssim: float = 0.2,
                     lambda_l1: float = 1.0) -> Tuple[Tensor, Tensor, Tensor]:
    """image [3,H,W] (any view whose pixels are equally strided, e.g. the `rgb` entry of the
    renderer's result), gt [3,H,W], mask [1,H,W] bool or None -> (loss, Ll1, ssim_value), all
    0-dim device tensors; only `loss` carries gradient (train.py:183-186 without LPIPS)."""
    return _Photometric.apply(image, gt, mask, lambda_dssim, lambda_l1)


def l1_loss(network_output: Tensor, gt: Tensor, mask: Optional[Tensor] = None) -> Tensor:
    """loss_utils.py:21-38"""
    return _Photometric.apply(network_output, gt, mask, 0.0, 1.0)[0]


def ssim(img1: Tensor, img2: Tensor, window_size: int = 11, size_average: bool = True,
         mask: Optional[Tensor] = None) -> Tensor:
    """loss_utils.py:96-131 (window 11, size_average=True: the only configuration train.py uses)"""
    if window_size != 11 or not size_average:
        raise NotImplementedError("the reference only calls ssim(window_size=11, size_average=True)")
    return 1.0 - _Photometric.apply(img1, img2, mask, 1.0, 1.0)[0]
