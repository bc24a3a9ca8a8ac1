"""TEST INFRASTRUCTURE (oracle), not product: CPU restatement of the reference's photometric
loss terms, SURVEY.md section 8(f) row N3.

Follows /root/reference/street_gaussian/utils/loss_utils.py:
  l1_loss   :21-38   mean |output - gt| over the masked pixels and all channels
  gaussian / create_window :84-93   11-tap Gaussian, sigma 1.5, normalised, outer product
  ssim / _ssim :96-131  zero the inputs outside the mask, depthwise conv2d with zero padding 5,
                        C1 = 0.01^2, C2 = 0.03^2, mean of the SSIM map over ALL C*H*W entries
and the combination at /root/reference/train.py:183-186:
  loss = (1 - lambda_dssim) * lambda_l1 * L1 + lambda_dssim * (1 - ssim)        (+ LPIPS: a network, out of scope)

PINNED: tests/golden/loss_l1_ssim.npz holds outputs of the reference's own functions (imported
from /root/reference by tools/make_golden_loss.py); tests/test_oracle_cpu.py checks this file
against them."""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

WINDOW = 11
SIGMA = 1.5
C1 = 0.01 ** 2
C2 = 0.03 ** 2


def gaussian_window_1d() -> torch.Tensor:
    g = torch.tensor([math.exp(-(x - WINDOW // 2) ** 2 / float(2 * SIGMA ** 2)) for x in range(WINDOW)])
    return g / g.sum()


def l1_loss(output: torch.Tensor, gt: torch.Tensor, mask: torch.Tensor | None = None) -> torch.Tensor:
    """output, gt [C,H,W]; mask [1,H,W] bool"""
    d = (output - gt).abs().permute(1, 2, 0)
    if mask is not None:
        d = d[mask[0]]
    return d.mean()


def ssim(img1: torch.Tensor, img2: torch.Tensor, mask: torch.Tensor | None = None) -> torch.Tensor:
    """img1, img2 [C,H,W]; mask [1,H,W] bool -> scalar mean SSIM"""
    C = img1.shape[0]
    if mask is not None:
        img1 = torch.where(mask, img1, torch.zeros_like(img1))
        img2 = torch.where(mask, img2, torch.zeros_like(img2))
    g = gaussian_window_1d()  # float32, like the reference's torch.Tensor(...) window (:84-93)
    w = (g[:, None] @ g[None, :]).float()[None, None].expand(C, 1, WINDOW, WINDOW).contiguous().to(img1.dtype)
    conv = lambda t: F.conv2d(t[None], w, padding=WINDOW // 2, groups=C)[0]
    mu1, mu2 = conv(img1), conv(img2)
    s1 = conv(img1 * img1) - mu1 * mu1
    s2 = conv(img2 * img2) - mu2 * mu2
    s12 = conv(img1 * img2) - mu1 * mu2
    m = ((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1 * mu1 + mu2 * mu2 + C1) * (s1 + s2 + C2))
    return m.mean()


def photometric_loss(image, gt, mask=None, lambda_dssim: float = 0.2, lambda_l1: float = 1.0):
    """-> (loss, l1, ssim) as at train.py:183-186 without the LPIPS term"""
    l1 = l1_loss(image, gt, mask)
    s = ssim(image, gt, mask)
    return (1.0 - lambda_dssim) * lambda_l1 * l1 + lambda_dssim * (1.0 - s), l1, s
