"""Generates tests/golden/loss_l1_ssim.npz by IMPORTING the reference's own loss functions
(/root/reference/street_gaussian/utils/loss_utils.py: l1_loss, ssim) in this container.
    python tools/make_golden_loss.py"""
import importlib.util
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("ref_loss_utils", "/root/reference/street_gaussian/utils/loss_utils.py")
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

out = {}
g = torch.Generator().manual_seed(7)
for name, (H, W, use_mask) in {"a": (40, 56, True), "b": (33, 47, False), "c": (16, 16, True)}.items():
    img = torch.rand(3, H, W, generator=g, dtype=torch.float64).requires_grad_(True)
    gt = torch.rand(3, H, W, generator=g, dtype=torch.float64)
    gt = (0.6 * gt + 0.4 * img.detach()).clamp(0, 1)          # correlated, like a render vs its target
    mask = (torch.rand(1, H, W, generator=g) > 0.3) if use_mask else None
    l1 = ref.l1_loss(img, gt, mask)
    s = ref.ssim(img, gt, mask=mask)
    loss = (1.0 - 0.2) * 1.0 * l1 + 0.2 * (1.0 - s)            # train.py:183-186, lambda_dssim 0.2, lambda_l1 1
    loss.backward()
    out[f"{name}_img"] = img.detach().numpy(); out[f"{name}_gt"] = gt.numpy()
    out[f"{name}_mask"] = mask.numpy() if mask is not None else np.zeros((0,), dtype=bool)
    out[f"{name}_l1"] = l1.detach().numpy(); out[f"{name}_ssim"] = s.detach().numpy()
    out[f"{name}_loss"] = loss.detach().numpy(); out[f"{name}_grad"] = img.grad.numpy()
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "loss_l1_ssim.npz"), **out)
print({k: v.shape for k, v in out.items()})
