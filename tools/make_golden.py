"""Generate tests/golden/*.npz from the REFERENCE's own importable code (run in the build
container only; /root/reference does not exist on the GPU box).

  python tools/make_golden.py

sh_eval.npz : street_gaussian/utils/sh_utils.py:57-112 `eval_sh` on seeded inputs, deg 0..3
              (+ deg 4), coefficient layout transposed to gsplat's [N,K,3].
"""
import importlib.util
import os
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location(
    "ref_sh_utils", "/root/reference/street_gaussian/utils/sh_utils.py")
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

g = torch.Generator().manual_seed(0)
N = 257
dirs = torch.randn(N, 3, generator=g) * 3.0          # un-normalised, as the renderer passes them
coeffs = torch.randn(N, 25, 3, generator=g)          # [N,K,3] gsplat layout
unit = dirs / dirs.norm(dim=-1, keepdim=True)
out = {"dirs": dirs.numpy(), "coeffs": coeffs.numpy()}
for deg in range(5):
    K = (deg + 1) ** 2
    sh = coeffs[:, :K, :].permute(0, 2, 1)            # reference layout [N,C,K]
    out[f"colors_deg{deg}"] = ref.eval_sh(deg, sh, unit).numpy()
rgb = torch.rand(16, 3, generator=g)
out["rgb"] = rgb.numpy()
out["rgb2sh"] = ref.RGB2SH(rgb).numpy()
np.savez(os.path.join(ROOT, "tests", "golden", "sh_eval.npz"), **out)
print("wrote tests/golden/sh_eval.npz", {k: v.shape for k, v in out.items()})
