"""Generates tests/golden/densify_stats.npz by EXECUTING the reference's own
`StreetGaussianModel.set_max_radii2D` / `add_densification_stats`
(/root/reference/street_gaussian/models/street_gaussian_model.py:486-519) in this container.

The module itself cannot be imported here (it pulls the whole training stack), so the two method
definitions are cut out of the reference file with `ast` — unmodified source text, compiled where it
lies, nothing copied into this repository — and called on a stand-in `self` that carries exactly the
attributes they touch.       python tools/make_golden_densify.py"""
import ast
import os
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference/street_gaussian/models/street_gaussian_model.py"
tree = ast.parse(open(SRC).read())
cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "StreetGaussianModel")
funcs = [n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name in ("set_max_radii2D", "add_densification_stats")]
assert len(funcs) == 2
mod = ast.Module(body=funcs, type_ignores=[])
ns = {"torch": torch, "GaussianModel": object}
exec(compile(mod, SRC, "exec"), ns)

g = torch.Generator().manual_seed(11)
N, W, H = 700, 1920, 1280
ranges = {"background": (0, 400), "obj_001": (400, 700)}
self_ = types.SimpleNamespace(graph_gaussian_range=ranges)
init = {}
for name, (a, b) in ranges.items():
    m = types.SimpleNamespace(max_radii2D=torch.rand(b - a, generator=g, dtype=torch.float64) * 0.01,
                              xyz_gradient_accum=torch.rand(b - a, 2, generator=g, dtype=torch.float64),
                              denom=torch.randint(0, 5, (b - a, 1), generator=g).double())
    setattr(self_, name, m)
    init[name] = (m.max_radii2D.clone(), m.xyz_gradient_accum.clone(), m.denom.clone())
vis = torch.rand(N, generator=g) < 0.4
radii = (torch.randint(0, 60, (N,), generator=g).double() / 1920.0) * vis
vp = torch.zeros(1, N, 2, dtype=torch.float64)
vp.grad = torch.randn(1, N, 2, generator=g, dtype=torch.float64) * 1e-4
vp.absgrad = torch.rand(1, N, 2, generator=g, dtype=torch.float64) * 1e-3
cam = types.SimpleNamespace(image_width=W, image_height=H)
ns["set_max_radii2D"](self_, radii, vis)
ns["add_densification_stats"](self_, vp, vis, cam)
out = {"vis": vis.numpy(), "radii": radii.numpy(), "grad": vp.grad.numpy(), "absgrad": vp.absgrad.numpy(),
       "W": np.int64(W), "H": np.int64(H)}
cat = lambda k: torch.cat([getattr(getattr(self_, n), k) for n in ranges]).numpy()
out["max_radii2D_out"], out["accum_out"], out["denom_out"] = cat("max_radii2D"), cat("xyz_gradient_accum"), cat("denom")
out["max_radii2D_in"] = torch.cat([init[n][0] for n in ranges]).numpy()
out["accum_in"] = torch.cat([init[n][1] for n in ranges]).numpy()
out["denom_in"] = torch.cat([init[n][2] for n in ranges]).numpy()
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "densify_stats.npz"), **out)
print({k: getattr(v, "shape", v) for k, v in out.items()})
