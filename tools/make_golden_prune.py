"""Generates tests/golden/prune_points.npz by EXECUTING the reference's own `GaussianModel.prune_optimizer` /
`prune_points` (/root/reference/street_gaussian/models/gaussian_model.py:363-382, :416-431) on a real
torch.optim.Adam with populated moments.  The method bodies are cut out of the reference file with `ast`
(unmodified, nothing copied into this repository) and bound to a stand-in object that carries the attributes
they touch.       python tools/make_golden_prune.py"""
import ast
import os
import types

import numpy as np
import torch
from torch import nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference/street_gaussian/models/gaussian_model.py"
tree = ast.parse(open(SRC).read())
cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "GaussianModel")
funcs = [n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name in ("prune_optimizer", "prune_points")]
assert len(funcs) == 2
ns = {"torch": torch, "nn": nn}
exec(compile(ast.Module(body=funcs, type_ignores=[]), SRC, "exec"), ns)

g = torch.Generator().manual_seed(23)
N = 300
shapes = {"xyz": (3,), "f_dc": (1, 3), "f_rest": (15, 3), "opacity": (1,), "scaling": (3,), "rotation": (4,),
          "semantic": (2,)}
params = {k: nn.Parameter(torch.randn((N,) + s, generator=g)) for k, s in shapes.items()}
opt = torch.optim.Adam([{"params": [p], "lr": 1e-3, "name": k} for k, p in params.items()], lr=0.0, eps=1e-15)
for _ in range(3):   # populate exp_avg / exp_avg_sq
    for p in params.values():
        p.grad = torch.randn(p.shape, generator=g)
    opt.step()
self_ = types.SimpleNamespace(optimizer=opt, xyz_gradient_accum=torch.rand(N, 2, generator=g),
                              denom=torch.randint(0, 9, (N, 1), generator=g).float(),
                              max_radii2D=torch.rand(N, generator=g))
self_.prune_optimizer = types.MethodType(ns["prune_optimizer"], self_)
out = {"mask": (torch.rand(N, generator=g) < 0.35).numpy()}   # True = PRUNE (prune_points takes the complement)
for k, p in params.items():
    st = opt.state[p]
    out[f"in_{k}"], out[f"in_m_{k}"], out[f"in_v_{k}"] = p.detach().numpy().copy(), st["exp_avg"].numpy().copy(), st["exp_avg_sq"].numpy().copy()
for k in ("xyz_gradient_accum", "denom", "max_radii2D"):
    out[f"in_{k}"] = getattr(self_, k).numpy().copy()
ns["prune_points"](self_, torch.from_numpy(out["mask"]))
attr = {"xyz": "_xyz", "f_dc": "_features_dc", "f_rest": "_features_rest", "opacity": "_opacity",
        "scaling": "_scaling", "rotation": "_rotation", "semantic": "_semantic"}
for k, a in attr.items():
    p = getattr(self_, a)
    st = opt.state[p]
    out[f"out_{k}"], out[f"out_m_{k}"], out[f"out_v_{k}"] = p.detach().numpy(), st["exp_avg"].numpy(), st["exp_avg_sq"].numpy()
for k in ("xyz_gradient_accum", "denom", "max_radii2D"):
    out[f"out_{k}"] = getattr(self_, k).numpy()
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "prune_points.npz"), **out)
print(len(out), "arrays; kept", int((~out["mask"]).sum()), "of", N)
