"""Generates tests/golden/densify_cycle.npz by EXECUTING the reference's own `GaussianModel.densify_and_prune`
(/root/reference/street_gaussian/models/gaussian_model.py:521-551) with everything it calls — `densify_and_clone`
(:493-519), `densify_and_split` (:452-491), `densification_postfix` (:433-450), `cat_optimizer` (:384-408),
`prune_points` / `prune_optimizer` (:363-382, :416-431) and `quaternion_to_matrix`
(/root/reference/street_gaussian/utils/general_utils.py:125-146) — on a real torch.optim.Adam with populated
moments.  The method bodies are cut out of the reference files with `ast` (unmodified, nothing copied into this
repository) and bound to a stand-in object.  The bodies hard-code device="cuda"; there is no GPU here, so they
see a `torch` proxy that drops the `device=` keyword of zeros(), makes cuda.empty_cache() a no-op and records the
standard-normal draw behind `torch.normal(mean, std)` (= randn * std + mean, ATen normal_out_impl) so that the
CUDA path can be fed the same noise.       python tools/make_golden_densify_cycle.py"""
import ast
import os
import types

import numpy as np
import torch
from torch import nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference/street_gaussian/models/gaussian_model.py"
SRC_U = "/root/reference/street_gaussian/utils/general_utils.py"
WANT = ("prune_optimizer", "cat_optimizer", "prune_points", "densification_postfix", "densify_and_split",
        "densify_and_clone", "densify_and_prune")


class TorchProxy:
    """Forwards to torch; only the device-bound calls of the method bodies are intercepted."""

    def __init__(self, gen):
        self.gen, self.noise = gen, []
        self.cuda = types.SimpleNamespace(empty_cache=lambda: None)

    def __getattr__(self, name):
        return getattr(torch, name)

    def zeros(self, *a, device=None, **k):
        return torch.zeros(*a, **k)

    def normal(self, mean, std):
        z = torch.randn(std.shape, generator=self.gen)
        self.noise.append(z)
        return z * std + mean


def load_methods(proxy):
    tree = ast.parse(open(SRC).read())
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "GaussianModel")
    funcs = [n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name in WANT]
    assert len(funcs) == len(WANT)
    tree_u = ast.parse(open(SRC_U).read())
    q2m = [n for n in tree_u.body if isinstance(n, ast.FunctionDef) and n.name == "quaternion_to_matrix"]
    assert len(q2m) == 1
    ns = {"torch": proxy, "nn": nn}
    exec(compile(ast.Module(body=q2m, type_ignores=[]), SRC_U, "exec"), ns)
    exec(compile(ast.Module(body=funcs, type_ignores=[]), SRC, "exec"), ns)
    return ns


SHAPES = {"xyz": (3,), "f_dc": (1, 3), "f_rest": (15, 3), "opacity": (1,), "scaling": (3,), "rotation": (4,),
          "semantic": (2,)}
ATTR = {"xyz": "_xyz", "f_dc": "_features_dc", "f_rest": "_features_rest", "opacity": "_opacity",
        "scaling": "_scaling", "rotation": "_rotation", "semantic": "_semantic"}


def make_case(seed, N, max_grad, min_opacity, extent, max_screen_size, percent_dense=0.01, percent_big_ws=0.1):
    g = torch.Generator().manual_seed(seed)
    proxy = TorchProxy(g)
    ns = load_methods(proxy)
    params = {k: nn.Parameter(torch.randn((N,) + s, generator=g)) for k, s in SHAPES.items()}
    with torch.no_grad():   # log-scales around log(percent_dense * extent) so that clone AND split both fire
        params["scaling"].copy_(np.log(percent_dense * extent) + 0.8 * torch.randn(N, 3, generator=g))
        params["opacity"].copy_(2.5 * torch.randn(N, 1, generator=g) - 1.0)
    opt = torch.optim.Adam([{"params": [p], "lr": 1e-3, "name": k} for k, p in params.items()], lr=0.0, eps=1e-15)
    for _ in range(3):   # populate exp_avg / exp_avg_sq
        for p in params.values():
            p.grad = torch.randn(p.shape, generator=g)
        opt.step()
    denom = torch.randint(0, 4, (N, 1), generator=g).float()          # zeros -> 0/0 = NaN -> 0 and x/0 = inf
    accum = torch.rand(N, 2, generator=g) * 2.5 * max_grad * denom.clamp(min=1)
    accum[denom[:, 0] == 0, :] *= (torch.rand(int((denom[:, 0] == 0).sum()), 1, generator=g) < 0.5)

    class Model:
        get_xyz = property(lambda s: s._xyz)
        get_scaling = property(lambda s: torch.exp(s._scaling))              # gaussian_model.py:215, :226-227
        get_opacity = property(lambda s: torch.sigmoid(s._opacity))          # :220, :251-252
    for name in WANT:
        setattr(Model, name, ns[name])
    m = Model()
    m.optimizer, m.scaling_inverse_activation = opt, torch.log               # :216
    m.percent_dense, m.percent_big_ws = percent_dense, percent_big_ws
    m.scalar_dict, m.tensor_dict = {}, {}
    for k, a in ATTR.items():
        setattr(m, a, params[k])
    m.xyz_gradient_accum, m.denom, m.max_radii2D = accum, denom, torch.rand(N, generator=g) * 40
    out = {"cfg": np.array([max_grad, min_opacity, extent, float(max_screen_size or 0), percent_dense,
                            percent_big_ws], dtype=np.float64)}
    for k, p in params.items():
        st = opt.state[p]
        out[f"in_{k}"], out[f"in_m_{k}"], out[f"in_v_{k}"] = (p.detach().numpy().copy(), st["exp_avg"].numpy().copy(),
                                                              st["exp_avg_sq"].numpy().copy())
    for k in ("xyz_gradient_accum", "denom", "max_radii2D"):
        out[f"in_{k}"] = getattr(m, k).numpy().copy()
    with torch.no_grad():
        scalars, _ = m.densify_and_prune(max_grad, min_opacity, True, extent, max_screen_size)
    assert len(proxy.noise) == 1
    out["noise"] = proxy.noise[0].numpy()
    for k, a in ATTR.items():
        p = getattr(m, a)
        st = opt.state[p]
        out[f"out_{k}"], out[f"out_m_{k}"], out[f"out_v_{k}"] = p.detach().numpy(), st["exp_avg"].numpy(), st["exp_avg_sq"].numpy()
    for k in ("xyz_gradient_accum", "denom", "max_radii2D"):
        out[f"out_{k}"] = getattr(m, k).numpy()
    out["scalars"] = np.array([scalars[k] for k in ("points_total", "points_clone", "points_split", "points_pruned")],
                              dtype=np.int64)
    return out


if __name__ == "__main__":
    allv = {}
    for tag, kw in (("a", dict(seed=31, N=400, max_grad=0.0002, min_opacity=0.005, extent=12.0, max_screen_size=20)),
                    ("b", dict(seed=32, N=257, max_grad=0.0006, min_opacity=0.05, extent=3.5, max_screen_size=None))):
        o = make_case(**kw)
        print(tag, "scalars (total, clone, split, pruned)", o["scalars"], "-> rows", o["out_xyz"].shape[0])
        allv.update({f"{tag}_{k}": v for k, v in o.items()})
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "densify_cycle.npz"), **allv)
