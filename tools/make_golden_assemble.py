"""Generates tests/golden/assemble.npz by EXECUTING the reference's own `StreetGaussianModel.get_xyz` /
`get_rotation` / `get_scaling` / `get_opacity` / `get_features` property bodies
(/root/reference/street_gaussian/models/street_gaussian_model.py:273-407; the sky model's clamped `get_scaling`,
gaussian_model_sky.py:63-66; the actors' `get_features_fourier`, gaussian_model_actor.py:67-76, with the reference's
`IDFT`, sh_utils.py:120-130) with
`quaternion_to_matrix` / `quaternion_raw_multiply` / `matrix_to_quaternion` and its helpers
(/root/reference/street_gaussian/utils/general_utils.py) — cut out of the reference files with `ast` (unmodified,
nothing copied into this repository) and bound to a stand-in scene graph: background + 3 actors (one without
flips, one empty) + sky, training mode with a 50 % flip mask, tracked poses as autograd leaves.  Outputs AND the
autograd gradients of a fixed random linear functional are stored.       python tools/make_golden_assemble.py"""
import ast
import os
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference/street_gaussian/models/street_gaussian_model.py"
SRC_U = "/root/reference/street_gaussian/utils/general_utils.py"


def load():
    tree_u = ast.parse(open(SRC_U).read())
    want_u = ("quaternion_to_matrix", "quaternion_raw_multiply", "matrix_to_quaternion", "_sqrt_positive_part")
    fu = [n for n in tree_u.body if isinstance(n, ast.FunctionDef) and n.name in want_u]
    assert len(fu) == len(want_u)
    for f in fu:
        f.decorator_list = []        # @torch.jit.script needs the source file; eager execution is the same math
    ns = {"torch": torch, "F": torch.nn.functional, "cfg": types.SimpleNamespace(mode="train")}
    exec(compile(ast.Module(body=fu, type_ignores=[]), SRC_U, "exec"), ns)
    tree = ast.parse(open(SRC).read())
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "StreetGaussianModel")
    fm = [n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name in ("get_xyz", "get_rotation", "get_scaling",
                                                                                "get_opacity", "get_features")]
    assert len(fm) == 5
    for f in fm:
        f.decorator_list = []        # call the property bodies as plain methods
    exec(compile(ast.Module(body=fm, type_ignores=[]), SRC, "exec"), ns)
    # the sky model's own get_scaling (clamped to its sphere radius), gaussian_model_sky.py:63-66
    src_sky = "/root/reference/street_gaussian/models/gaussian_model_sky.py"
    sky = next(n for n in ast.parse(open(src_sky).read()).body if isinstance(n, ast.ClassDef) and n.name == "GaussianModelSky")
    fs = [n for n in sky.body if isinstance(n, ast.FunctionDef) and n.name == "get_scaling"]
    assert len(fs) == 1
    fs[0].decorator_list, fs[0].name = [], "sky_get_scaling"
    exec(compile(ast.Module(body=fs, type_ignores=[]), src_sky, "exec"), ns)
    # ... and its shell-projected get_xyz, gaussian_model_sky.py:68-76
    fx = [n for n in sky.body if isinstance(n, ast.FunctionDef) and n.name == "get_xyz"]
    assert len(fx) == 1
    fx[0].decorator_list, fx[0].name = [], "sky_get_xyz"
    exec(compile(ast.Module(body=fx, type_ignores=[]), src_sky, "exec"), ns)
    # the actor model's Fourier DC features (gaussian_model_actor.py:67-76) and the base model's get_features
    # (gaussian_model.py:238-241); IDFT from the reference's own sh_utils (imports on CPU)
    import importlib.util
    spec = importlib.util.spec_from_file_location("ref_sh_utils", "/root/reference/street_gaussian/utils/sh_utils.py")
    shu = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(shu)
    ns["IDFT"] = shu.IDFT
    for path, cname, fname, alias in (("/root/reference/street_gaussian/models/gaussian_model_actor.py", "GaussianModelActor",
                                       "get_features_fourier", "actor_get_features_fourier"),
                                      ("/root/reference/street_gaussian/models/gaussian_model.py", "GaussianModel",
                                       "get_features", "base_get_features")):
        c = next(n for n in ast.parse(open(path).read()).body if isinstance(n, ast.ClassDef) and n.name == cname)
        f = [n for n in c.body if isinstance(n, ast.FunctionDef) and n.name == fname]
        assert len(f) == 1
        f[0].decorator_list, f[0].name = [], alias
        exec(compile(ast.Module(body=f, type_ignores=[]), path, "exec"), ns)
    return ns


class Leaf:
    """One model of the scene graph: raw parameters, the reference's getters (gaussian_model.py:229-235)."""

    def __init__(self, n, g):
        self._xyz = (torch.randn(n, 3, generator=g) * 3).requires_grad_(True)
        self._rotation = (torch.randn(n, 4, generator=g) * 1.7).requires_grad_(True)
        self._scaling = (torch.randn(n, 3, generator=g) - 1.0).requires_grad_(True)
        self._opacity = (torch.randn(n, 1, generator=g) * 2.5).requires_grad_(True)
        self._features_dc = torch.randn(n, self.fourier_dim, 3, generator=g).requires_grad_(True)
        self._features_rest = (torch.randn(n, 15, 3, generator=g) * 0.1).requires_grad_(True)

    fourier_dim = 1
    get_xyz = property(lambda s: s._xyz)
    get_rotation = property(lambda s: torch.nn.functional.normalize(s._rotation))
    get_scaling = property(lambda s: torch.exp(s._scaling))                  # gaussian_model.py:215, :226-227
    get_opacity = property(lambda s: torch.sigmoid(s._opacity))              # :220, :251-252


if __name__ == "__main__":
    ns = load()
    g = torch.Generator().manual_seed(41)
    sizes = {"background": 37, "obj_000": 19, "obj_001": 0, "obj_002": 33, "sky": 11}
    class Actor(Leaf):                                                        # gaussian_model_actor.py:67-76
        fourier_dim, fourier_scale, start_frame, end_frame = 5, 1.0, 10, 58

    def actor_features(self, frame=0):
        # the body moves the base to 'cuda'; a tensor subclass that ignores .to() keeps it on the CPU
        ns["IDFT"] = lambda t, d: _NoMove(IDFT_REF(t, d))
        return ns["actor_get_features_fourier"](self, frame)
    IDFT_REF = ns["IDFT"]

    class _NoMove:
        def __init__(self, t): self.t = t
        def __getitem__(self, i): return _NoMove(self.t[i])
        def to(self, *a, **k): return self.t
    Actor.get_features_fourier = actor_features
    Leaf.get_features = property(ns["base_get_features"])
    m = types.SimpleNamespace(**{k: (Actor if k.startswith("obj_") else Leaf)(n, g) for k, n in sizes.items()})
    m.frame = 31
    m.graph_obj_list = ["obj_000", "obj_001", "obj_002"]
    m.get_visibility = lambda name: True
    m.flip_axis = 1                                                           # street_gaussian_model.py:55-58
    fm = torch.eye(3) * -1
    fm[m.flip_axis, m.flip_axis] = 1
    m.flip_matrix = ns["matrix_to_quaternion"](fm.unsqueeze(0))
    pose_rot = {k: (torch.randn(4, generator=g)).requires_grad_(True) for k in m.graph_obj_list}   # RAW (not unit)
    pose_trans = {k: (torch.randn(3, generator=g) * 10).requires_grad_(True) for k in m.graph_obj_list}
    # parse_camera :245-271
    m.obj_rots = torch.cat([pose_rot[k].expand(sizes[k], -1) for k in m.graph_obj_list], 0)
    m.obj_trans = torch.cat([pose_trans[k].unsqueeze(0).expand(sizes[k], -1) for k in m.graph_obj_list], 0)
    flips = {"obj_000": torch.rand(sizes["obj_000"], generator=g) < 0.5, "obj_001": torch.zeros(0, dtype=torch.bool),
             "obj_002": torch.zeros(sizes["obj_002"], dtype=torch.bool)}
    m.flip_mask = torch.cat([flips[k] for k in m.graph_obj_list], 0)
    class Sky(Leaf):                                                          # gaussian_model_sky.py:63-76
        scaling_activation = staticmethod(torch.exp)
        sphere_radius = torch.Tensor([0.45])                                  # :24, a [1] float tensor
        sphere_center = torch.tensor([0.5, -0.3, 0.2])
        get_scaling = property(ns["sky_get_scaling"])
        get_xyz = property(ns["sky_get_xyz"])
    m.sky._xyz = (m.sky._xyz.detach() * 0.3).requires_grad_(True)            # some inside, some outside 2 R
    m.sky.__class__ = Sky
    xyz = ns["get_xyz"](m)
    rot = ns["get_rotation"](m)
    n = sum(sizes.values())
    assert xyz.shape == (n, 3) and rot.shape == (n, 4)
    w_xyz, w_rot = torch.randn(n, 3, generator=g), torch.randn(n, 4, generator=g)
    sca, opa = ns["get_scaling"](m), ns["get_opacity"](m)
    assert sca.shape == (n, 3) and opa.shape == (n, 1)
    w_sca, w_opa = torch.randn(n, 3, generator=g), torch.randn(n, 1, generator=g)
    fea = ns["get_features"](m)
    assert fea.shape == (n, 16, 3)
    w_fea = torch.randn(n, 16, 3, generator=g)
    ((xyz * w_xyz).sum() + (rot * w_rot).sum() + (sca * w_sca).sum() + (opa * w_opa).sum() + (fea * w_fea).sum()).backward()
    out = {"names": np.array(list(sizes)), "sizes": np.array(list(sizes.values())), "flip_quat": m.flip_matrix[0].numpy(),
           "flip_axis": np.array(m.flip_axis), "out_xyz": xyz.detach().numpy(), "out_rot": rot.detach().numpy(),
           "w_xyz": w_xyz.numpy(), "w_rot": w_rot.numpy(), "out_scaling": sca.detach().numpy(),
           "out_opacity": opa.detach().numpy(), "w_scaling": w_sca.numpy(), "w_opacity": w_opa.numpy(),
           "sky_clamp_max": np.array(float(Sky.sphere_radius)),
           "sky_center": Sky.sphere_center.numpy(), "sky_radius": np.array(float(Sky.sphere_radius), dtype=np.float32), "out_features": fea.detach().numpy(),
           "w_features": w_fea.numpy(),
           "actor_idft_base": IDFT_REF(Actor.fourier_scale * (m.frame - Actor.start_frame) / (Actor.end_frame - Actor.start_frame),
                                       Actor.fourier_dim)[0].numpy()}
    for k in sizes:
        leaf = getattr(m, k)
        out[f"{k}_xyz"], out[f"{k}_rot"] = leaf._xyz.detach().numpy(), leaf._rotation.detach().numpy()
        out[f"{k}_v_xyz"], out[f"{k}_v_rot"] = leaf._xyz.grad.numpy(), leaf._rotation.grad.numpy()
        out[f"{k}_scaling"], out[f"{k}_opacity"] = leaf._scaling.detach().numpy(), leaf._opacity.detach().numpy()
        out[f"{k}_v_scaling"], out[f"{k}_v_opacity"] = leaf._scaling.grad.numpy(), leaf._opacity.grad.numpy()
        out[f"{k}_f_dc"], out[f"{k}_f_rest"] = leaf._features_dc.detach().numpy(), leaf._features_rest.detach().numpy()
        out[f"{k}_v_f_dc"], out[f"{k}_v_f_rest"] = leaf._features_dc.grad.numpy(), leaf._features_rest.grad.numpy()
    for k in m.graph_obj_list:
        out[f"{k}_pose_rot"], out[f"{k}_pose_trans"] = pose_rot[k].detach().numpy(), pose_trans[k].detach().numpy()
        z4, z3 = torch.zeros(4), torch.zeros(3)
        out[f"{k}_v_pose_rot"] = (pose_rot[k].grad if pose_rot[k].grad is not None else z4).numpy()
        out[f"{k}_v_pose_trans"] = (pose_trans[k].grad if pose_trans[k].grad is not None else z3).numpy()
        out[f"{k}_flip"] = flips[k].numpy()
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "assemble.npz"), **out)
    d = (m.sky._xyz.detach() - Sky.sphere_center).norm(dim=1)
    print("sky points inside / outside the 2R shell:", int((d < 0.9).sum()), int((d >= 0.9).sum()))
    print(len(out), "arrays;", n, "Gaussians; flip quaternion", out["flip_quat"], "; flipped", int(m.flip_mask.sum()))
