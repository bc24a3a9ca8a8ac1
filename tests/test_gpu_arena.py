"""Flat scene arena (-m gpu): fused Adam vs torch.optim.Adam as the reference configures it, checkpoint / PLY
formats against the reference's own GaussianModel method bodies (bytecode table of oracle/build_ref_callers.py),
densify / prune on the arena with live gradient views."""
import os
import types

import numpy as np
import pytest
import torch

from oracle import build_ref_callers as BRC
from oracle.ref_standins import bind

pytestmark = pytest.mark.gpu

LRS = {"xyz": 0.00016 * 5.0, "f_dc": 0.0025, "f_rest": 0.0025 / 20.0, "opacity": 0.05, "scaling": 0.005, "rotation": 0.001,
       "semantic": 0.0}     # street_gaussian/config/config.py:55-62, gaussian_model.py:293-301


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def table():
    t, meta = BRC.load()
    if t is None:
        pytest.skip(meta)
    return t


def _fill_random(arena, seed=0):
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name in arena.specs:
            for a, p in arena.params(name).items():
                p.copy_(torch.randn(p.shape, generator=g).to(p.device) * (0.2 if a != "xyz" else 10.0))


def _torch_adam_like_reference(params):
    groups = [{"params": [params[a]], "lr": LRS[a], "name": a} for a in params]
    return torch.optim.Adam(groups, lr=0.0, eps=1e-15)       # gaussian_model.py:305


def test_fused_adam_matches_torch_adam_and_shards(dev):
    from street_crafter_b200.model_arena import ModelSpec, SceneArena
    specs = {"background": ModelSpec(5003, 3), "obj_000": ModelSpec(257, 3, fourier_dim=5), "sky": ModelSpec(64, 1)}
    arena = SceneArena(specs, dev)
    _fill_random(arena)
    ref = {n: {a: p.detach().clone().requires_grad_(True) for a, p in arena.params(n).items() if p.numel()} for n in specs}
    opts = {n: _torch_adam_like_reference(ref[n]) for n in specs}
    g = torch.Generator().manual_seed(7)
    lrs = {n: LRS for n in specs}
    for step in range(4):
        arena.zero_grad()
        for n in specs:
            for a, p in ref[n].items():
                gr = torch.randn(p.shape, generator=g).to(dev) * (10.0 ** float(torch.randint(-4, 1, (1,), generator=g)))
                gr[torch.rand(p.shape[0], generator=g).to(dev) < 0.3] = 0.0      # invisible Gaussians: zero rows
                p.grad = gr.clone()
                arena.grads(n)[a].copy_(gr)
            opts[n].step()
        if step % 2 == 0:
            arena.adam_step(lrs)
        else:   # the sharded form: every "rank" steps its slice of the flat buffers
            world = 3
            for r in range(world):
                lo, hi = arena.shard_range(r, world)
                arena.adam_step(lrs, lo, hi, advance=(r == 0))
    for n in specs:
        m, v = arena.moments(n)
        for a, p in ref[n].items():
            st = opts[n].state[p]
            assert torch.allclose(arena.params(n)[a], p, rtol=1e-5, atol=1e-6), (n, a, (arena.params(n)[a] - p).abs().max())
            # (torch's lerp / addcmul kernels contract their multiply-adds differently: last-bit differences)
            assert torch.allclose(m[a], st["exp_avg"], rtol=1e-5, atol=1e-9), (n, a, (m[a] - st["exp_avg"]).abs().max())
            assert torch.allclose(v[a], st["exp_avg_sq"], rtol=1e-5, atol=1e-12), (n, a)
        assert arena.step[n] == 4


class _PlyElem:
    """what GaussianModel.load_ply(input_ply=...) reads of a plyfile element: [name] columns and .properties"""

    def __init__(self, arr):
        self.arr = arr
        self.properties = [types.SimpleNamespace(name=f) for f in arr.dtype.names]

    def __getitem__(self, k):
        return self.arr[k]


def _ref_model(table, n, deg, dev, seed=3):
    cfg = types.SimpleNamespace(mode="train", optim=types.SimpleNamespace(
        position_lr_init=0.00016, position_lr_final=0.0000016, position_lr_delay_mult=0.01, position_lr_max_steps=30000,
        feature_lr=0.0025, opacity_lr=0.05, scaling_lr=0.005, rotation_lr=0.001, semantic_lr=0.0, percent_dense=0.01,
        percent_big_ws=0.1))
    ns = bind(table, ["gm_state_dict", "gm_load_state_dict", "gm_training_setup", "gm_make_ply", "gm_load_ply",
                      "gm_construct_list_of_attributes"],
              {"torch": torch, "nn": torch.nn, "np": np, "cfg": cfg, "get_expon_lr_func": lambda **k: (lambda it: 0.0),
               "PlyData": None})

    class M:
        max_sh_degree, spatial_lr_scale, model_name, num_classes = deg, 5.0, "background", 0
        get_xyz = property(lambda s: s._xyz)
        state_dict = ns["gm_state_dict"]
        load_state_dict = ns["gm_load_state_dict"]
        training_setup = ns["gm_training_setup"]
        make_ply = ns["gm_make_ply"]
        load_ply = ns["gm_load_ply"]
        construct_list_of_attributes = ns["gm_construct_list_of_attributes"]
    g = torch.Generator().manual_seed(seed)
    K = (deg + 1) ** 2
    m = M()
    mk = lambda *s: torch.nn.Parameter(torch.randn(*s, generator=g).to(dev))
    m._xyz, m._features_dc, m._features_rest = mk(n, 3), mk(n, 1, 3), mk(n, K - 1, 3)
    m._scaling, m._rotation, m._opacity, m._semantic = mk(n, 3), mk(n, 4), mk(n, 1), mk(n, 0)
    m.max_radii2D = torch.rand(n, generator=g).to(dev)
    return m, M, g


def test_checkpoint_and_ply_vs_reference_methods(table, dev, tmp_path):
    """GaussianModel.state_dict / load_state_dict / make_ply / load_ply (UNMODIFIED bodies) <-> SceneArena"""
    from street_crafter_b200.model_arena import ModelSpec, SceneArena, read_ply
    n, deg = 3001, 3
    m, M, g = _ref_model(table, n, deg, dev)
    m.training_setup()
    m.xyz_gradient_accum += torch.rand(n, 2, generator=g).to(dev)
    m.denom += 2.0
    for _ in range(3):
        for grp in m.optimizer.param_groups:
            p = grp["params"][0]
            p.grad = torch.randn(p.shape, generator=g).to(dev) * 0.01
        m.optimizer.step()
    sd_ref = m.state_dict(is_final=False)
    arena = SceneArena({"background": ModelSpec(1, deg)}, dev)
    arena.load_state_dict("background", sd_ref)
    assert arena.specs["background"].n == n and arena.step["background"] == 3
    lrs = {grp["name"]: grp["lr"] for grp in m.optimizer.param_groups}
    sd = arena.state_dict("background", is_final=False, lrs=lrs, spatial_lr_scale=5.0, active_sh_degree=0)
    assert list(sd) == list(sd_ref), (list(sd), list(sd_ref))
    for k in sd_ref:
        if k == "optimizer":
            a, b = sd[k], sd_ref[k]
            assert a["param_groups"] == b["param_groups"]
            assert list(a["state"]) == list(b["state"])
            for i in b["state"]:
                for kk in ("step", "exp_avg", "exp_avg_sq"):
                    assert torch.equal(a["state"][i][kk].cpu(), b["state"][i][kk].cpu()), (i, kk)
        elif torch.is_tensor(sd_ref[k]):
            assert torch.equal(sd[k].detach(), sd_ref[k].detach()), k
        else:
            assert sd[k] == sd_ref[k], k
    assert list(arena.state_dict("background", is_final=True)) == list(m.state_dict(is_final=True))
    # the reference loads what the arena wrote (load_state_dict in train mode: training_setup + optimizer.load_state_dict)
    m2, _, _ = _ref_model(table, 5, deg, dev, seed=9)
    m2.load_state_dict(sd)
    assert torch.equal(m2._xyz, m._xyz) and torch.equal(m2.denom, m.denom)
    # one more optimizer step on both sides
    arena.zero_grad()
    for grp in m.optimizer.param_groups:
        p = grp["params"][0]
        p.grad = torch.randn(p.shape, generator=g).to(dev) * 0.01
        arena.grads("background")[grp["name"]].copy_(p.grad)
    m.optimizer.step()
    arena.adam_step({"background": lrs})
    for grp in m.optimizer.param_groups:
        a = arena.params("background")[grp["name"]]
        assert torch.allclose(a, grp["params"][0], rtol=1e-5, atol=1e-6), grp["name"]
    # ---- PLY rows (on bit-identical parameters: the two optimizers above differ in the last bits)
    with torch.no_grad():
        for k, t in arena.model_tensors("background").items():
            t.copy_(getattr(m, k))
    el_ref = m.make_ply()
    el = arena.make_ply("background")
    assert el.dtype == el_ref.dtype and el.shape == el_ref.shape
    assert el.tobytes() == el_ref.tobytes(), "PLY rows differ from GaussianModel.make_ply"
    path = str(tmp_path / "point_cloud.ply")
    arena.save_ply(path)
    elems = read_ply(path)
    assert [e[0] for e in elems] == ["vertex_background"] and elems[0][1].tobytes() == el_ref.tobytes()
    m3, _, _ = _ref_model(table, 7, deg, dev, seed=11)
    m3.load_ply(path=None, input_ply=_PlyElem(elems[0][1]))            # the reference's own parser of the element
    for k in ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation"):
        assert torch.equal(getattr(m3, k).detach(), getattr(m, k).detach()), k
    arena2 = SceneArena({"background": ModelSpec(2, deg), "sky": ModelSpec(3, deg)}, dev)
    assert arena2.load_ply(path) == ["background"]
    for k, t in arena2.model_tensors("background").items():
        assert torch.equal(t.detach(), getattr(m, k).detach()), k
    assert arena2.specs["sky"].n == 3


def test_densify_and_prune_on_the_arena_keeps_gradient_views(dev):
    """render -> backward into the arena -> statistics -> densify ON the arena -> render again: the parameter views,
    their .grad links and the other model's slice survive the re-allocation; results equal the dict-based cycle"""
    from street_crafter_b200.densify import densification_stats, densify_and_prune
    from street_crafter_b200.model_arena import ModelSpec, SceneArena
    from street_crafter_b200.renderer import render_kernel_gsplat
    from street_crafter_b200.scene_graph import assemble_features, assemble_scaling_opacity, assemble_xyz_rotation
    from tools.scene import make_camera, make_gaussians
    n, n_sky, W, H, deg = 20000, 500, 320, 208, 3
    s = make_gaussians(n + n_sky, sh_degree=deg, seed=3)
    arena = SceneArena({"background": ModelSpec(n, deg), "sky": ModelSpec(n_sky, deg)}, dev)
    with torch.no_grad():
        for name, sl in (("background", slice(0, n)), ("sky", slice(n, n + n_sky))):
            p = arena.params(name)
            p["xyz"].copy_(s["means"][sl]); p["rotation"].copy_(s["quats"][sl])
            p["scaling"].copy_(torch.log(s["scales"][sl])); p["opacity"].copy_(torch.logit(s["opacities"][sl])[:, None])
            p["f_dc"].copy_(s["shs"][sl, :1]); p["f_rest"].copy_(s["shs"][sl, 1:])
    V, K = make_camera(W, H)
    cc = (-(V[:3, :3].T @ V[:3, 3])).to(dev)

    def render():
        segs = [dict(arena.params(k)) for k in arena.specs]
        segs = [{"xyz": p["xyz"], "rotation": p["rotation"], "scaling": p["scaling"], "opacity": p["opacity"],
                 "f_dc": p["f_dc"], "f_rest": p["f_rest"]} for p in segs]
        xyz, rot = assemble_xyz_rotation(segs)
        sca, opa = assemble_scaling_opacity(segs)
        fea = assemble_features(segs)
        return render_kernel_gsplat(xyz, fea, sca, rot, opa, V[None].to(dev), K[None].to(dev), cc, W, H, deg)

    arena.zero_grad()
    out = render()
    (out["rgb"].square().sum() + out["acc"].sum()).backward()
    gb = arena.buffers["grads"]
    assert float(gb.abs().sum()) > 0 and arena.params("background")["xyz"].grad.data_ptr() == arena.grads("background")["xyz"].data_ptr()
    vp = out["viewspace_points"]
    for name, (a, b) in (("background", (0, n)), ("sky", (n, n + n_sky))):
        st = arena.stats[name]
        densification_stats(vp, out["visibility_filter"], out["radii"], W, H, st["xyz_gradient_accum"], st["denom"],
                            st["max_radii2D"], a, b)
    # the dict-based cycle on copies = the reference-pinned path (tests/golden/densify_cycle.npz)
    p0 = {a: t.detach().clone() for a, t in arena.params("background").items() if t.numel()}
    m0, v0 = ({a: t.clone() for a, t in d.items() if a in p0} for d in arena.moments("background"))
    st0 = {k: t.clone() for k, t in arena.stats["background"].items()}
    sky_before = {a: t.detach().clone() for a, t in arena.params("sky").items()}
    torch.manual_seed(5)
    want_p, want_m, want_v, _, want_scal = densify_and_prune(p0, m0, v0, st0["xyz_gradient_accum"], st0["denom"], 1e-7, 0.005,
                                                              30.0, None)
    torch.manual_seed(5)
    scal = arena.densify_and_prune("background", 1e-7, 0.005, 30.0, None)
    assert scal == want_scal and scal["points_split"] + scal["points_clone"] > 0
    n2 = arena.specs["background"].n
    assert n2 == want_p["xyz"].shape[0] != n
    for a in want_p:
        assert torch.equal(arena.params("background")[a].detach().reshape(want_p[a].shape), want_p[a]), a
        assert torch.equal(arena.moments("background")[0][a].reshape(want_m[a].shape), want_m[a]), a
    for a, t in sky_before.items():
        assert torch.equal(arena.params("sky")[a].detach(), t), a
    assert arena.stats["background"]["denom"].shape == (n2, 1) and float(arena.stats["background"]["denom"].sum()) == 0
    # the views are live: a second render + backward lands in the NEW gradient buffer
    arena.zero_grad()
    out2 = render()
    (out2["rgb"].square().sum()).backward()
    assert arena.buffers["grads"].data_ptr() != gb.data_ptr() and float(arena.buffers["grads"].abs().sum()) > 0
    for name in arena.specs:
        for a, p in arena.params(name).items():
            assert p.grad.data_ptr() == arena.grads(name)[a].data_ptr(), (name, a)
    # prune on the arena
    keep = torch.rand(n2, device=dev) < 0.7
    before = arena.params("background")["xyz"].detach().clone()
    arena.prune("background", keep)
    assert torch.equal(arena.params("background")["xyz"].detach(), before[keep])
    assert arena.stats["background"]["max_radii2D"].shape == (int(keep.sum()),)
