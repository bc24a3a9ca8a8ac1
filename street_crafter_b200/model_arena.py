"""The Gaussian models of a scene on ONE flat arena.

The reference keeps every model (background, each actor, sky) as seven `nn.Parameter`s with a
`torch.optim.Adam` of seven groups (/root/reference/street_gaussian/models/gaussian_model.py:56-80, :287-305),
re-allocated at every densification (:363-551).  Here all models share FOUR identically laid-out flat fp32
buffers — parameters, gradients, Adam exp_avg, Adam exp_avg_sq — so that

  * the view-sharded trainer reduces ONE contiguous gradient buffer (all-reduce, or reduce-scatter + sharded
    fused Adam + all-gather), with the raw parameter tensors the reference code sees being views of the buffers;
  * `densify_and_prune` / `prune` move kernels write their outputs straight into a FRESH arena and the views,
    `.grad` links and gradient sinks are re-derived — a multi-GPU run can densify and stay in lock-step;
  * `state_dict` / `load_state_dict` speak the reference's checkpoint keys (gaussian_model.py:158-206) and
    `save_ply` / `load_ply` its PLY rows (gaussian_model.py:82-156, street_gaussian_model.py:88-112).

Layout: model-major, attribute-major inside a model (xyz | f_dc | f_rest | opacity | scaling | rotation |
semantic — the reference's optimizer group order), every slice padded to 4 floats (16-byte aligned views).
No CPU path."""
from __future__ import annotations

import ctypes as C
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.distributed as dist
from torch import Tensor

from ._lib import check, lib
from .compaction import compact_rows
from .densify import densify_and_prune as _densify_and_prune

ATTRS = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation", "semantic")
# checkpoint key of every attribute (GaussianModel.state_dict, gaussian_model.py:183-192)
STATE_KEYS = {"xyz": "xyz", "f_dc": "feature_dc", "f_rest": "feature_rest", "opacity": "opacity", "scaling": "scaling",
              "rotation": "rotation", "semantic": "semantic"}
BUFFERS = ("params", "grads", "exp_avg", "exp_avg_sq")


class ModelSpec:
    """shape of one model: n Gaussians, K SH coefficients per channel, fourier_dim DC coefficients (actors,
    gaussian_model_actor.py:41-46), num_classes semantic columns (0 unless data.use_semantic)"""

    def __init__(self, n: int, sh_degree: int = 3, fourier_dim: int = 1, num_classes: int = 0):
        self.n, self.sh_degree, self.fourier_dim, self.num_classes = int(n), int(sh_degree), int(fourier_dim), int(num_classes)

    @property
    def K(self) -> int:
        return (self.sh_degree + 1) ** 2

    def shapes(self) -> Dict[str, Tuple[int, ...]]:
        n = self.n
        return {"xyz": (n, 3), "f_dc": (n, self.fourier_dim, 3), "f_rest": (n, self.K - 1, 3), "opacity": (n, 1),
                "scaling": (n, 3), "rotation": (n, 4), "semantic": (n, self.num_classes)}

    def with_n(self, n: int) -> "ModelSpec":
        return ModelSpec(n, self.sh_degree, self.fourier_dim, self.num_classes)


def _pad4(x: int) -> int:
    return (x + 3) // 4 * 4


class SceneArena:
    def __init__(self, specs: Dict[str, ModelSpec], device):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("SceneArena lives on a CUDA device: street_crafter_b200 has no CPU path")
        self.specs: Dict[str, ModelSpec] = dict(specs)
        self.step: Dict[str, int] = {k: 0 for k in specs}
        self.stats: Dict[str, Dict[str, Tensor]] = {}
        self._allocate()
        for name, sp in self.specs.items():
            self.stats[name] = self._zero_stats(sp.n)

    # ------------------------------------------------------------------ layout
    def _zero_stats(self, n: int) -> Dict[str, Tensor]:
        z = lambda *s: torch.zeros(s, dtype=torch.float32, device=self.device)
        return {"xyz_gradient_accum": z(n, 2), "denom": z(n, 1), "max_radii2D": z(n)}

    def _allocate(self) -> None:
        self.layout: Dict[str, Dict[str, Tuple[int, int, Tuple[int, ...]]]] = {}
        off = 0
        for name, sp in self.specs.items():
            self.layout[name] = {}
            for a, shape in sp.shapes().items():
                numel = int(np.prod(shape))
                self.layout[name][a] = (off, numel, shape)
                off += _pad4(numel)
        self.total = off
        self.buffers = {b: torch.zeros(max(self.total, 4), dtype=torch.float32, device=self.device) for b in BUFFERS}
        self._views: Dict[str, Dict[str, Dict[str, Tensor]]] = {b: {} for b in BUFFERS}
        for name in self.specs:
            for b in BUFFERS:
                self._views[b][name] = {a: self.buffers[b][o:o + n].view(shape)
                                        for a, (o, n, shape) in self.layout[name].items()}
            for a, p in self._views["params"][name].items():
                p.requires_grad_(True)
                p.grad = self._views["grads"][name][a]

    def span(self, name: str) -> Tuple[int, int]:
        """element range of a model inside the flat buffers"""
        lay = self.layout[name]
        first = lay[ATTRS[0]][0]
        last_o, last_n, _ = lay[ATTRS[-1]]
        return first, last_o + _pad4(last_n)

    # ------------------------------------------------------------------ views
    def params(self, name: str) -> Dict[str, Tensor]:
        """the model's RAW parameters under the reference's optimizer group names; leaf tensors (views of the
        parameter buffer) whose .grad are views of the gradient buffer"""
        return self._views["params"][name]

    def grads(self, name: str) -> Dict[str, Tensor]:
        return self._views["grads"][name]

    def moments(self, name: str) -> Tuple[Dict[str, Tensor], Dict[str, Tensor]]:
        return self._views["exp_avg"][name], self._views["exp_avg_sq"][name]

    def model_tensors(self, name: str) -> Dict[str, Tensor]:
        """the attribute names of the reference's GaussianModel (gaussian_model.py:72-78)"""
        p = self.params(name)
        return {"_xyz": p["xyz"], "_features_dc": p["f_dc"], "_features_rest": p["f_rest"], "_opacity": p["opacity"],
                "_scaling": p["scaling"], "_rotation": p["rotation"], "_semantic": p["semantic"]}

    def zero_grad(self) -> None:
        self.buffers["grads"].zero_()
        for name in self.specs:
            for a, p in self._views["params"][name].items():
                g = self._views["grads"][name][a]
                if p.grad is None or p.grad.data_ptr() != g.data_ptr():
                    p.grad = g

    # ------------------------------------------------------------------ collectives (8(e))
    def allreduce_gradients(self, group=None) -> None:
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(self.buffers["grads"], op=dist.ReduceOp.SUM, group=group)

    def shard_range(self, rank: int, world: int) -> Tuple[int, int]:
        """[lo, hi) of the flat buffers that `rank` owns (multiples of 4 elements, equal sizes: the buffers are
        padded so that reduce_scatter_tensor / all_gather_into_tensor apply)"""
        per = _pad4((self.total + world - 1) // world)
        return min(self.total, rank * per), min(self.total, (rank + 1) * per)

    # ------------------------------------------------------------------ optimiser
    def adam_step(self, lrs: Dict[str, Dict[str, float]], lo: int = 0, hi: Optional[int] = None, betas=(0.9, 0.999),
                  eps: float = 1e-15, advance: bool = True) -> None:
        """One torch.optim.Adam step of every model (gaussian_model.py:293-305: lr per attribute, eps 1e-15) on
        the flat buffers, restricted to elements [lo, hi).  lrs[model][attr]; a missing attr has lr 0."""
        hi = self.total if hi is None else hi
        st = torch.cuda.current_stream(self.device).cuda_stream
        b = self.buffers
        if advance:
            for name in self.specs:
                self.step[name] += 1
        # consecutive models (layout order) with the same step count share bias corrections: up to 9 models
        # (63 groups) per launch
        runs: List[Tuple[int, List[str]]] = []
        for name in self.specs:
            if runs and runs[-1][0] == self.step[name] and len(runs[-1][1]) < 9:
                runs[-1][1].append(name)
            else:
                runs.append((self.step[name], [name]))
        with torch.cuda.device(self.device):
            for step, chunk in runs:
                starts, lr = [], []
                for name in chunk:
                    for a in ATTRS:
                        starts.append(self.layout[name][a][0])
                        lr.append(float(lrs.get(name, {}).get(a, 0.0)))
                starts.append(self.span(chunk[-1])[1])
                clo, chi = max(lo, starts[0]), min(hi, starts[-1])
                if chi <= clo:
                    continue
                n = len(lr)
                check(lib.sc_adam_step(n, (C.c_int64 * (n + 1))(*starts), (C.c_float * n)(*lr), clo, chi,
                                       b["params"].data_ptr(), b["grads"].data_ptr(), b["exp_avg"].data_ptr(),
                                       b["exp_avg_sq"].data_ptr(), float(betas[0]), float(betas[1]), float(eps),
                                       int(step), st), "sc_adam_step")

    # ------------------------------------------------------------------ densification on the arena (N4)
    def _begin_rebuild(self, name: str, new_n: int) -> None:
        """fresh buffers for `name` having `new_n` Gaussians; the other models' slices are copied over (one copy
        per model and buffer); views and .grad links re-derived.  The caller fills the resized model's views."""
        old_buffers = self.buffers
        old_span = {k: self.span(k) for k in self.specs}
        self.specs[name] = self.specs[name].with_n(new_n)
        self._allocate()
        for k in self.specs:
            if k == name:
                continue
            (o0, o1), (n0, n1) = old_span[k], self.span(k)
            for b in ("params", "exp_avg", "exp_avg_sq"):
                self.buffers[b][n0:n1].copy_(old_buffers[b][o0:o1])

    def rebuild(self, name: str, new_n: int, fill: Callable[[Callable[[str, str], Tensor]], None],
                new_stats: Optional[Dict[str, Tensor]] = None) -> None:
        """Re-allocate for `name` with `new_n` Gaussians, then `fill(dst)` writes the model, where dst(buffer, attr)
        is the destination view in the fresh arena."""
        self._begin_rebuild(name, new_n)
        fill(lambda b, a: self._views[b][name][a])
        self.stats[name] = new_stats if new_stats is not None else self._zero_stats(new_n)

    @torch.no_grad()
    def densify_and_prune(self, name: str, max_grad: float, min_opacity: float, extent: float, max_screen_size,
                          **kw) -> Dict[str, int]:
        """GaussianModel.densify_and_prune (gaussian_model.py:521-551) of one model: the move kernel writes the
        clone / split / prune result of the parameters and both Adam moments INTO a fresh arena (no staging copy);
        gradient views and sinks are re-derived, the other models' slices copied."""
        old_p = {a: t.detach() for a, t in self.params(name).items() if t.numel()}
        m, v = self.moments(name)
        old_m = {a: m[a] for a in old_p}
        old_v = {a: v[a] for a in old_p}
        st = self.stats[name]
        begun = []

        def alloc(role, a, shape):
            if not begun:
                self._begin_rebuild(name, shape[0])     # the old views above keep the old buffers alive
                begun.append(True)
            return self._views[role][name][a]
        _, _, _, stats, scal = _densify_and_prune(old_p, old_m, old_v, st["xyz_gradient_accum"], st["denom"], max_grad,
                                                   min_opacity, extent, max_screen_size, alloc=alloc, **kw)
        if not begun:                                   # empty model: nothing moved
            return scal
        self.stats[name] = stats
        return scal

    @torch.no_grad()
    def prune(self, name: str, keep: Tensor) -> None:
        """GaussianModel.prune_points(~keep) (gaussian_model.py:416-431): parameters, moments and statistics of
        one model compacted by ONE launch, the parameters and moments straight into a fresh arena"""
        attrs = [a for a in ATTRS if self.params(name)[a].numel()]
        m, v = self.moments(name)
        st = self.stats[name]
        flat = [self.params(name)[a].detach() for a in attrs] + [m[a] for a in attrs] + [v[a] for a in attrs] + \
               [st["xyz_gradient_accum"], st["denom"], st["max_radii2D"]]
        k = len(attrs)
        roles = [("params", a) for a in attrs] + [("exp_avg", a) for a in attrs] + [("exp_avg_sq", a) for a in attrs]
        begun = []

        def alloc(i, shape):
            if not begun:
                self._begin_rebuild(name, shape[0])
                begun.append(True)
            if i < 3 * k:
                return self._views[roles[i][0]][name][roles[i][1]]
            return torch.empty(shape, dtype=torch.float32, device=self.device)
        out = compact_rows(keep, flat, alloc=alloc)
        if not begun:
            return
        self.stats[name] = {"xyz_gradient_accum": out[3 * k], "denom": out[3 * k + 1], "max_radii2D": out[3 * k + 2]}

    # ------------------------------------------------------------------ checkpoints (gaussian_model.py:158-206)
    def _torch_adam(self, name: str, lrs: Dict[str, float], eps: float = 1e-15) -> torch.optim.Adam:
        """a real torch.optim.Adam over the model's parameter views carrying the arena's moments and step — its
        state_dict() is, by construction, what the reference's `self.optimizer.state_dict()` holds"""
        p = self.params(name)
        groups = [{"params": [p[a]], "lr": float(lrs.get(a, 0.0)), "name": a} for a in ATTRS]
        opt = torch.optim.Adam(groups, lr=0.0, eps=eps)
        if self.step[name] > 0:
            m, v = self.moments(name)
            for a in ATTRS:
                opt.state[p[a]] = {"step": torch.tensor(float(self.step[name])), "exp_avg": m[a].clone(),
                                   "exp_avg_sq": v[a].clone()}
        return opt

    def state_dict(self, name: str, is_final: bool = False, lrs: Optional[Dict[str, float]] = None,
                   spatial_lr_scale: float = 1.0, active_sh_degree: Optional[int] = None) -> Dict:
        """GaussianModel.state_dict(is_final) key for key (gaussian_model.py:183-206)"""
        p = self.params(name)
        # key order of the reference dict: xyz, feature_dc, feature_rest, scaling, rotation, opacity, semantic
        sd = {STATE_KEYS[a]: p[a] for a in ("xyz", "f_dc", "f_rest", "scaling", "rotation", "opacity", "semantic")}
        if not is_final:
            st = self.stats[name]
            sd.update({"spatial_lr_scale": spatial_lr_scale, "denom": st["denom"], "max_radii2D": st["max_radii2D"],
                       "xyz_gradient_accum": st["xyz_gradient_accum"],
                       "active_sh_degree": self.specs[name].sh_degree if active_sh_degree is None else active_sh_degree,
                       "optimizer": self._torch_adam(name, lrs or {}).state_dict()})
        return sd

    @torch.no_grad()
    def load_state_dict(self, name: str, sd: Dict) -> None:
        """a checkpoint entry written by the reference's GaussianModel.state_dict (or by state_dict above):
        parameters, statistics and — when present — the Adam moments / step of its optimizer"""
        n = sd["xyz"].shape[0]
        sp = self.specs[name]
        new_spec = ModelSpec(n, int(round((sd["feature_rest"].shape[1] + 1) ** 0.5)) - 1, sd["feature_dc"].shape[1],
                             sd["semantic"].shape[1] if "semantic" in sd else 0)
        self.specs[name] = new_spec
        opt = sd.get("optimizer")

        def fill(dst):
            for i, a in enumerate(ATTRS):
                src = sd.get(STATE_KEYS[a])
                if src is not None and src.numel():
                    dst("params", a).copy_(src.detach().to(self.device).view(dst("params", a).shape))
                if opt is not None and i in opt["state"]:
                    s = opt["state"][i]
                    dst("exp_avg", a).copy_(s["exp_avg"].to(self.device).view(dst("exp_avg", a).shape))
                    dst("exp_avg_sq", a).copy_(s["exp_avg_sq"].to(self.device).view(dst("exp_avg_sq", a).shape))
        stats = None
        if "denom" in sd:
            stats = {"xyz_gradient_accum": sd["xyz_gradient_accum"].to(self.device).float().reshape(n, 2).clone(),
                     "denom": sd["denom"].to(self.device).float().reshape(n, 1).clone(),
                     "max_radii2D": sd["max_radii2D"].to(self.device).float().reshape(n).clone()}
        self.rebuild(name, n, fill, stats)
        if opt is not None and opt["state"]:
            self.step[name] = int(float(next(iter(opt["state"].values()))["step"]))
        del sp

    # ------------------------------------------------------------------ PLY (gaussian_model.py:82-156)
    def ply_attributes(self, name: str) -> List[str]:
        """construct_list_of_attributes (gaussian_model.py:328-342)"""
        sp = self.specs[name]
        l = ["x", "y", "z", "nx", "ny", "nz"]
        l += [f"f_dc_{i}" for i in range(sp.fourier_dim * 3)]
        l += [f"f_rest_{i}" for i in range((sp.K - 1) * 3)]
        l.append("opacity")
        l += [f"scale_{i}" for i in range(3)] + [f"rot_{i}" for i in range(4)]
        l += [f"semantic_{i}" for i in range(sp.num_classes)]
        return l

    @torch.no_grad()
    def make_ply(self, name: str) -> np.ndarray:
        """GaussianModel.make_ply: a structured array, one f4 field per attribute; the rows are packed on the
        device by one kernel (the f_dc / f_rest transposes included) and copied to the host once"""
        sp = self.specs[name]
        if sp.fourier_dim != 1:
            raise NotImplementedError("PLY rows of Fourier actors (f_dc with fourier_dim > 1) keep the reference's "
                                      "transpose(1, 2).flatten layout: not needed for the background / sky models")
        p = self.params(name)
        width = int(lib.sc_ply_row_width(sp.K, sp.num_classes))
        rows = torch.empty((sp.n, width), dtype=torch.float32, device=self.device)
        ptr = lambda t: t.data_ptr() if t.numel() else None
        with torch.cuda.device(self.device):
            check(lib.sc_ply_pack(sp.n, sp.K, sp.num_classes, ptr(p["xyz"]), ptr(p["f_dc"]), ptr(p["f_rest"]),
                                  ptr(p["opacity"]), ptr(p["scaling"]), ptr(p["rotation"]), ptr(p["semantic"]),
                                  rows.data_ptr(), torch.cuda.current_stream(self.device).cuda_stream), "sc_ply_pack")
        host = rows.cpu().numpy()
        names = self.ply_attributes(name)
        assert len(names) == width
        return host.view(np.dtype([(a, "<f4") for a in names])).reshape(-1)

    def save_ply(self, path: str, names: Optional[Sequence[str]] = None) -> None:
        """StreetGaussianModel.save_ply (street_gaussian_model.py:88-100): one `vertex_<model>` element per model in
        a binary little-endian PLY — the file plyfile's PlyData([...]).write produces for these elements"""
        names = list(self.specs) if names is None else list(names)
        elems = [(f"vertex_{n}", self.make_ply(n)) for n in names]
        header = ["ply", "format binary_little_endian 1.0"]
        for en, arr in elems:
            header.append(f"element {en} {arr.shape[0]}")
            header += [f"property float {f}" for f in arr.dtype.names]
        header.append("end_header")
        with open(path, "wb") as f:
            f.write(("\n".join(header) + "\n").encode("ascii"))
            for _, arr in elems:
                f.write(arr.tobytes())

    @torch.no_grad()
    def load_ply(self, path: str) -> List[str]:
        """StreetGaussianModel.load_ply (:102-112): every `vertex_<model>` element whose model exists here"""
        elems = read_ply(path)
        loaded = []
        for en, arr in elems:
            name = en[7:]
            if name not in self.specs:
                continue
            self.load_ply_element(name, arr)
            loaded.append(name)
        return loaded

    @torch.no_grad()
    def load_ply_element(self, name: str, arr: np.ndarray) -> None:
        """GaussianModel.load_ply(input_ply=...) (gaussian_model.py:107-156) on a structured vertex array"""
        fields = arr.dtype.names
        n_dc = len([f for f in fields if f.startswith("f_dc_")])
        n_rest = len([f for f in fields if f.startswith("f_rest_")])
        S = len([f for f in fields if f.startswith("semantic_")])
        K = n_rest // 3 + 1
        assert n_dc == 3, "Fourier actors are not stored through this path"
        sp = ModelSpec(arr.shape[0], int(round(K ** 0.5)) - 1, 1, S)
        self.specs[name] = sp
        order = ["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(3)] + [f"f_rest_{i}" for i in range(n_rest)] + \
                ["opacity"] + [f"scale_{i}" for i in range(3)] + [f"rot_{i}" for i in range(4)] + [f"semantic_{i}" for i in range(S)]
        host = np.stack([np.asarray(arr[f], dtype=np.float32) for f in order], axis=1) if arr.shape[0] else \
            np.zeros((0, len(order)), np.float32)
        rows = torch.from_numpy(np.ascontiguousarray(host)).to(self.device)

        def fill(dst):
            ptr = lambda t: t.data_ptr() if t.numel() else None
            with torch.cuda.device(self.device):
                check(lib.sc_ply_unpack(sp.n, sp.K, S, rows.data_ptr(), ptr(dst("params", "xyz")), ptr(dst("params", "f_dc")),
                                        ptr(dst("params", "f_rest")), ptr(dst("params", "opacity")),
                                        ptr(dst("params", "scaling")), ptr(dst("params", "rotation")),
                                        ptr(dst("params", "semantic")),
                                        torch.cuda.current_stream(self.device).cuda_stream), "sc_ply_unpack")
        self.rebuild(name, sp.n, fill)
        self.step[name] = 0


def read_ply(path: str) -> List[Tuple[str, np.ndarray]]:
    """binary little-endian PLY with float properties -> [(element name, structured array)]"""
    with open(path, "rb") as f:
        data = f.read()
    end = data.index(b"end_header\n") + len(b"end_header\n")
    lines = data[:end].decode("ascii").split("\n")
    assert lines[0] == "ply" and lines[1].startswith("format binary_little_endian"), lines[:2]
    elems: List[Tuple[str, int, List[str]]] = []
    for ln in lines[2:]:
        t = ln.split()
        if not t:
            continue
        if t[0] == "element":
            elems.append((t[1], int(t[2]), []))
        elif t[0] == "property":
            assert t[1] in ("float", "float32"), ln
            elems[-1][2].append(t[2])
    out, off = [], end
    for name, count, props in elems:
        dt = np.dtype([(p, "<f4") for p in props])
        out.append((name, np.frombuffer(data, dtype=dt, count=count, offset=off).copy()))
        off += count * dt.itemsize
    return out
