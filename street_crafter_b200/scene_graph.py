"""Scene-graph assembly of the projection kernel's `means` / `quats` inputs (SURVEY.md section 8(f) row N2,
geometric half): the reference's `StreetGaussianModel.get_xyz` / `get_rotation`
(/root/reference/street_gaussian/models/street_gaussian_model.py:296-364) — per-model cat, flip-mask clone /
masked index, quaternion_to_matrix + einsum, two quaternion_raw_multiply and two normalize, each with its autograd
node — as ONE forward and ONE backward kernel over the concatenated Gaussian set (csrc/assemble.cu).  No CPU path."""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Sequence

import torch
from torch import Tensor

from ._lib import check, lib

MAX_SEGMENTS = 32      # SC_ASSEMBLE_MAX_SEGMENTS
FLIP_QUAT_Y = (0.0, 0.0, 1.0, 0.0)   # matrix_to_quaternion(diag(-1, 1, -1)), street_gaussian_model.py:55-58


def _tables(segs, starts, flip_quat, shells):
    n = len(segs)
    ptr = lambda t: t.data_ptr() if t is not None and t.numel() else None
    arr = lambda xs: (C.c_void_p * n)(*xs)
    shell = (C.c_float * (4 * n))(*[v for sh in shells for v in (sh if sh is not None else (0.0, 0.0, 0.0, 0.0))])
    return ((C.c_int64 * (n + 1))(*starts), arr([ptr(s[0]) for s in segs]), arr([ptr(s[1]) for s in segs]),
            arr([s[2].data_ptr() if s[2] is not None else None for s in segs]),
            arr([s[3].data_ptr() if s[3] is not None else None for s in segs]), arr([ptr(s[4]) for s in segs]),
            (C.c_float * 4)(*flip_quat), shell)


def _segments_of(kinds, flips, tensors):
    segs, it = [], iter(tensors)
    for actor, flip in zip(kinds, flips):
        xyz, rot = next(it), next(it)
        pr, pt = (next(it), next(it)) if actor else (None, None)
        segs.append((xyz, rot, pr, pt, flip))
    return segs


class _Assemble(torch.autograd.Function):
    @staticmethod
    def forward(ctx, meta, *tensors):
        kinds, flips, flip_axis, flip_quat, shells = meta
        dev = tensors[0].device
        tensors = tuple(t.contiguous() for t in tensors)
        segs = _segments_of(kinds, flips, tensors)
        starts = [0]
        for s in segs:
            starts.append(starts[-1] + s[0].shape[0])
        N = starts[-1]
        out_xyz = torch.empty((N, 3), dtype=torch.float32, device=dev)
        out_rot = torch.empty((N, 4), dtype=torch.float32, device=dev)
        st = torch.cuda.current_stream().cuda_stream
        with torch.cuda.device(dev):
            for a in range(0, len(segs), MAX_SEGMENTS):
                part = segs[a:a + MAX_SEGMENTS]
                t = _tables(part, starts[a:a + len(part) + 1], flip_quat, shells[a:a + len(part)])
                check(lib.sc_assemble_fwd(len(part), t[0], t[1], t[2], t[3], t[4], t[5], flip_axis, t[6], t[7],
                                          out_xyz.data_ptr(), out_rot.data_ptr(), st), "sc_assemble_fwd")
        # inputs go through save_for_backward (autograd's version counters guard against an in-place parameter
        # or pose update between forward and backward); only pointer-free metadata stays on ctx
        ctx.save_for_backward(*tensors)
        ctx.starts, ctx.meta = starts, meta
        return out_xyz, out_rot

    @staticmethod
    def backward(ctx, v_xyz, v_rot):
        kinds, flips, flip_axis, flip_quat, shells = ctx.meta
        segs, starts = _segments_of(kinds, flips, ctx.saved_tensors), ctx.starts
        dev = segs[0][0].device
        v_xyz = (v_xyz if v_xyz is not None else torch.zeros((starts[-1], 3), device=dev)).float().contiguous()
        v_rot = (v_rot if v_rot is not None else torch.zeros((starts[-1], 4), device=dev)).float().contiguous()
        need = iter(ctx.needs_input_grad[1:])
        # three allocations for all segments: per-Gaussian gradients are row slices of two [N, .] buffers, the pose
        # gradients rows of one zeroed [S, 7] buffer (the kernel accumulates into them)
        N = starts[-1]
        all_gx = torch.empty((N, 3), dtype=torch.float32, device=dev)
        all_gr = torch.empty((N, 4), dtype=torch.float32, device=dev)
        all_gp = torch.zeros((len(segs), 7), dtype=torch.float32, device=dev)
        grads, gsegs = [None], []
        for i, ((xyz, rot, pr, pt, _), actor) in enumerate(zip(segs, kinds)):
            gx = all_gx[starts[i]:starts[i + 1]] if next(need) else None
            gr = all_gr[starts[i]:starts[i + 1]] if next(need) else None
            gpr = gpt = None
            if actor:
                gpr = all_gp[i, :4] if next(need) else None
                gpt = all_gp[i, 4:] if next(need) else None
            gsegs.append((gx, gr, gpr, gpt))
            grads += [gx, gr] + ([gpr.view_as(pr) if gpr is not None else None,
                                  gpt.view_as(pt) if gpt is not None else None] if actor else [])
        st = torch.cuda.current_stream().cuda_stream
        ptr = lambda t: t.data_ptr() if t is not None and t.numel() else None
        with torch.cuda.device(dev):
            for a in range(0, len(segs), MAX_SEGMENTS):
                part, gpart = segs[a:a + MAX_SEGMENTS], gsegs[a:a + MAX_SEGMENTS]
                n = len(part)
                t = _tables(part, starts[a:a + n + 1], flip_quat, shells[a:a + n])
                g = [(C.c_void_p * n)(*[ptr(gs[j]) for gs in gpart]) for j in range(4)]
                check(lib.sc_assemble_bwd(n, t[0], t[1], t[2], t[3], t[4], t[5], flip_axis, t[6], t[7], v_xyz.data_ptr(),
                                          v_rot.data_ptr(), g[0], g[1], g[2], g[3], st), "sc_assemble_bwd")
        return tuple(grads)


def assemble_xyz_rotation(segments: Sequence[Dict[str, Optional[Tensor]]], flip_quat: Sequence[float] = FLIP_QUAT_Y,
                          flip_axis: int = 1):
    """segments, in scene-graph order (background, visible actors, sky — `parse_camera`,
    street_gaussian_model.py:216-243): dicts with `xyz` [n,3] and `rotation` [n,4] (the model's RAW `_rotation`);
    actors add `pose_rot` [4] (ActorPose.get_tracking_rotation, raw wxyz), `pose_trans` [3] and optionally `flip`
    [n] bool (the training-time flip mask, :262-271); the sky model adds `shell` = (centre xyz, radius): its
    positions inside twice that radius are pushed out onto the sphere of radius 2 R (GaussianModelSky.get_xyz,
    gaussian_model_sky.py:68-76).  -> (`get_xyz` [N,3], `get_rotation` [N,4]), differentiable w.r.t. every float
    input (tracked poses included)."""
    if not segments:
        raise ValueError("assemble_xyz_rotation needs at least one segment")
    kinds, flips, tensors, shells = [], [], [], []
    for s in segments:
        xyz, rot = s["xyz"], s["rotation"]
        sh = s.get("shell")
        if sh is not None:
            c, r = sh
            c = [float(v) for v in (c.tolist() if torch.is_tensor(c) else c)]
            r = float(r.item() if torch.is_tensor(r) else r)
            # 2 * sphere_radius as the reference forms it: a float32 product (gaussian_model_sky.py:73)
            sh = (c[0], c[1], c[2], float(torch.tensor(r, dtype=torch.float32) * 2))
            assert s.get("pose_rot") is None, "the sky shell applies to a plain (unposed) segment"
        shells.append(sh)
        if not (xyz.is_cuda and rot.is_cuda):
            raise RuntimeError("assemble_xyz_rotation needs CUDA tensors: street_crafter_b200 has no CPU path")
        assert xyz.dtype == torch.float32 and rot.dtype == torch.float32 and xyz.shape[1:] == (3,) and \
            rot.shape == (xyz.shape[0], 4), (xyz.shape, rot.shape)
        actor = s.get("pose_rot") is not None
        kinds.append(actor)
        tensors += [xyz, rot]
        flip = None
        if actor:
            pr, pt = s["pose_rot"], s["pose_trans"]
            assert pr.numel() == 4 and pt.numel() == 3 and pr.is_cuda and pt.is_cuda
            tensors += [pr, pt]
            if s.get("flip") is not None:
                flip = s["flip"]
                flip = (flip if flip.dtype == torch.bool else flip != 0).reshape(-1).contiguous().view(torch.uint8)
                assert flip.numel() == xyz.shape[0] and flip.is_cuda
        flips.append(flip)
    meta = (tuple(kinds), tuple(flips), int(flip_axis), tuple(float(v) for v in flip_quat), tuple(shells))
    return _Assemble.apply(meta, *tensors)


class _AssembleAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, clamp_max, *tensors):
        dev = tensors[0].device
        segs = [(tensors[i].contiguous(), tensors[i + 1].contiguous()) for i in range(0, len(tensors), 2)]
        starts = [0]
        for s in segs:
            starts.append(starts[-1] + s[0].shape[0])
        N = starts[-1]
        out_s = torch.empty((N, 3), dtype=torch.float32, device=dev)
        out_o = torch.empty((N, 1), dtype=torch.float32, device=dev)
        st = torch.cuda.current_stream().cuda_stream
        with torch.cuda.device(dev):
            for a in range(0, len(segs), MAX_SEGMENTS):
                t = _AssembleAct._tables(segs[a:a + MAX_SEGMENTS], starts[a:a + MAX_SEGMENTS + 1], clamp_max[a:a + MAX_SEGMENTS])
                check(lib.sc_assemble_act_fwd(t[0], t[1], t[2], t[3], t[4], out_s.data_ptr(), out_o.data_ptr(), st),
                      "sc_assemble_act_fwd")
        ctx.save_for_backward(*[t for sg in segs for t in sg])   # version-counter checked (not ctx attributes)
        ctx.starts, ctx.clamp_max = starts, clamp_max
        return out_s, out_o

    @staticmethod
    def _tables(part, starts, clamp):
        n = len(part)
        ptr = lambda t: t.data_ptr() if t is not None and t.numel() else None
        return (n, (C.c_int64 * (n + 1))(*starts), (C.c_void_p * n)(*[ptr(s[0]) for s in part]),
                (C.c_void_p * n)(*[ptr(s[1]) for s in part]), (C.c_float * n)(*clamp))

    @staticmethod
    def backward(ctx, v_s, v_o):
        saved = ctx.saved_tensors
        segs = [(saved[i], saved[i + 1]) for i in range(0, len(saved), 2)]
        starts, clamp_max = ctx.starts, ctx.clamp_max
        dev, N = segs[0][0].device, starts[-1]
        v_s = (v_s if v_s is not None else torch.zeros((N, 3), device=dev)).float().contiguous()
        v_o = (v_o if v_o is not None else torch.zeros((N, 1), device=dev)).float().contiguous()
        all_gs = torch.empty((N, 3), dtype=torch.float32, device=dev)
        all_go = torch.empty((N,), dtype=torch.float32, device=dev)
        need = iter(ctx.needs_input_grad[1:])
        gsegs = []
        for i, (sc, op) in enumerate(segs):
            gs = all_gs[starts[i]:starts[i + 1]] if next(need) else None
            go = all_go[starts[i]:starts[i + 1]].view_as(op) if next(need) else None
            gsegs.append((gs, go))
        st = torch.cuda.current_stream().cuda_stream
        ptr = lambda t: t.data_ptr() if t is not None and t.numel() else None
        with torch.cuda.device(dev):
            for a in range(0, len(segs), MAX_SEGMENTS):
                part, gpart = segs[a:a + MAX_SEGMENTS], gsegs[a:a + MAX_SEGMENTS]
                t = _AssembleAct._tables(part, starts[a:a + len(part) + 1], clamp_max[a:a + MAX_SEGMENTS])
                g = [(C.c_void_p * len(part))(*[ptr(gs[j]) for gs in gpart]) for j in range(2)]
                check(lib.sc_assemble_act_bwd(t[0], t[1], t[2], t[3], t[4], v_s.data_ptr(), v_o.data_ptr(), g[0], g[1], st),
                      "sc_assemble_act_bwd")
        return (None,) + tuple(x for gs in gsegs for x in gs)


def assemble_scaling_opacity(segments: Sequence[Dict[str, Optional[Tensor]]]):
    """segments in scene-graph order: dicts with the models' RAW `scaling` [n,3] (log) and `opacity` [n,1] (logit);
    the sky model adds `scaling_clamp_max` (its sphere radius, gaussian_model_sky.py:63-66).
    -> (`get_scaling` [N,3], `get_opacity` [N,1]) of street_gaussian_model.py:273-294 / :389-407, differentiable."""
    if not segments:
        raise ValueError("assemble_scaling_opacity needs at least one segment")
    tensors, clamp = [], []
    for s in segments:
        sc, op = s["scaling"], s["opacity"]
        if not (sc.is_cuda and op.is_cuda):
            raise RuntimeError("assemble_scaling_opacity needs CUDA tensors: street_crafter_b200 has no CPU path")
        assert sc.dtype == torch.float32 and op.dtype == torch.float32 and sc.shape[1:] == (3,) and \
            op.numel() == sc.shape[0], (sc.shape, op.shape)
        tensors += [sc, op]
        cm = s.get("scaling_clamp_max")
        clamp.append(float("inf") if cm is None else float(cm))
    return _AssembleAct.apply(tuple(clamp), *tensors)


MAX_FOURIER = 8        # SC_ASSEMBLE_MAX_FOURIER


class _AssembleFeatures(torch.autograd.Function):
    @staticmethod
    def _tables(starts, n_fourier, bases):
        n = len(n_fourier)
        flat = []
        for b in bases:
            flat += list(b) + [0.0] * (MAX_FOURIER - len(b))
        return n, (C.c_int64 * (n + 1))(*starts), (C.c_int * n)(*n_fourier), (C.c_float * (n * MAX_FOURIER))(*flat)

    @staticmethod
    def forward(ctx, bases, *tensors):
        dev = tensors[0].device
        segs = [(tensors[i].contiguous(), tensors[i + 1].contiguous()) for i in range(0, len(tensors), 2)]
        K = segs[0][1].shape[1] + 1
        starts = [0]
        for dc, _ in segs:
            starts.append(starts[-1] + dc.shape[0])
        out = torch.empty((starts[-1], K, 3), dtype=torch.float32, device=dev)
        st = torch.cuda.current_stream().cuda_stream
        ptr = lambda t: t.data_ptr() if t.numel() else None
        with torch.cuda.device(dev):
            for a in range(0, len(segs), MAX_SEGMENTS):
                part = segs[a:a + MAX_SEGMENTS]
                n, c_starts, c_nf, c_base = _AssembleFeatures._tables(starts[a:a + len(part) + 1],
                                                                     [s[0].shape[1] for s in part], bases[a:a + len(part)])
                check(lib.sc_assemble_features_fwd(n, c_starts, K, (C.c_void_p * n)(*[ptr(s[0]) for s in part]),
                                                   (C.c_void_p * n)(*[ptr(s[1]) for s in part]), c_nf, c_base,
                                                   out.data_ptr(), st), "sc_assemble_features_fwd")
        ctx.shapes, ctx.starts, ctx.bases, ctx.K = [(s[0].shape, s[1].shape) for s in segs], starts, bases, K
        return out

    @staticmethod
    def backward(ctx, v_out):
        v_out = v_out.float().contiguous()
        dev, starts, K = v_out.device, ctx.starts, ctx.K
        need = iter(ctx.needs_input_grad[1:])
        gsegs = []
        for dc_shape, rest_shape in ctx.shapes:
            gdc = torch.empty(dc_shape, dtype=torch.float32, device=dev) if next(need) else None
            grest = torch.empty(rest_shape, dtype=torch.float32, device=dev) if next(need) else None
            gsegs.append((gdc, grest))
        st = torch.cuda.current_stream().cuda_stream
        ptr = lambda t: t.data_ptr() if t is not None and t.numel() else None
        with torch.cuda.device(dev):
            for a in range(0, len(gsegs), MAX_SEGMENTS):
                part = gsegs[a:a + MAX_SEGMENTS]
                n, c_starts, c_nf, c_base = _AssembleFeatures._tables(
                    starts[a:a + len(part) + 1], [s[0][1] for s in ctx.shapes[a:a + len(part)]], ctx.bases[a:a + len(part)])
                check(lib.sc_assemble_features_bwd(n, c_starts, K, c_nf, c_base, v_out.data_ptr(),
                                                   (C.c_void_p * n)(*[ptr(g[0]) for g in part]),
                                                   (C.c_void_p * n)(*[ptr(g[1]) for g in part]), st),
                      "sc_assemble_features_bwd")
        return (None,) + tuple(x for g in gsegs for x in g)


def assemble_features(segments: Sequence[Dict[str, Optional[Tensor]]]):
    """segments in scene-graph order: dicts with `f_dc` [n,C,3] and `f_rest` [n,K-1,3]; actors (C = fourier_dim) add
    `idft_base` — the C floats `IDFT(time, C)[0]` of gaussian_model_actor.py:67-71 (plain models: C = 1, base 1).
    -> `get_features` [N,K,3] of street_gaussian_model.py:366-386, differentiable."""
    if not segments:
        raise ValueError("assemble_features needs at least one segment")
    tensors, bases = [], []
    K = segments[0]["f_rest"].shape[1] + 1
    for s in segments:
        dc, rest = s["f_dc"], s["f_rest"]
        if not (dc.is_cuda and rest.is_cuda):
            raise RuntimeError("assemble_features needs CUDA tensors: street_crafter_b200 has no CPU path")
        assert dc.dtype == torch.float32 and rest.dtype == torch.float32 and dc.dim() == 3 and dc.shape[2] == 3 and \
            rest.shape == (dc.shape[0], K - 1, 3) and 1 <= dc.shape[1] <= MAX_FOURIER, (dc.shape, rest.shape)
        base = s.get("idft_base")
        base = [1.0] if base is None else [float(v) for v in (base.tolist() if isinstance(base, Tensor) else base)]
        assert len(base) == dc.shape[1], (len(base), dc.shape)
        tensors += [dc, rest]
        bases.append(tuple(base))
    return _AssembleFeatures.apply(tuple(bases), *tensors)
