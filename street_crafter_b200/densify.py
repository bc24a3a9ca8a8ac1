"""Per-iteration densification statistics: the reference's
`set_max_radii2D` + `add_densification_stats`
(/root/reference/street_gaussian/models/street_gaussian_model.py:486-497, :503-519; called at
/root/reference/train.py:283-284) as ONE kernel over the Gaussians instead of ~10 boolean-index /
cat / norm / max launches.  No CPU path."""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional

import torch
from torch import Tensor

from ._lib import check, lib

MAX_ARRAYS = 32      # SC_COMPACT_MAX_ARRAYS


@torch.no_grad()
def densification_stats(viewspace_point_tensor: Tensor, visibility_filter: Tensor, radii: Tensor, width: int,
                        height: int, xyz_gradient_accum: Tensor, denom: Tensor, max_radii2D: Tensor,
                        start: int = 0, end: int = None) -> None:
    """viewspace_point_tensor: the renderer's `viewspace_points` ([1,N,2] or [N,2]) AFTER backward, carrying
    `.grad` and `.absgrad`; visibility_filter [N] bool; radii [N] float (the renderer's `radii`);
    xyz_gradient_accum [N,2], denom [N,1] or [N], max_radii2D [N] are updated IN PLACE (contiguous fp32 —
    the arrays of ONE model of the reference's scene graph).  `start:end` selects that model's rows of the
    renderer outputs (`graph_gaussian_range[model_name]`); default: all rows."""
    g = viewspace_point_tensor.grad
    a = getattr(viewspace_point_tensor, "absgrad", None)
    if g is None or a is None:
        raise RuntimeError("viewspace_point_tensor needs .grad and .absgrad (render with absgrad=True, then backward)")
    if not g.is_cuda:
        raise RuntimeError("densification_stats needs CUDA tensors: street_crafter_b200 has no CPU path")
    g = g.reshape(-1, 2)[start:end].float().contiguous()
    a = a.reshape(-1, 2)[start:end].float().contiguous()
    N = g.shape[0]
    vis = visibility_filter.reshape(-1)[start:end]
    vis = (vis if vis.dtype == torch.bool else vis != 0).contiguous().view(torch.uint8)
    rad = radii.reshape(-1)[start:end].float().contiguous()
    assert vis.numel() == N and rad.numel() == N
    for t, n in ((xyz_gradient_accum, 2 * N), (denom, N), (max_radii2D, N)):
        assert t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.numel() == n, (t.shape, t.dtype)
    with torch.cuda.device(g.device):
        check(lib.sc_densification_stats(N, a.data_ptr(), g.data_ptr(), vis.data_ptr(), rad.data_ptr(), int(width),
                                         int(height), xyz_gradient_accum.data_ptr(), denom.data_ptr(),
                                         max_radii2D.data_ptr(), torch.cuda.current_stream().cuda_stream),
              "sc_densification_stats")


_KINDS = {"xyz": 2, "scaling": 3}        # SC_DENSIFY_XYZ / SC_DENSIFY_SCALING; every other parameter is SC_DENSIFY_PARAM


@torch.no_grad()
def densify_and_prune(params: Dict[str, Tensor], exp_avg: Dict[str, Tensor], exp_avg_sq: Dict[str, Tensor],
                      xyz_gradient_accum: Tensor, denom: Tensor, max_grad: float, min_opacity: float, extent: float,
                      max_screen_size, percent_dense: float = 0.01, percent_big_ws: float = 0.1, n_copies: int = 2,
                      noise: Optional[Tensor] = None, generator: Optional[torch.Generator] = None,
                      grad_column: int = 0, alloc=None):
    """The reference's `GaussianModel.densify_and_prune(max_grad, min_opacity, prune_big_points, extent,
    max_screen_size)` (/root/reference/street_gaussian/models/gaussian_model.py:521-551: densify_and_clone
    :493-519, densify_and_split :452-491, cat_optimizer :384-408, prune_points :416-431) on plain tensors, as
    classify -> scan -> ONE move launch over parameters and both Adam moments (csrc/densify_cycle.cu).

    params / exp_avg / exp_avg_sq: dicts keyed like the reference's optimizer groups (`xyz`, `f_dc`, `f_rest`,
    `opacity`, `scaling`, `rotation`, optionally `semantic`, ...) of contiguous fp32 CUDA tensors [N, ...];
    a parameter without optimizer state is simply absent from the moment dicts.  `noise` [n_copies*n_split, 3]
    replaces the standard-normal draw behind `torch.normal(mean=0, std=stds)` (:469-470); by default it is drawn
    here with `torch.randn` on the device — the same Philox stream the reference's call consumes.
    `grad_column=1` takes the gradients from `xyz_gradient_accum[:, 1]` (the background / actor models'
    `densify_grad_abs_*` option, gaussian_model_bkgd.py:102-105).
    `alloc(role, name, shape)` (role in params / exp_avg / exp_avg_sq; optional) supplies the destination of every
    output array — model_arena.SceneArena hands out views of a fresh flat arena, so the move kernel writes there.
    -> (params, exp_avg, exp_avg_sq, stats, scalars): stats = the three zeroed statistics (:546-548),
    scalars = the reference's `scalar_dict` (points_total / clone / split / pruned)."""
    xyz, scaling, rotation, opacity = params["xyz"], params["scaling"], params["rotation"], params["opacity"]
    if not xyz.is_cuda:
        raise RuntimeError("densify_and_prune needs CUDA tensors: street_crafter_b200 has no CPU path")
    dev, N = xyz.device, xyz.shape[0]
    names = list(params)
    table = [(params[k], _KINDS.get(k, 0)) for k in names]
    table += [(exp_avg[k], 1) for k in names if k in exp_avg] + [(exp_avg_sq[k], 1) for k in names if k in exp_avg_sq]
    for t, _ in table + [(xyz_gradient_accum, 0), (denom, 0)]:
        assert t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.shape[0] == N, (t.shape, t.dtype)
    assert xyz.shape[1:] == (3,) and scaling.shape[1:] == (3,) and rotation.shape[1:] == (4,) and opacity.numel() == N
    assert xyz_gradient_accum.numel() == 2 * N and denom.numel() == N
    # the float32 thresholds the reference's comparisons see (python-double products vs float32 tensor products)
    clone_max = percent_dense * extent                                                           # :498
    split_min = float(torch.tensor(extent, dtype=torch.float32) * percent_dense)                 # :459, :463
    big_ws = extent * percent_big_ws                                                             # :537
    shrink = 0.8 * n_copies                                                                      # :473
    st = torch.cuda.current_stream().cuda_stream
    zeros = lambda n, *s: torch.zeros((n,) + s, dtype=torch.float32, device=dev)
    if N == 0:
        scal = {"points_total": 0, "points_clone": 0, "points_split": 0, "points_pruned": 0}
        return (dict(params), dict(exp_avg), dict(exp_avg_sq),
                {"xyz_gradient_accum": zeros(0, 2), "denom": zeros(0, 1), "max_radii2D": zeros(0)}, scal)
    with torch.cuda.device(dev):
        flags = torch.empty(4 * N, dtype=torch.int32, device=dev)
        dest = torch.empty(4 * N + 2, dtype=torch.int64, device=dev)     # [4N] scan, total, n_clone_selected
        n_clone_sel = dest[4 * N + 1:].view(torch.int32)
        check(lib.sc_densify_classify(N, xyz_gradient_accum.data_ptr(), denom.data_ptr(), scaling.data_ptr(),
                                      opacity.data_ptr(), max_grad, clone_max, split_min, min_opacity,
                                      1 if max_screen_size else 0, big_ws, shrink, int(grad_column), flags.data_ptr(),
                                      n_clone_sel.data_ptr(), st), "sc_densify_classify")
        ws_bytes = int(lib.sc_scan_workspace_bytes(4 * N))
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        check(lib.sc_exclusive_scan_i32(4 * N, flags.data_ptr(), dest.data_ptr(), dest.data_ptr() + 32 * N,
                                        ws.data_ptr(), ws_bytes, st), "sc_exclusive_scan_i32")
        # one read-back sizes the outputs and the noise (the reference's boolean indexing syncs ~20 times)
        e = dest[[N, 2 * N, 3 * N, 4 * N, 4 * N + 1]].tolist()
        n_stay, n_clone, n_child, n_split = e[0], e[1] - e[0], e[2] - e[1], e[3] - e[2]
        n_clone_selected = e[4] & 0xFFFFFFFF
        if noise is None:
            noise = torch.randn((n_copies * n_split, 3), dtype=torch.float32, device=dev, generator=generator)
        noise = noise.contiguous()
        assert noise.is_cuda and noise.dtype == torch.float32 and tuple(noise.shape) == (n_copies * n_split, 3), noise.shape
        n_out = n_stay + n_clone + n_copies * n_child
        index_ws = torch.empty(max(1, n_clone + 2 * n_child), dtype=torch.int32, device=dev)
        roles = [("params", k) for k in names] + [("exp_avg", k) for k in names if k in exp_avg] + \
                [("exp_avg_sq", k) for k in names if k in exp_avg_sq]
        if alloc is None:
            outs = [torch.empty((n_out,) + tuple(t.shape[1:]), dtype=torch.float32, device=dev) for t, _ in table]
        else:
            outs = [alloc(r, k, (n_out,) + tuple(t.shape[1:])) for (r, k), (t, _) in zip(roles, table)]
            for o, (t, _) in zip(outs, table):
                assert o.is_cuda and o.dtype == torch.float32 and o.is_contiguous() and \
                    tuple(o.shape) == (n_out,) + tuple(t.shape[1:]), (o.shape, t.shape)
        for a in range(0, len(table), MAX_ARRAYS):
            part, part_o = table[a:a + MAX_ARRAYS], outs[a:a + MAX_ARRAYS]
            n = len(part)
            src_arr = (C.c_void_p * n)(*[t.data_ptr() for t, _ in part])
            dst_arr = (C.c_void_p * n)(*[t.data_ptr() for t in part_o])
            w_arr = (C.c_int * n)(*[max(1, t[0].numel()) for t, _ in part])
            k_arr = (C.c_int * n)(*[k for _, k in part])
            if n_out:
                check(lib.sc_densify_move(N, flags.data_ptr(), dest.data_ptr(), n_stay, n_clone, n_child, n_split,
                                          n_copies, noise.data_ptr(), scaling.data_ptr(), rotation.data_ptr(), shrink,
                                          index_ws.data_ptr(), n, src_arr, dst_arr, w_arr, k_arr, st), "sc_densify_move")
    it = iter(outs)
    new_p = {k: next(it) for k in names}
    new_m = {k: next(it) for k in names if k in exp_avg}
    new_v = {k: next(it) for k in names if k in exp_avg_sq}
    stats = {"xyz_gradient_accum": zeros(n_out, 2), "denom": zeros(n_out, 1), "max_radii2D": zeros(n_out)}
    scal = {"points_total": N, "points_clone": n_clone_selected, "points_split": n_split,
            "points_pruned": (N - n_split - n_stay) + (n_clone_selected - n_clone) + n_copies * (n_split - n_child)}
    return new_p, new_m, new_v, stats, scal
