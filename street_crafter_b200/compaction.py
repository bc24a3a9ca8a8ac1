"""Prune / select compaction for the Gaussian set and its optimizer state (SURVEY.md section 8(f) row N4,
first half): the reference rebuilds every parameter, its Adam `exp_avg` / `exp_avg_sq` and the densification
statistics with one boolean-index chain per tensor
(/root/reference/street_gaussian/models/gaussian_model.py:363-382 `prune_optimizer`, :416-431 `prune_points`;
the `[selected_pts_mask]` gathers of `densify_and_clone` / `densify_and_split` are the same operation).  Here
ONE prefix sum of the mask and ONE kernel launch move all of them, rows in their original order, so parameters
and moments stay aligned.  No CPU path."""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Sequence

import torch
from torch import Tensor

from ._lib import check, lib

MAX_ARRAYS = 32


@torch.no_grad()
def compact_rows(keep: Tensor, tensors: Sequence[Tensor], alloc=None) -> List[Tensor]:
    """keep [N] bool; tensors: fp32 CUDA tensors whose first dimension is N (any trailing shape).
    -> new contiguous tensors holding the rows with keep == True, in order (== `t[keep]`)."""
    if not keep.is_cuda:
        raise RuntimeError("compact_rows needs CUDA tensors: street_crafter_b200 has no CPU path")
    N = keep.numel()
    dev = keep.device
    k8 = (keep if keep.dtype == torch.bool else keep != 0).reshape(-1).contiguous().view(torch.uint8)
    srcs = []
    for t in tensors:
        assert t.is_cuda and t.dtype == torch.float32 and t.shape[0] == N, (t.shape, t.dtype)
        srcs.append(t.contiguous())
    if N == 0:
        return [t.clone() for t in srcs]
    st = torch.cuda.current_stream().cuda_stream
    with torch.cuda.device(dev):
        ws_bytes = int(lib.sc_scan_workspace_bytes(N))
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        dest = torch.empty(N + 1, dtype=torch.int64, device=dev)
        check(lib.sc_exclusive_scan_i32(N, k8.to(torch.int32).data_ptr(), dest.data_ptr(), dest.data_ptr() + 8 * N,
                                        ws.data_ptr(), ws_bytes, st), "sc_exclusive_scan_i32")
        n_keep = int(dest[N].item())  # sizes the outputs (the reference's boolean indexing syncs the same way)
        if alloc is None:
            outs = [torch.empty((n_keep,) + tuple(t.shape[1:]), dtype=torch.float32, device=dev) for t in srcs]
        else:   # alloc(i, shape): destination of output i (model_arena.SceneArena: views of a fresh flat arena)
            outs = [alloc(i, (n_keep,) + tuple(t.shape[1:])) for i, t in enumerate(srcs)]
        for a in range(0, len(srcs), MAX_ARRAYS):
            part_s, part_d = srcs[a:a + MAX_ARRAYS], outs[a:a + MAX_ARRAYS]
            n = len(part_s)
            src_arr = (C.c_void_p * n)(*[t.data_ptr() for t in part_s])
            dst_arr = (C.c_void_p * n)(*[t.data_ptr() for t in part_d])
            w_arr = (C.c_int * n)(*[max(1, t[0].numel()) if N else 1 for t in part_s])
            if N and n_keep:
                check(lib.sc_compact_rows(N, k8.data_ptr(), dest.data_ptr(), n, src_arr, dst_arr, w_arr, st),
                      "sc_compact_rows")
    return outs


@torch.no_grad()
def prune_state(keep: Tensor, params: Dict[str, Tensor], exp_avg: Dict[str, Tensor], exp_avg_sq: Dict[str, Tensor],
                stats: Dict[str, Tensor]):
    """`prune_points(~keep)` on plain tensors: parameters, both Adam moments and the per-Gaussian statistics
    (xyz_gradient_accum, denom, max_radii2D) are compacted by ONE launch.  Returns the four dicts rebuilt."""
    names = list(params)
    groups = [params, exp_avg, exp_avg_sq]
    flat = [g[n] for g in groups for n in names if n in g] + [stats[k] for k in stats]
    out = compact_rows(keep, flat)
    res, i = [], 0
    for g in groups:
        d = {}
        for n in names:
            if n in g:
                d[n] = out[i]
                i += 1
        res.append(d)
    res.append({k: out[i + j] for j, k in enumerate(stats)})
    return tuple(res)
