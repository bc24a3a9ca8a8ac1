"""View-sharded multi-GPU step (a NEW capability — the reference trainer is
single-GPU, train.py:112-154).

Every rank holds the full Gaussian set in one flat, attribute-major parameter arena
[means 3N | quats 4N | scales 3N | opacities N | shs 3KN] with an identically laid out
gradient arena.  The parameter tensors handed to the renderer are views of the former; their
`.grad` are views of the latter, so after the local views' backward passes ONE
ncclAllReduce(SUM, fp32) over the contiguous gradient arena (59 floats per Gaussian at K=16,
the reference's Adam groups xyz/f_dc/f_rest/opacity/scaling/rotation,
street_gaussian/models/gaussian_model.py:293-301) finishes the step — no pack kernel, no
per-tensor collectives.  Views are assigned round-robin: view v -> rank v mod world."""
from __future__ import annotations

from typing import Dict, List, Sequence

import torch
import torch.distributed as dist
from torch import Tensor

FIELDS = (("means", 3), ("quats", 4), ("scales", 3), ("opacities", 1))


def arena_layout(n: int, k_sh: int):
    """[(name, offset, numel, shape)] of the attribute-major arena; offsets are multiples of 4
    floats so every view is 16-byte aligned (float4 kernels)."""
    out, off = [], 0
    for name, w in FIELDS:
        out.append((name, off, n * w, (n, w)))
        off += (n * w + 3) // 4 * 4
    out.append(("shs", off, n * k_sh * 3, (n, k_sh, 3)))
    off += (n * k_sh * 3 + 3) // 4 * 4
    return out, off


class GaussianArena:
    def __init__(self, scene: Dict[str, Tensor], device, pad_to: int = 4):
        """pad_to: the buffers' length is rounded up to a multiple of it (4 * world for ShardedAdam, whose
        reduce-scatter / all-gather need equal shards of whole float4s)"""
        n = scene["means"].shape[0]
        k_sh = scene["shs"].shape[1]
        self.n, self.k_sh = n, k_sh
        self.layout, total = arena_layout(n, k_sh)
        self.used = total
        total = (total + pad_to - 1) // pad_to * pad_to
        self.params = torch.zeros(total, dtype=torch.float32, device=device)
        self.grads = torch.zeros(total, dtype=torch.float32, device=device)
        for name, off, numel, shape in self.layout:
            src = scene[name].reshape(-1).to(device)
            self.params[off:off + numel].copy_(src)
            p = self.params[off:off + numel].view(shape).requires_grad_(True)
            p.grad = self.grads[off:off + numel].view(shape)
            setattr(self, name, p)

    def sinks(self) -> Dict[str, Tensor]:
        """views of the gradient arena, keyed by parameter: pass as `grad_sinks=` so that the
        projection / SH backward kernels add straight into the arena (no per-view gradient
        tensors, no AccumulateGrad adds, untouched entries for culled Gaussians)"""
        return {name: self.grads[off:off + numel].view(shape if name != "opacities" else (-1,))
                for name, off, numel, shape in self.layout}

    def tensors(self) -> List[Tensor]:
        return [getattr(self, name) for name, *_ in self.layout]

    def zero_grad(self):
        self.grads.zero_()
        for name, off, numel, shape in self.layout:
            p = getattr(self, name)
            if p.grad is None or p.grad.data_ptr() != self.grads.data_ptr() + 4 * off:
                p.grad = self.grads[off:off + numel].view(shape)


def allreduce_gradients(arena: GaussianArena, group=None) -> None:
    """one SUM all-reduce over the flat gradient arena (NCCL over NVLink on GPUs, gloo in tests)"""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(arena.grads, op=dist.ReduceOp.SUM, group=group)


def shard_views(n_views: int, rank: int, world: int) -> List[int]:
    """view v -> rank v mod world (round-robin keeps the per-rank camera mix balanced)"""
    return [v for v in range(n_views) if v % world == rank]


def allreduce_densification_stats(xyz_gradient_accum: Tensor, denom: Tensor, max_radii2D: Tensor,
                                  group=None) -> None:
    """densification statistics of street_gaussian_model.py:486-519 reduced so that every rank
    takes identical clone/split/prune decisions: accumulators SUM, max_radii2D MAX."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(xyz_gradient_accum, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(denom, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(max_radii2D, op=dist.ReduceOp.MAX, group=group)


class HostViewFeeder:
    """Stages per-view host inputs (pinned camera matrices, target images) onto the device on a
    side stream, so that the copies of view v+1 overlap the render of view v.  Usage:

        feeder = HostViewFeeder(device)
        staged = [feeder.stage(host_tensors_of_view) for view in views]   # enqueue all copies
        for tensors, ready in staged:
            feeder.acquire(tensors, ready)    # compute stream waits for THIS view's copies only
            ... render / loss / backward with `tensors` ...
    """

    def __init__(self, device):
        self.device = torch.device(device)
        self.stream = torch.cuda.Stream(self.device)

    def stage(self, host_tensors: Sequence[Tensor]):
        self.stream.wait_stream(torch.cuda.current_stream(self.device))  # order after prior use of freed blocks
        with torch.cuda.stream(self.stream):
            dev = [t.to(self.device, non_blocking=True) for t in host_tensors]
            ready = torch.cuda.Event()
            ready.record(self.stream)
        return dev, ready

    def acquire(self, dev_tensors: Sequence[Tensor], ready) -> None:
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(ready)
        for t in dev_tensors:
            t.record_stream(cur)  # the caching allocator must not recycle them under the compute stream


class ShardedAdam:
    """Optimizer step of the view-sharded trainer without a replicated update: the gradient
    arena is REDUCE-SCATTERED (half the traffic of the all-reduce), every rank runs the fused Adam kernel
    (csrc/optim.cu: torch.optim.Adam semantics, one learning rate per attribute as
    street_gaussian/models/gaussian_model.py:293-305) on its 1/world slice — exp_avg / exp_avg_sq exist only for
    that slice — and the updated parameters are ALL-GATHERED, optionally on a side stream so that the transfer
    runs under the next step's first kernels (`wait()` before the parameters are read).  With world == 1 it is a
    plain fused Adam over the whole arena."""

    def __init__(self, arena: GaussianArena, lrs: Dict[str, float], betas=(0.9, 0.999), eps: float = 1e-15, group=None):
        self.arena, self.lrs, self.betas, self.eps, self.group = arena, dict(lrs), betas, float(eps), group
        on = dist.is_available() and dist.is_initialized()
        self.world = dist.get_world_size(group) if on else 1
        self.rank = dist.get_rank(group) if on else 0
        total = arena.params.numel()
        if total % (4 * self.world):
            raise ValueError(f"arena length {total} is not a multiple of 4 * world: build it with pad_to=4 * world")
        self.per = total // self.world
        self.lo, self.hi = self.rank * self.per, (self.rank + 1) * self.per
        dev = arena.params.device
        self.exp_avg = torch.zeros(self.per, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(self.per, dtype=torch.float32, device=dev)
        self.t = 0
        self.side = torch.cuda.Stream(dev) if dev.type == "cuda" else None
        self._gathered = None
        starts = [off for _, off, _, _ in arena.layout] + [total]
        self._starts = starts
        self._lr = [float(self.lrs.get(name, 0.0)) for name, *_ in arena.layout]

    def step(self, async_gather: bool = False) -> None:
        import ctypes as C
        from ._lib import check, lib
        a = self.arena
        self.t += 1
        if self.world > 1:
            # in place: the reduced slice of rank r lands at its own offset of the gradient arena
            dist.reduce_scatter_tensor(a.grads[self.lo:self.hi], a.grads, op=dist.ReduceOp.SUM, group=self.group)
        n = len(self._lr)
        st = torch.cuda.current_stream(a.params.device).cuda_stream
        with torch.cuda.device(a.params.device):
            # the moment buffers hold the slice only: hand the kernel their virtual base (element 0 of the arena)
            check(lib.sc_adam_step(n, (C.c_int64 * (n + 1))(*self._starts), (C.c_float * n)(*self._lr), self.lo, self.hi,
                                   a.params.data_ptr(), a.grads.data_ptr(), self.exp_avg.data_ptr() - 4 * self.lo,
                                   self.exp_avg_sq.data_ptr() - 4 * self.lo, float(self.betas[0]), float(self.betas[1]),
                                   self.eps, self.t, st), "sc_adam_step")
        if self.world > 1:
            if async_gather and self.side is not None:
                self.side.wait_stream(torch.cuda.current_stream(a.params.device))
                with torch.cuda.stream(self.side):
                    dist.all_gather_into_tensor(a.params.detach(), a.params.detach()[self.lo:self.hi], group=self.group)
                    self._gathered = torch.cuda.Event()
                    self._gathered.record(self.side)
            else:
                dist.all_gather_into_tensor(a.params.detach(), a.params.detach()[self.lo:self.hi], group=self.group)

    def wait(self) -> None:
        """the compute stream waits for an asynchronous parameter all-gather"""
        if self._gathered is not None:
            torch.cuda.current_stream(self.arena.params.device).wait_event(self._gathered)
            self._gathered = None
