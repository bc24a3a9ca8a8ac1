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
    def __init__(self, scene: Dict[str, Tensor], device):
        n = scene["means"].shape[0]
        k_sh = scene["shs"].shape[1]
        self.n, self.k_sh = n, k_sh
        self.layout, total = arena_layout(n, k_sh)
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
