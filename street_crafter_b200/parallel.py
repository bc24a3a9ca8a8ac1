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


def arena_layout(n: int, k_sh: int, align: int = 4):
    """[(name, offset, numel, shape)] of the attribute-major arena; offsets are multiples of 4
    floats so every view is 16-byte aligned (float4 kernels).  The SH block starts (and the arena ends) on
    a multiple of `align`: the geometric block [0, shs offset) and the SH block are then each divisible into
    `align / 4` equal float4 slices (ShardedAdam shards and gathers them separately)."""
    out, off = [], 0
    for name, w in FIELDS:
        out.append((name, off, n * w, (n, w)))
        off += (n * w + 3) // 4 * 4
    off = (off + align - 1) // align * align
    out.append(("shs", off, n * k_sh * 3, (n, k_sh, 3)))
    off += n * k_sh * 3
    off = (off + align - 1) // align * align
    return out, off


class GaussianArena:
    def __init__(self, scene: Dict[str, Tensor], device, pad_to: int = 4):
        """pad_to: the buffers' length is rounded up to a multiple of it (4 * world for ShardedAdam, whose
        reduce-scatter / all-gather need equal shards of whole float4s)"""
        n = scene["means"].shape[0]
        k_sh = scene["shs"].shape[1]
        self.n, self.k_sh = n, k_sh
        self.layout, total = arena_layout(n, k_sh, max(4, pad_to))
        # [geometric block | SH block]: the two ranges ShardedAdam exchanges separately
        self.blocks = [(0, self.layout[-1][1]), (self.layout[-1][1], total)]
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

    def zero_grad(self, stream=None):
        """stream: a side stream to run the fill on (ordered after the current stream's pending work) -> the event
        to wait for before the first backward; the fill then overlaps whatever does not touch the gradients"""
        ev = None
        if stream is not None:
            stream.wait_stream(torch.cuda.current_stream(self.grads.device))
            with torch.cuda.stream(stream):
                self.grads.zero_()
                ev = torch.cuda.Event()
                ev.record(stream)
        else:
            self.grads.zero_()
        for name, off, numel, shape in self.layout:
            p = getattr(self, name)
            if p.grad is None or p.grad.data_ptr() != self.grads.data_ptr() + 4 * off:
                p.grad = self.grads[off:off + numel].view(shape)
        return ev


def allreduce_gradients(arena: GaussianArena, group=None) -> None:
    """one SUM all-reduce over the flat gradient arena (NCCL over NVLink on GPUs, gloo in tests)"""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(arena.grads, op=dist.ReduceOp.SUM, group=group)


class TwoPhaseAllReduce:
    """The all-reduce of the flat gradient arena in its two blocks: `begin_sh()` starts the SH block (81 % of the
    bytes) on a side stream as soon as the last view's SH backward has run, `finish()` reduces the geometric block and
    makes the compute stream wait for both.  Between the two the caller runs the deferred projection-backward
    kernels of its views (rendering.fused_view(deferred=...)), which only write the geometric block."""

    def __init__(self, arena: GaussianArena, group=None):
        self.arena, self.group = arena, group
        on = dist.is_available() and dist.is_initialized()
        self.world = dist.get_world_size(group) if on else 1
        self.side = torch.cuda.Stream(arena.grads.device) if arena.grads.is_cuda and self.world > 1 else None
        self._ev = None

    def begin_sh(self) -> None:
        if self.world == 1:
            return
        a = self.arena
        b0, b1 = a.blocks[1]
        if self.side is None:
            dist.all_reduce(a.grads[b0:b1], op=dist.ReduceOp.SUM, group=self.group)
            return
        self.side.wait_stream(torch.cuda.current_stream(a.grads.device))
        with torch.cuda.stream(self.side):
            dist.all_reduce(a.grads[b0:b1], op=dist.ReduceOp.SUM, group=self.group)
            self._ev = torch.cuda.Event()
            self._ev.record(self.side)

    def finish(self) -> None:
        if self.world == 1:
            return
        a = self.arena
        b0, b1 = a.blocks[0]
        dist.all_reduce(a.grads[b0:b1], op=dist.ReduceOp.SUM, group=self.group)
        if self._ev is not None:
            torch.cuda.current_stream(a.grads.device).wait_event(self._ev)
            self._ev = None


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
    that slice — and the updated parameters are ALL-GATHERED.  The arena's two blocks are exchanged separately:
    the geometric block (means, quats, scales, opacities: 19 % of the bytes) is what the next step's projection and
    binning read; the SH block (81 %) is first read by the SH kernel, so with `async_gather` its all-gather runs on
    a side stream UNDER the projection + binning of the next step's views (rendering.prepare_view).  `wait_geometry()`
    / `wait_sh()` make the compute stream wait for the respective gather.  world == 1: a plain fused Adam."""

    def __init__(self, arena: GaussianArena, lrs: Dict[str, float], betas=(0.9, 0.999), eps: float = 1e-15, group=None,
                 reserve_sms: int = 32):
        self.arena, self.lrs, self.betas, self.eps, self.group = arena, dict(lrs), betas, float(eps), group
        on = dist.is_available() and dist.is_initialized()
        self.world = dist.get_world_size(group) if on else 1
        self.rank = dist.get_rank(group) if on else 0
        if self.world > 1 and arena.params.is_cuda:
            # the asynchronous parameter all-gather runs beside the next step's binning, whose cooperative kernel
            # needs its whole grid resident: keep it off the SMs the collective's CTAs occupy
            from ._lib import check, lib
            check(lib.sc_bin_reserve_sms(int(reserve_sms)), "sc_bin_reserve_sms")
        total = arena.params.numel()
        self.ranges = []          # (block lo, block hi, my lo, my hi, offset of my slice inside the moment buffers)
        mom = 0
        for b0, b1 in arena.blocks:
            if (b1 - b0) % (4 * self.world):
                raise ValueError(f"arena block [{b0}, {b1}) is not a multiple of 4 * world: build the arena with "
                                 f"pad_to=4 * world")
            per = (b1 - b0) // self.world
            self.ranges.append((b0, b1, b0 + self.rank * per, b0 + (self.rank + 1) * per, mom))
            mom += per
        dev = arena.params.device
        self.exp_avg = torch.zeros(max(mom, 4), dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(max(mom, 4), dtype=torch.float32, device=dev)
        self.t = 0
        self.side = torch.cuda.Stream(dev) if dev.type == "cuda" else None
        self._ev = [None, None]
        self._rs_ev = None
        self._starts = [off for _, off, _, _ in arena.layout] + [total]
        self._lr = [float(self.lrs.get(name, 0.0)) for name, *_ in arena.layout]

    def begin_sh_exchange(self) -> None:
        """Reduce-scatter of the SH block on the side stream, to be called as soon as the SH gradients are final —
        after the SH backward of the rank's last view, with the projection backward of its views still to run
        (rendering.fused_view(deferred=...)): 81 % of the gradient bytes then travel under those kernels."""
        if self.world == 1 or self.side is None:
            return
        a = self.arena
        b0, b1, lo, hi, _ = self.ranges[1]
        self.side.wait_stream(torch.cuda.current_stream(a.params.device))
        with torch.cuda.stream(self.side):
            dist.reduce_scatter_tensor(a.grads[lo:hi], a.grads[b0:b1], op=dist.ReduceOp.SUM, group=self.group)
            self._rs_ev = torch.cuda.Event()
            self._rs_ev.record(self.side)

    def step(self, async_gather: bool = False) -> None:
        import ctypes as C
        from ._lib import check, lib
        a = self.arena
        dev = a.params.device
        self.t += 1
        params = a.params.detach()
        n = len(self._lr)
        st = torch.cuda.current_stream(dev).cuda_stream
        for i, (b0, b1, lo, hi, mo) in enumerate(self.ranges):
            if self.world > 1:
                if i == 1 and self._rs_ev is not None:      # begin_sh_exchange already reduced this block
                    torch.cuda.current_stream(dev).wait_event(self._rs_ev)
                    self._rs_ev = None
                else:
                    # in place: the reduced slice of rank r lands at its own offset of the gradient arena
                    dist.reduce_scatter_tensor(a.grads[lo:hi], a.grads[b0:b1], op=dist.ReduceOp.SUM, group=self.group)
            with torch.cuda.device(dev):
                # the moment buffers hold this rank's slices only: hand the kernel their virtual base
                check(lib.sc_adam_step(n, (C.c_int64 * (n + 1))(*self._starts), (C.c_float * n)(*self._lr), lo, hi,
                                       a.params.data_ptr(), a.grads.data_ptr(), self.exp_avg.data_ptr() + 4 * (mo - lo),
                                       self.exp_avg_sq.data_ptr() + 4 * (mo - lo), float(self.betas[0]),
                                       float(self.betas[1]), self.eps, self.t, st), "sc_adam_step")
        if self.world > 1:
            for i, (b0, b1, lo, hi, _) in enumerate(self.ranges):
                if async_gather and self.side is not None:
                    self.side.wait_stream(torch.cuda.current_stream(dev))
                    with torch.cuda.stream(self.side):
                        dist.all_gather_into_tensor(params[b0:b1], params[lo:hi], group=self.group)
                        self._ev[i] = torch.cuda.Event()
                        self._ev[i].record(self.side)
                else:
                    dist.all_gather_into_tensor(params[b0:b1], params[lo:hi], group=self.group)

    def _wait(self, i: int) -> None:
        if self._ev[i] is not None:
            torch.cuda.current_stream(self.arena.params.device).wait_event(self._ev[i])
            self._ev[i] = None

    def wait_geometry(self) -> None:
        """the compute stream waits for the asynchronous all-gather of the geometric block"""
        self._wait(0)

    def wait_sh(self) -> None:
        self._wait(1)

    def wait(self) -> None:
        self._wait(0)
        self._wait(1)
