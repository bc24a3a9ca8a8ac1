"""Two-stream schedule of a multi-view training step on ONE GPU (what a rank does with the views
it owns).

The reference renders one camera per iteration, every stage behind the previous one
(/root/reference/street_gaussian/models/street_gaussian_renderer.py:219-294, train.py:150-200).  The stages of a
camera pass fall into two classes that saturate DIFFERENT resources of a B200:

  * compositing forward + backward        FP32 issue bound, ~0 HBM traffic   (48 % of a cfg2 step)
  * projection, binning, SH, SH backward,
    projection backward                   HBM / latency bound, ~idle FP32    (52 %)

and inside a step nothing of view v+1's forward depends on view v's backward (the parameters change once per
step).  `ViewPipeline.step` therefore runs the compositing kernels of all views back to back on the COMPUTE stream
and everything else on a MEMORY stream, one view ahead for the forward stages and one view behind for the
backward tail:

    compute:                     RF0 RB0       RF1 RB1       RF2 RB2
    memory:   P0 B0 S0 | P1 B1 S1 | T0 | P2 B2 S2 | T1 |             T2

  P projection  B binning  S SH colours  RF / RB compositing forward / backward  T = SH backward + projection
  backward of the view (they ADD into the gradient arena: all T run on the one memory stream, in order).

Results are those of the sequential schedule (same kernels on the same inputs; the floating-point atomics of RB make
gradients differ in the last bits from run to run either way).  Cross-stream tensors are registered with the caching
allocator (`record_stream`)."""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch
from torch import Tensor

from .renderer import render_kernel_gsplat
from .rendering import prepare_view_begin, prepare_view_finish, prepared_tensors


class ViewPipeline:
    def __init__(self, device, ahead: int = 1, memory_priority: int = -1, pipeline_sh: bool = True,
                 pipeline_tail: bool = True):
        """ahead: how many views the memory stream prepares before the first compositing kernel starts (>= 1;
        len(views) = all of them up front).  memory_priority: stream priority of the memory stream (-1 = high: its short
        CTAs take the slots that retiring compositing CTAs free)."""
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("ViewPipeline needs a CUDA device: street_crafter_b200 has no CPU path")
        self.ahead = max(1, int(ahead))
        self.pipeline_sh = bool(pipeline_sh)
        self.pipeline_tail = bool(pipeline_tail)
        self.mem = torch.cuda.Stream(self.device, priority=memory_priority)

    def step(self, means: Tensor, shs: Tensor, scales: Tensor, quats: Tensor, opacities: Tensor,
             views: Sequence[Tuple[Tensor, Tensor, Tensor]], width: int, height: int, sh_degree: int,
             grad_sinks: Optional[Dict[str, Tensor]], backward: Optional[Callable[[int, Dict[str, Tensor]], None]],
             znear: float = 0.001, zfar: float = 1000.0, antialiasing: bool = True, absgrad: bool = True,
             capacities: Optional[Sequence[int]] = None, counts_host: Optional[Sequence[Tensor]] = None,
             collect_projection_backward: Optional[List] = None) -> List[Dict[str, Tensor]]:
        """views: (w2c[1,4,4], K[1,3,3], camera_center[3]) on the device.  `backward(v, out)` attaches the caller's loss
        to the render of view v and calls .backward() on it (on the compute stream = the current stream); None (with
        `grad_sinks=None`, under torch.no_grad()): a render-only pass (evaluation: colours clamped to [0, 1]).  The gradient
        arena behind `grad_sinks` must have been zeroed on the current stream.  On return the compute stream is ordered
        after every kernel of the step.  `capacities` (one per view) selects the sync-free binning
        (rendering.prepare_view_begin): no host wait anywhere in the step; `counts_host[v]` (pinned int64[3]) receives
        the view's counts for the capacity check.  `collect_projection_backward` (a list): the views' projection
        backward launches are appended to it instead of being run — the view-sharded trainer runs them while the SH
        gradients, final without them, are already being exchanged (parallel.TwoPhaseAllReduce / ShardedAdam)."""
        comp = torch.cuda.current_stream(self.device)
        mem = self.mem
        n = len(views)
        op1 = opacities if opacities.dim() == 1 else opacities[:, 0]
        op2 = opacities if opacities.dim() == 2 else opacities[:, None]
        mem.wait_stream(comp)  # parameters / zeroed gradients written on the compute stream

        def begin(v):
            V, K, C = views[v]
            with torch.cuda.stream(mem):
                return prepare_view_begin(means, quats, scales, op1, V, K, width, height, near_plane=znear,
                                          far_plane=zfar, antialiased=antialiasing,
                                          shs=shs if self.pipeline_sh else None, campos=C, sh_degree=sh_degree,
                                          capacity=None if capacities is None else int(capacities[v]),
                                          counts_host=None if counts_host is None else counts_host[v])

        def finish(pb):
            with torch.cuda.stream(mem):
                pb = prepare_view_finish(pb, False)
                ev = torch.cuda.Event()
                ev.record(mem)
            for t in prepared_tensors(pb):
                t.record_stream(comp)
            return pb, ev

        ready: List = []      # (prepared state, event) per view, in order
        pending = None        # a begun, not yet finished view
        nxt = 0
        for _ in range(min(self.ahead, n)):
            ready.append(finish(begin(nxt)))
            nxt += 1
        outs = []
        for v in range(n):
            if nxt < n:
                pending = begin(nxt)     # enqueued BEFORE the host blocks on anything of view v
                nxt += 1
            pb, ev = ready.pop(0)
            comp.wait_event(ev)
            V, K, C = views[v]
            deferred: List = []
            out = render_kernel_gsplat(means, shs, scales, quats, op2, V, K, C, width, height, sh_degree, znear=znear,
                                       zfar=zfar, antialiasing=antialiasing, absgrad=absgrad, grad_sinks=grad_sinks,
                                       aux_ids=False, prepared=pb, deferred=deferred, train=backward is not None,
                                       defer_sh=self.pipeline_tail and collect_projection_backward is None)
            if backward is not None:
                backward(v, out)
            outs.append(out)
            if pending is not None:
                # the counts of the next view have long arrived: the compute stream holds RF + RB of this view
                ready.append(finish(pending))
                pending = None
            if collect_projection_backward is not None:
                collect_projection_backward.extend(deferred)
            elif self.pipeline_tail:
                done = torch.cuda.Event()
                done.record(comp)
                with torch.cuda.stream(mem):
                    mem.wait_event(done)
                    for fn in deferred:
                        fn()
            else:
                for fn in deferred:
                    fn()
        comp.wait_stream(mem)
        return outs


class GraphedStep:
    """A multi-view step of `ViewPipeline` captured ONCE in a CUDA graph and replayed: the host launches one graph per
    step instead of ~50 kernels, allocations, autograd nodes and two read-backs per view (measured on cfg2: 3.3 ms of
    host work for a 4.0 ms step — the sequential step is host-bound, the overlapped one would be entirely).

    Needs the sync-free binning, i.e. a CAPACITY per view for its intersection list: measured here by one eager pass,
    times `margin`.  `replay()` returns False when a view outgrew its capacity in that replay (the scene moved or
    densified): the step's results are then incomplete — call `recapture()` (it re-measures) and replay again.
    Camera matrices, upstream tensors etc. are STATIC tensors: write new values into them before `replay()`.
    Re-capture after every densification (the parameter tensors change)."""

    def __init__(self, pipe: ViewPipeline, zero_grad: Optional[Callable[[], None]],
                 params: Callable[[], Tuple[Tensor, ...]],
                 views: Sequence[Tuple[Tensor, Tensor, Tensor]], width: int, height: int, sh_degree: int,
                 grad_sinks: Optional[Callable[[], Dict[str, Tensor]]],
                 backward: Optional[Callable[[int, Dict[str, Tensor]], None]],
                 margin: float = 1.25, before: Optional[Callable[[], None]] = None,
                 after: Optional[Callable[[], None]] = None, split_projection_backward: bool = False,
                 **render_kwargs):
        """params() -> (means, shs, scales, quats, opacities); zero_grad() clears the gradient arena on the current
        stream; `zero_grad = grad_sinks = backward = None`: a RENDER-ONLY graph (trajectory / novel-view rendering:
        rewrite the camera tensors, replay, read `outs[v]["rgb"]`); before() / after() (optional) are captured in front of / behind the views (host -> device copies of
        the step's inputs from pinned memory; the optimiser step on the arena, the loss read-back).
        split_projection_backward: TWO graphs — `replay_main()` ends with the last view's SH backward (the SH
        gradients, 81 % of the arena, are final), `replay_tail()` holds the views' projection backward: a
        multi-GPU step starts the exchange of the SH gradients between the two."""
        self.pipe, self.zero_grad, self.params, self.views = pipe, zero_grad, params, list(views)
        self.width, self.height, self.sh_degree = int(width), int(height), int(sh_degree)
        self.grad_sinks, self.backward, self.margin, self.after = grad_sinks, backward, float(margin), after
        self.before = before
        self.split = bool(split_projection_backward)
        self.graph_tail: Optional[torch.cuda.CUDAGraph] = None
        self.render_kwargs = render_kwargs
        self.counts_host = [torch.zeros(3, dtype=torch.int64).pin_memory() for _ in self.views]
        self.capacities: List[int] = []
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.outs: List[Dict[str, Tensor]] = []
        self.recapture()

    def _run(self):
        if self.before is not None:
            self.before()
        if self.backward is None:   # render only
            with torch.no_grad():
                means, shs, scales, quats, opacities = (p.detach() for p in self.params())
                return self.pipe.step(means, shs, scales, quats, opacities, self.views, self.width, self.height,
                                      self.sh_degree, None, None, capacities=self.capacities,
                                      counts_host=self.counts_host, **self.render_kwargs)
        self.zero_grad()
        # fresh leaves over the same storage: autograd ties a leaf's gradient accumulator to the stream the leaf was
        # first used on; the caller's tensors may have met the default stream, which a capture must not wait for
        # (gradients go to the sinks, the accumulators stay empty)
        means, shs, scales, quats, opacities = (p.detach().requires_grad_(True) for p in self.params())
        self._tail = [] if self.split else None
        outs = self.pipe.step(means, shs, scales, quats, opacities, self.views, self.width, self.height, self.sh_degree,
                              self.grad_sinks(), self.backward, capacities=self.capacities,
                              counts_host=self.counts_host, collect_projection_backward=self._tail,
                              **self.render_kwargs)
        if not self.split and self.after is not None:
            self.after()
        return outs

    def _run_tail(self):
        for fn in self._tail:
            fn()
        if self.after is not None:
            self.after()

    def measure(self) -> List[int]:
        """intersection counts of the views as they are now (one eager, synchronising projection + binning each)"""
        from .rendering import prepare_view
        if self.before is not None:
            self.before()
        means, shs, scales, quats, opacities = self.params()
        op1 = opacities if opacities.dim() == 1 else opacities[:, 0]
        kw = self.render_kwargs
        n = []
        for V, K, _ in self.views:
            pb = prepare_view(means, quats, scales, op1, V, K, self.width, self.height,
                              near_plane=kw.get("znear", 0.001), far_plane=kw.get("zfar", 1000.0),
                              antialiased=kw.get("antialiasing", True))
            n.append(int(pb["flatten_ids"].numel()))
        return n

    def recapture(self) -> None:
        dev = self.pipe.device
        self.graph, self.outs = None, []
        self.capacities = [max(4096, int(c * self.margin) + 4096) for c in self.measure()]
        # warm-up on a side stream (allocator pools, lazy module loading, cudaFuncSetAttribute calls), then capture
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            self._run()
            if self.split:
                self._run_tail()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        from ._lib import launch_count
        l0 = launch_count()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            outs = self._run()
        g2 = None
        if self.split:
            g2 = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g2, pool=g.pool()):
                self._run_tail()
            self._tail = None
        self.launches_per_replay = int(launch_count() - l0)   # this library's kernels inside one replay
        self.graph, self.graph_tail, self.outs = g, g2, outs

    def overflowed(self) -> List[int]:
        """views whose list outgrew its capacity in the LAST COMPLETED replay (synchronise first)"""
        return [v for v, (h, c) in enumerate(zip(self.counts_host, self.capacities)) if int(h[1]) > c]

    def replay(self) -> List[Dict[str, Tensor]]:
        self.graph.replay()
        if self.graph_tail is not None:
            self.graph_tail.replay()
        return self.outs

    def replay_main(self) -> List[Dict[str, Tensor]]:
        self.graph.replay()
        return self.outs

    def replay_tail(self) -> None:
        self.graph_tail.replay()
