"""NCCL correctness of the view-sharded step; run under torchrun on >= 2 GPUs:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 tools/nccl_parity.py

  1. N-rank sharded views + ONE all-reduce of the flat gradient arena == one rank rendering every view sequentially;
  2. all-reduced densification statistics leave every rank with IDENTICAL clone / split / prune flags;
  3. reduce-scatter + sharded fused Adam + all-gather == all-reduce + torch.optim.Adam on the full arena;
  4. the sharded step replayed from two CUDA graphs around the start of the SH-block all-reduce == sequential.
Exit code 0 and "nccl parity ok" on rank 0 when all four hold."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

GRAD_RTOL = 1e-3


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    dist.init_process_group("nccl", device_id=dev)
    from street_crafter_b200._lib import check, lib
    from street_crafter_b200.densify import densification_stats
    from street_crafter_b200.parallel import (GaussianArena, ShardedAdam, TwoPhaseAllReduce, allreduce_densification_stats,
                                              shard_views)
    from street_crafter_b200.renderer import render_kernel_gsplat
    from tools.scene import make_camera, make_gaussians

    n, W, H, deg, n_views = 60000, 640, 400, 3, 4
    scene = make_gaussians(n, sh_degree=deg, seed=0)
    cams = [make_camera(W, H, yaw_deg=y, z_shift=z) for y, z in ((0.0, 0.0), (30.0, 1.5), (-30.0, 3.0), (10.0, 4.5))]
    g = torch.Generator().manual_seed(3)
    ups = [(torch.randn(1, H, W, 4, generator=g).to(dev) * torch.tensor([1, 1, 1, 0.02], device=dev),
            torch.randn(1, H, W, 1, generator=g).to(dev)) for _ in range(n_views)]

    def run_views(arena, views, stats=None, deferred=None):
        arena.zero_grad()
        sinks = arena.sinks()
        for v in views:
            Vm, Km = cams[v]
            cc = (-(Vm[:3, :3].T @ Vm[:3, 3])).to(dev)
            out = render_kernel_gsplat(arena.means, arena.shs, arena.scales, arena.quats, arena.opacities, Vm[None].to(dev),
                                       Km[None].to(dev), cc, W, H, deg, grad_sinks=sinks, deferred=deferred)
            torch.autograd.backward([out["render_colors"], out["render_alphas"]], list(ups[v]))
            if stats is not None:
                densification_stats(out["viewspace_points"], out["visibility_filter"], out["radii"], W, H, *stats)

    # ---- 1. sharded + all-reduce == sequential
    arena = GaussianArena(scene, dev, pad_to=4 * world)
    stats = [torch.zeros(n, 2, device=dev), torch.zeros(n, 1, device=dev), torch.zeros(n, device=dev)]
    # (the bench's schedule: projection backward of all views deferred, SH block reduced under it)
    deferred = []
    run_views(arena, shard_views(n_views, rank, world), stats, deferred)
    exchange = TwoPhaseAllReduce(arena)
    exchange.begin_sh()
    for fn in deferred:
        fn()
    exchange.finish()
    allreduce_densification_stats(*stats)
    seq = GaussianArena(scene, dev, pad_to=4 * world)
    seq_stats = [torch.zeros(n, 2, device=dev), torch.zeros(n, 1, device=dev), torch.zeros(n, device=dev)]
    run_views(seq, range(n_views), seq_stats)
    ok = True
    for name, off, numel, _ in arena.layout:
        a, b = arena.grads[off:off + numel].double(), seq.grads[off:off + numel].double()
        err = float((a - b).abs().max() / (b.abs().max() + 1e-30))
        bad = int(((a - b).abs() > GRAD_RTOL * b.abs() + 1e-6 * b.abs().max()).sum())
        if rank == 0:
            print(f"grad {name:10s} max-rel {err:.2e}, {bad} of {numel} entries miss the per-element bar")
        ok &= err <= GRAD_RTOL and bad <= max(4, 1e-4 * numel)
    for s_a, s_b, nm in zip(stats, seq_stats, ("xyz_gradient_accum", "denom", "max_radii2D")):
        ok &= bool(torch.allclose(s_a, s_b, rtol=1e-4, atol=1e-7))
        if rank == 0:
            print(f"stats {nm}: max diff {float((s_a - s_b).abs().max()):.2e}")
    # ---- 2. identical densification decisions on every rank
    flags = torch.empty(4 * n, dtype=torch.int32, device=dev)
    nsel = torch.zeros(1, dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    check(lib.sc_densify_classify(n, stats[0].data_ptr(), stats[1].data_ptr(), arena.scales.data_ptr(),
                                  arena.opacities.data_ptr(), 2e-4, 0.3, 0.3, 0.005, 0, 3.0, 1.6, 0, flags.data_ptr(),
                                  nsel.data_ptr(), st), "sc_densify_classify")
    allf = [torch.empty_like(flags) for _ in range(world)]
    dist.all_gather(allf, flags)
    same = all(bool(torch.equal(allf[0], f)) for f in allf)
    if rank == 0:
        print("densification flags identical on all ranks:", same, "| flagged:", int(flags.sum()))
    ok &= same and int(flags.sum()) > 0
    # ---- 3. reduce-scatter + sharded Adam + all-gather == all-reduce + torch Adam
    lrs = {"means": 1.6e-4, "quats": 1e-3, "scales": 5e-3, "opacities": 5e-2, "shs": 2.5e-3}
    sharded = GaussianArena(scene, dev, pad_to=4 * world)
    opt = ShardedAdam(sharded, lrs)
    ref_p = {name: getattr(seq, name).detach().clone().requires_grad_(True) for name, *_ in seq.layout}
    ref_opt = torch.optim.Adam([{"params": [ref_p[k]], "lr": lrs[k]} for k in ref_p], lr=0.0, eps=1e-15)
    for it in range(3):
        deferred = [] if it != 0 else None
        run_views(sharded, shard_views(n_views, rank, world), None, deferred)
        if deferred is not None:
            opt.begin_sh_exchange()
            for fn in deferred:
                fn()
        opt.step(async_gather=(it == 1))
        opt.wait()
        run_views(seq, range(n_views))
        for name, off, numel, shape in seq.layout:
            ref_p[name].grad = seq.grads[off:off + numel].view(shape).clone()
        ref_opt.step()
        with torch.no_grad():
            for name in ref_p:
                getattr(seq, name).copy_(ref_p[name])
    for name, off, numel, shape in sharded.layout:
        a, b = sharded.params[off:off + numel], ref_p[name].detach().reshape(-1)
        d = float((a - b).abs().max())
        moved = float((b - scene[name].reshape(-1).to(dev)).abs().max())
        if rank == 0:
            print(f"adam {name:10s} max |param diff| {d:.2e} (parameters moved by up to {moved:.2e})")
        ok &= d <= 2e-3 * moved + 1e-6
    # ---- 4. the step as TWO CUDA graphs with the SH-block all-reduce started between them (bench.py's N > 1 schedule)
    from street_crafter_b200.pipeline import GraphedStep, ViewPipeline
    garena = GaussianArena(scene, dev, pad_to=4 * world)
    gsinks = garena.sinks()
    mine = list(shard_views(n_views, rank, world))
    gviews = []
    for v in mine:
        Vm, Km = cams[v]
        gviews.append((Vm[None].to(dev), Km[None].to(dev), (-(Vm[:3, :3].T @ Vm[:3, 3])).to(dev)))

    def gbwd(i, out):
        torch.autograd.backward([out["render_colors"], out["render_alphas"]], list(ups[mine[i]]))

    gstep = GraphedStep(ViewPipeline(dev), garena.zero_grad,
                        lambda: (garena.means, garena.shs, garena.scales, garena.quats, garena.opacities), gviews, W, H, deg,
                        lambda: gsinks, gbwd, split_projection_backward=True)
    gex = TwoPhaseAllReduce(garena)
    seq = GaussianArena(scene, dev, pad_to=4 * world)   # (the Adam check above moved the old one's parameters)
    run_views(seq, range(n_views))
    for _ in range(2):
        gstep.replay_main()
        gex.begin_sh()
        gstep.replay_tail()
        gex.finish()
        torch.cuda.synchronize()
        ok &= gstep.overflowed() == []
        for name, off, numel, _ in garena.layout:
            a, b = garena.grads[off:off + numel].double(), seq.grads[off:off + numel].double()
            err = float((a - b).abs().max() / (b.abs().max() + 1e-30))
            ok &= err <= GRAD_RTOL
            if rank == 0:
                print(f"graph {name:10s} max-rel {err:.2e}")
    del gstep
    t = torch.tensor([1.0 if ok else 0.0], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("nccl parity ok" if t.item() == 1.0 else "nccl parity FAILED")
    sys.exit(0 if t.item() == 1.0 else 1)


if __name__ == "__main__":
    main()
