"""Host-side logic of the view-sharded multi-GPU step (street_crafter_b200/parallel.py) on CPU:
world-size-2 gloo.  The renderer is replaced by a differentiable stand-in of the same
signature (the CUDA kernels need a GPU); what is checked is the plumbing — arena layout, view
assignment, .grad views, and that ONE all-reduce over the flat gradient arena reproduces the
single-process gradients."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from street_crafter_b200.parallel import (GaussianArena, allreduce_densification_stats, allreduce_gradients,
                                          arena_layout, shard_views)
from tools.scene import make_gaussians


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _fake_render(arena, view):
    """differentiable stand-in touching every parameter, view-dependent"""
    w = 1.0 + 0.25 * view
    return ((arena.means * w).sin().sum() + (arena.quats ** 2).sum() * w + arena.scales.exp().sum()
            + (arena.opacities * w).sum() + (arena.shs * (w + 0.5)).pow(2).sum())


def _single_process_grads(scene, n_views):
    arena = GaussianArena(scene, torch.device("cpu"))
    arena.zero_grad()
    for v in range(n_views):
        _fake_render(arena, v).backward()
    return arena.grads.clone()


def _worker(rank, world, port, n_views, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    scene = make_gaussians(257, sh_degree=3, seed=0)
    arena = GaussianArena(scene, torch.device("cpu"))
    arena.zero_grad()
    for v in shard_views(n_views, rank, world):
        _fake_render(arena, v).backward()
    # every parameter's .grad must still be a view of the flat arena (no pack step)
    for name, off, numel, shape in arena.layout:
        p = getattr(arena, name)
        assert p.grad.data_ptr() == arena.grads.data_ptr() + 4 * off, name
    allreduce_gradients(arena)
    acc = torch.full((257, 2), float(rank + 1))
    den = torch.full((257, 1), 1.0)
    mx = torch.full((257,), float(rank))
    allreduce_densification_stats(acc, den, mx)
    torch.save({"grads": arena.grads, "acc": acc, "den": den, "mx": mx}, os.path.join(out_dir, f"r{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_arena_layout_alignment():
    layout, total = arena_layout(1001, 16)
    names = [l[0] for l in layout]
    assert names == ["means", "quats", "scales", "opacities", "shs"]
    for name, off, numel, shape in layout:
        assert off % 4 == 0, name  # 16-byte aligned slices (float4 kernels)
    assert total >= 1001 * 59


def test_shard_views_round_robin():
    assert shard_views(40, 0, 8) == [0, 8, 16, 24, 32]
    assert shard_views(40, 7, 8) == [7, 15, 23, 31, 39]
    seen = sorted(v for r in range(3) for v in shard_views(10, r, 3))
    assert seen == list(range(10))


def test_view_sharded_allreduce_world2(tmp_path):
    n_views, world = 5, 2
    want = _single_process_grads(make_gaussians(257, sh_degree=3, seed=0), n_views)
    mp.spawn(_worker, args=(world, _free_port(), n_views, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        got = torch.load(os.path.join(str(tmp_path), f"r{r}.pt"))
        assert torch.allclose(got["grads"], want, rtol=1e-5, atol=1e-6), f"rank {r}"
        assert torch.equal(got["acc"], torch.full((257, 2), 3.0))     # SUM of 1 and 2
        assert torch.equal(got["den"], torch.full((257, 1), 2.0))
        assert torch.equal(got["mx"], torch.full((257,), 1.0))        # MAX
