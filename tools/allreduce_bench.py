"""Times the gradient-arena all-reduce (fp32 SUM, 59 floats x 2 M Gaussians = 472 MB) on N GPUs.
    torchrun --nproc-per-node N tools/allreduce_bench.py"""
import os, sys, time
import torch, torch.distributed as dist
rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
dev = torch.device("cuda", local); torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
n = 59 * 2_000_000
x = torch.randn(n, device=dev)
def t_ms(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)
ms = t_ms(lambda: dist.all_reduce(x))
if rank == 0:
    print(f"NCCL all_reduce {n*4/1e6:.0f} MB fp32, {world} GPUs: {ms:.3f} ms  busbw {2*(world-1)/world*n*4/ms/1e6:.0f} GB/s "
          f"[NCCL_ALGO={os.environ.get('NCCL_ALGO')} NCCL_NVLS_ENABLE={os.environ.get('NCCL_NVLS_ENABLE')}]", flush=True)
# chunked: 8 slices back to back (what a bucketed reduce would cost)
parts = x.chunk(8)
ms = t_ms(lambda: [dist.all_reduce(p) for p in parts])
if rank == 0:
    print(f"  same buffer as 8 chunks: {ms:.3f} ms", flush=True)
try:
    import torch.distributed._symmetric_memory as symm
    y = symm.empty(n, dtype=torch.float32, device=dev)
    y.copy_(x)
    h = symm.rendezvous(y, dist.group.WORLD.group_name)
    for name in ("multimem_all_reduce_", "two_shot_all_reduce_", "one_shot_all_reduce"):
        op = getattr(torch.ops.symm_mem, name, None)
        if op is None:
            continue
        try:
            ms = t_ms(lambda: op(y, "sum", dist.group.WORLD.group_name))
            if rank == 0:
                print(f"  symm_mem.{name}: {ms:.3f} ms  busbw {2*(world-1)/world*n*4/ms/1e6:.0f} GB/s", flush=True)
        except Exception as e:
            if rank == 0:
                print(f"  symm_mem.{name}: failed: {str(e)[:200]}", flush=True)
except Exception as e:
    if rank == 0:
        print("symmetric memory unavailable:", str(e)[:300], flush=True)
dist.destroy_process_group()
