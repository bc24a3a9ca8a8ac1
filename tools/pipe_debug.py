"""per-step host / device times and allocator activity of pipeline.ViewPipeline on the cfg2 workload
    python tools/pipe_debug.py [ahead] [steps]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from street_crafter_b200.parallel import GaussianArena
from street_crafter_b200.pipeline import ViewPipeline
from street_crafter_b200.renderer import render_kernel_gsplat
from tools.scene import make_camera, make_gaussians

ahead = int(sys.argv[1]) if len(sys.argv) > 1 else 1
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 12
mode = sys.argv[3] if len(sys.argv) > 3 else "pipeline"
dev = torch.device("cuda:0")
W, H, deg, n = 1920, 1280, 3, 2_000_000
arena = GaussianArena(make_gaussians(n, sh_degree=deg, seed=0), dev)
sinks = arena.sinks()
views = []
for yaw in (-45.0, 0.0, 45.0):
    V, K = make_camera(W, H, yaw_deg=yaw)
    views.append((V[None].to(dev), K[None].to(dev), (-(V[:3, :3].T @ V[:3, 3])).to(dev)))
g = torch.Generator().manual_seed(1)
v_rc = [torch.randn(1, H, W, 4, generator=g).to(dev) for _ in views]
v_ra = [torch.randn(1, H, W, 1, generator=g).to(dev) for _ in views]
def bwd(v, out):
    torch.autograd.backward([out["render_colors"], out["render_alphas"]], [v_rc[v], v_ra[v]])
pipe = ViewPipeline(dev, ahead=ahead)
def step():
    arena.zero_grad()
    if mode == "pipeline":
        pipe.step(arena.means, arena.shs, arena.scales, arena.quats, arena.opacities, views, W, H, deg, sinks, bwd)
    else:
        for v, (V, K, C) in enumerate(views):
            out = render_kernel_gsplat(arena.means, arena.shs, arena.scales, arena.quats, arena.opacities, V, K, C, W, H, deg,
                                       grad_sinks=sinks, aux_ids=False)
            bwd(v, out)
for i in range(steps):
    torch.cuda.synchronize()
    s0 = torch.cuda.memory_stats(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    step()
    e1.record(); t1 = time.perf_counter()
    torch.cuda.synchronize(); t2 = time.perf_counter()
    s1 = torch.cuda.memory_stats(dev)
    print("step %2d host-enqueue %.3f ms  host-total %.3f ms  device %.3f ms  segments +%d  reserved %.0f MB  retries %d" % (
        i, (t1 - t0) * 1e3, (t2 - t0) * 1e3, e0.elapsed_time(e1), s1["segment.all.allocated"] - s0["segment.all.allocated"],
        s1["reserved_bytes.all.current"] / 1e6, s1["num_alloc_retries"]), flush=True)
