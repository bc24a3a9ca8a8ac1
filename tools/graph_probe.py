"""can a camera pass with the cooperative binning kernel be captured in a CUDA graph?  (probe, run on the GPU box)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from street_crafter_b200.parallel import GaussianArena
from street_crafter_b200.pipeline import GraphedStep, ViewPipeline
from street_crafter_b200.renderer import render_kernel_gsplat
from tools.scene import make_camera, make_gaussians

dev = torch.device("cuda:0")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
ahead = int(sys.argv[2]) if len(sys.argv) > 2 else 1
W, H, deg = 1920, 1280, 3
arena = GaussianArena(make_gaussians(n, sh_degree=deg, seed=0), dev)
views = []
for yaw in (-45.0, 0.0, 45.0):
    V, K = make_camera(W, H, yaw_deg=yaw)
    views.append((V[None].to(dev), K[None].to(dev), (-(V[:3, :3].T @ V[:3, 3])).to(dev)))
g = torch.Generator().manual_seed(1)
v_rc = [torch.randn(1, H, W, 4, generator=g).to(dev) for _ in views]
v_ra = [torch.randn(1, H, W, 1, generator=g).to(dev) for _ in views]
def bwd(v, out):
    torch.autograd.backward([out["render_colors"], out["render_alphas"]], [v_rc[v], v_ra[v]])
sinks = arena.sinks()
# eager sequential reference
arena.zero_grad()
for v, (V, K, C) in enumerate(views):
    out = render_kernel_gsplat(arena.means, arena.shs, arena.scales, arena.quats, arena.opacities, V, K, C, W, H, deg,
                               grad_sinks=sinks, aux_ids=False)
    bwd(v, out)
torch.cuda.synchronize()
ref = {k: t.clone() for k, t in sinks.items()}
ref_img = out["render_colors"].clone()
from street_crafter_b200._lib import lib
lib.sc_rasterize_limit_warps(int(os.environ.get("RASTER_WARPS", "0")))
pipe = ViewPipeline(dev, ahead=ahead, memory_priority=int(os.environ.get("MEM_PRIO", "-1")))
gs = GraphedStep(pipe, arena.zero_grad, lambda: (arena.means, arena.shs, arena.scales, arena.quats, arena.opacities),
                 views, W, H, deg, lambda: sinks, bwd)
print("captured; capacities", gs.capacities, flush=True)
for i in range(6):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record(); outs = gs.replay(); e1.record(); t1 = time.perf_counter()
    torch.cuda.synchronize()
    print("replay %d: host %.3f ms device %.3f ms overflow %s counts %s" % (i, (t1 - t0) * 1e3, e0.elapsed_time(e1), gs.overflowed(),
          [h.tolist() for h in gs.counts_host]), flush=True)
print("image equal:", torch.equal(outs[-1]["render_colors"], ref_img))
for k in ref:
    d = (sinks[k] - ref[k]).abs().max().item() / ref[k].abs().max().item()
    print("grad", k, "max diff / max", d)
