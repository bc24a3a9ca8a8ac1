"""GPU timeline of one bench step (torch.profiler / CUPTI): kernel time, idle gaps and who
precedes them.   python tools/timeline.py [steps] [e2e]   (e2e: host inputs + the fused loss)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from torch.profiler import profile, ProfilerActivity
from street_crafter_b200.parallel import GaussianArena
from street_crafter_b200.renderer import render_kernel_gsplat
from tools.scene import make_camera, make_gaussians

dev = torch.device("cuda:0")
W, H, deg, n = 1920, 1280, 3, 2_000_000
arena = GaussianArena(make_gaussians(n, sh_degree=deg, seed=0), dev)
sinks = arena.sinks()
cams = []
g = torch.Generator().manual_seed(1234)
for yaw in (0.0, 45.0, -45.0):
    V, K = make_camera(W, H, yaw_deg=yaw)
    cams.append((V[None].to(dev), K[None].to(dev), (-(V[:3, :3].T @ V[:3, 3])).to(dev),
                 torch.randn(1, H, W, 4, generator=g).to(dev), torch.randn(1, H, W, 1, generator=g).to(dev)))

def step():
    arena.zero_grad()
    for V, K, cc, a, b in cams:
        out = render_kernel_gsplat(arena.means, arena.shs, arena.scales, arena.quats, arena.opacities, V, K, cc,
                                   W, H, deg, grad_sinks=sinks)
        torch.autograd.backward([out["render_colors"], out["render_alphas"]], [a, b])

if len(sys.argv) > 2 and sys.argv[2] == "e2e":
    from street_crafter_b200.losses import photometric_loss_render
    from street_crafter_b200.parallel import HostViewFeeder
    feeder = HostViewFeeder(dev)
    host = [(c[0].cpu().pin_memory(), c[1].cpu().pin_memory(), c[2].cpu().pin_memory(),
             torch.rand(3, H, W).pin_memory()) for c in cams]

    def step():
        arena.zero_grad()
        staged = [feeder.stage(h) for h in host]
        losses = []
        for tensors, ready in staged:
            feeder.acquire(tensors, ready)
            Vd, Kd, Cd, gt = tensors
            out = render_kernel_gsplat(arena.means, arena.shs, arena.scales, arena.quats, arena.opacities, Vd, Kd,
                                       Cd, W, H, deg, grad_sinks=sinks)
            loss = photometric_loss_render(out["render_colors"], gt, None, 0.2, 1.0)[0]
            loss.backward()
            losses.append(loss.detach())
        return torch.stack(losses).cpu()

for _ in range(3):
    step()
torch.cuda.synchronize()
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(steps):
        step()
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
t0, t1 = evs[0].time_range.start, max(e.time_range.end for e in evs)
busy = sum(e.time_range.end - e.time_range.start for e in evs)
print(f"{steps} steps: span {(t1-t0)/1e3/steps:.3f} ms/step, kernel+memset busy {busy/1e3/steps:.3f} ms/step, "
      f"{len(evs)/steps:.0f} GPU ops/step")
import collections
agg = collections.defaultdict(lambda: [0, 0.0])
gaps = collections.defaultdict(lambda: [0, 0.0])
prev = None
for e in evs:
    agg[e.name[:70]][0] += 1; agg[e.name[:70]][1] += e.time_range.end - e.time_range.start
    if prev is not None:
        gap = e.time_range.start - prev.time_range.end
        if gap > 0:
            k = prev.name[:40] + " -> " + e.name[:40]
            gaps[k][0] += 1; gaps[k][1] += gap
    prev = e
print("-- ops (us/step)")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:40]:
    print(f"  {v[1]/steps:9.1f} us  n={v[0]/steps:5.1f}  {k}")
print("-- idle gaps by (previous -> next) (us/step)")
tot = sum(v[1] for v in gaps.values())
print(f"  total idle {tot/1e3/steps:.3f} ms/step")
for k, v in sorted(gaps.items(), key=lambda kv: -kv[1][1])[:30]:
    print(f"  {v[1]/steps:9.1f} us  n={v[0]/steps:5.1f}  {k}")
