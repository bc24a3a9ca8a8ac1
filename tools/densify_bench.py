"""Times the densification cycle (densify_and_prune: clone + split + prune) on the cfg2 Gaussian set: 2 M rows of
(xyz 3 | f_dc 3 | f_rest 45 | opacity 1 | scaling 3 | rotation 4) floats for the parameters and both Adam moments
(18 arrays, 708 B / Gaussian) against the pinned oracle's torch op chain on the same device tensors (what the
reference runs).  Algorithmic bytes: 28 B classify read per row + every input float read once + every output
float written once.       python tools/densify_bench.py"""
import math, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle.densify_cycle_torch import densify_and_prune as oracle_cycle      # the reference-side timing only
from street_crafter_b200.densify import densify_and_prune
dev = torch.device("cuda:0")
N = 2_000_000
g = torch.Generator().manual_seed(0)
extent, max_grad, min_op = 20.0, 0.0002, 0.005
shapes = {"xyz": (3,), "f_dc": (1, 3), "f_rest": (15, 3), "opacity": (1,), "scaling": (3,), "rotation": (4,)}
P = {k: torch.randn((N,) + s, generator=g).to(dev) for k, s in shapes.items()}
# a training-like mix: ~4 % cloned, ~4 % split, ~6 % pruned
P["scaling"] = (math.log(0.01 * extent) + 0.8 * torch.randn(N, 3, generator=g)).to(dev)
P["opacity"] = (2.0 * torch.randn(N, 1, generator=g) + 1.0).to(dev)
M = {k: torch.randn_like(t) for k, t in P.items()}
V = {k: torch.rand_like(t) for k, t in P.items()}
denom = torch.randint(1, 4, (N, 1), generator=g).float().to(dev)
accum = (torch.rand(N, 2, generator=g) * 1.18 * max_grad).to(dev) * denom
gen = torch.Generator(device=dev)
def t_ms(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
out = densify_and_prune(P, M, V, accum, denom, max_grad, min_op, extent, 20, generator=gen.manual_seed(1))
sc, n_out = out[4], out[0]["xyz"].shape[0]
noise = torch.randn((2 * sc["points_split"], 3), device=dev)
ours = t_ms(lambda: densify_and_prune(P, M, V, accum, denom, max_grad, min_op, extent, 20, noise=noise))
ref = t_ms(lambda: oracle_cycle(P, M, V, accum, denom, noise, max_grad, min_op, extent, 20))
row = sum(t[0].numel() for t in P.values()) * 4 * 3
moved = N * (28 + 16 + 16 + 32) + N * row + n_out * row
print(f"scalars {sc} -> {n_out} rows")
print(f"densify_and_prune: {ours:.3f} ms ({moved/ours/1e6:.0f} GB/s algorithmic, incl. the scan and the one host "
      f"read-back); torch op chain of the reference: {ref:.3f} ms ({ref/ours:.1f}x)")
