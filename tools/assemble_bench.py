"""Times the scene-graph assembly (get_xyz + get_rotation, forward + backward) on the cfg2 scene graph: 1.9 M
background Gaussians + 20 actors x 5000 (half of each flipped) against the pinned oracle's torch op chain with
autograd on the same device tensors (what the reference runs every render call).  Algorithmic bytes: 28 B read +
28 B written per Gaussian forward; 28 B upstream + 28 B parameters read + 28 B gradients written backward.
      python tools/assemble_bench.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle.assemble_torch import assemble as oracle_assemble          # the reference-side timing only
from street_crafter_b200.scene_graph import FLIP_QUAT_Y, assemble_xyz_rotation
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
leaf = lambda *s, k=1.0: (torch.randn(*s, generator=g) * k).to(dev).requires_grad_(True)
segs = [{"xyz": leaf(1_900_000, 3, k=30.0), "rotation": leaf(1_900_000, 4)}]
for a in range(20):
    segs.append({"xyz": leaf(5000, 3, k=2.0), "rotation": leaf(5000, 4), "pose_rot": leaf(4), "pose_trans": leaf(3, k=20.0),
                 "flip": (torch.rand(5000, generator=g) < 0.5).to(dev)})
N = sum(s["xyz"].shape[0] for s in segs)
w_xyz, w_rot = torch.randn(N, 3, device=dev), torch.randn(N, 4, device=dev)
fq = torch.tensor(FLIP_QUAT_Y, device=dev)
leaves = [t for s in segs for t in s.values() if t is not None and t.requires_grad]
def step(fn):
    xyz, rot = fn()
    torch.autograd.backward([xyz, rot], [w_xyz, w_rot])
    for t in leaves: t.grad = None
def t_ms(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
ours = t_ms(lambda: step(lambda: assemble_xyz_rotation(segs)))
ref = t_ms(lambda: step(lambda: oracle_assemble(segs, fq, 1)))
moved = N * (56 + 84)
print(f"assemble fwd+bwd, {N} Gaussians in {len(segs)} segments: {ours:.3f} ms ({moved/ours/1e6:.0f} GB/s algorithmic, "
      f"autograd glue included); torch op chain of the reference: {ref:.3f} ms ({ref/ours:.1f}x)")

# ---- activations and features of the same scene graph (fourier_dim 1, SH degree 3) ----
from oracle.assemble_torch import assemble_activations as oracle_act, assemble_features as oracle_feat
from street_crafter_b200.scene_graph import assemble_features, assemble_scaling_opacity
for s in segs:
    n = s["xyz"].shape[0]
    s.update(scaling=leaf(n, 3), opacity=leaf(n, 1), f_dc=leaf(n, 1, 3), f_rest=leaf(n, 15, 3))
    if "pose_rot" in s:
        s["idft_base"] = [1.0]
leaves = [t for s in segs for t in s.values() if isinstance(t, torch.Tensor) and t.requires_grad]
w_s, w_o, w_f = torch.randn(N, 3, device=dev), torch.randn(N, 1, device=dev), torch.randn(N, 16, 3, device=dev)
def step2(fn, ws):
    outs = fn()
    outs = list(outs) if isinstance(outs, tuple) else [outs]
    torch.autograd.backward(outs, ws)
    for t in leaves: t.grad = None
a_ours = t_ms(lambda: step2(lambda: assemble_scaling_opacity(segs), [w_s, w_o]))
a_ref = t_ms(lambda: step2(lambda: oracle_act(segs), [w_s, w_o]))
f_ours = t_ms(lambda: step2(lambda: assemble_features(segs), [w_f]))
f_ref = t_ms(lambda: step2(lambda: oracle_feat(segs), [w_f]))
print(f"scaling/opacity fwd+bwd: {a_ours:.3f} ms ({N*64/a_ours/1e6:.0f} GB/s algorithmic); torch op chain: {a_ref:.3f} ms")
print(f"features fwd+bwd (K=16): {f_ours:.3f} ms ({N*768/f_ours/1e6:.0f} GB/s algorithmic); torch op chain: {f_ref:.3f} ms")
