"""Times the compositing kernels alone on one cfg2 camera for every pixels-per-thread variant.
    python tools/raster_sweep.py [n_gaussians]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from street_crafter_b200 import rendering as R
from street_crafter_b200.renderer import render_kernel_gsplat
from tools.scene import make_camera, make_gaussians

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
W, H = 1920, 1280
dev = torch.device("cuda:0")
s = {k: v.to(dev) for k, v in make_gaussians(n, sh_degree=3, seed=0).items()}
for yaw in (0.0, 45.0):
    V, K = make_camera(W, H, yaw_deg=yaw)
    cc = (-(V[:3, :3].T @ V[:3, 3])).to(dev)
    with torch.no_grad():
        out = render_kernel_gsplat(s["means"], s["shs"], s["scales"], s["quats"], s["opacities"][:, None],
                                   V[None].to(dev), K[None].to(dev), cc, W, H, 3, train=False)
    a = out["aux"]
    g = torch.Generator().manual_seed(1)
    v_rc = torch.randn(1, H, W, 4, generator=g).to(dev)
    v_ra = torch.randn(1, H, W, 1, generator=g).to(dev)
    print(f"yaw {yaw}: I={a['isect_ids'].numel()}")
    ref = None
    for ppt in (8,):
        os.environ["SCB200_FWD_PPT"] = str(ppt)
        os.environ["SCB200_BWD_PPT"] = str(ppt)
        tf, tb = [], []
        for it in range(6):
            ins = [a[k].detach().clone().requires_grad_(True) for k in ("means2d", "conics", "colors", "opacities")]
            e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            e[0].record()
            rc, ra = R.rasterize_to_pixels(ins[0], ins[1], ins[2], ins[3], W, H, 16, a["isect_offsets"],
                                           a["flatten_ids"], absgrad=True)
            e[1].record()
            e[2].record()
            torch.autograd.backward([rc, ra], [v_rc, v_ra])
            e[3].record()
            torch.cuda.synchronize()
            tf.append(e[0].elapsed_time(e[1])); tb.append(e[2].elapsed_time(e[3]))
        grads = [t.grad for t in ins] + [ins[0].absgrad]
        if ref is None:
            ref = (rc, ra, grads)
            dmsg = ""
        else:
            dimg = (rc - ref[0]).abs().max().item()
            dg = max(((g1 - g0).abs().max() / (g0.abs().max() + 1e-30)).item() for g1, g0 in zip(grads, ref[2]))
            dmsg = f" | vs ppt1: img {dimg:.2e} grad-rel {dg:.2e}"
        print(f"  ppt {ppt}: fwd {min(tf):.3f} ms  bwd(+zero-fill) {min(tb):.3f} ms{dmsg}")
