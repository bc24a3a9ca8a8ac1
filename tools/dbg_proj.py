import sys, torch
sys.path.insert(0, '.')
from oracle import gsplat_torch as G
from tools.scene import *
import street_crafter_b200 as S
dev = torch.device('cuda:0')
n, W, H = 20000, 960, 640
s = make_gaussians(n, seed=n); V, K = make_camera(W, H)
want = G.fully_fused_projection(s["means"], None, s["quats"], s["scales"], V[None], K[None], W, H, near_plane=0.001, far_plane=1000.0, calc_compensations=True)
got = S.fully_fused_projection(s["means"].to(dev), None, s["quats"].to(dev), s["scales"].to(dev), V[None].to(dev), K[None].to(dev), W, H, near_plane=0.001, far_plane=1000.0, calc_compensations=True)
for nm, g, w in zip(["radii", "means2d", "depths", "conics", "comp"], got, want):
    g = g.cpu(); bad = (g != w)
    print(nm, bad.sum().item())
    if bad.any() and g.dtype.is_floating_point:
        idx = bad.nonzero()[:8]
        for i in idx:
            i = tuple(i.tolist())
            print(i, g[i].item(), w[i].item(), (g[i].view(torch.int32) - w[i].view(torch.int32)).item())
