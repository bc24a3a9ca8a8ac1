"""Times the stand-alone radix sorter over n (fixed cost vs per-element cost).
    python tools/sort_bench.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from street_crafter_b200._lib import check, lib

dev = torch.device("cuda:0")
st = torch.cuda.current_stream().cuda_stream

def run(fn, n, keys, begin, end, reps=10):
    vals = torch.arange(n, dtype=torch.int32, device=dev)
    kb, vb = torch.empty_like(keys), torch.empty_like(vals)
    wsb = int(lib.sc_sort_workspace_bytes(n))
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    ts = []
    for _ in range(reps):
        ka = keys.clone(); va = vals.clone()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        check(fn(n, ka.data_ptr(), va.data_ptr(), kb.data_ptr(), vb.data_ptr(), begin, end, ws.data_ptr(), wsb, st), "sort")
        e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return min(ts)

g = torch.Generator(device="cpu").manual_seed(0)
print("u32 keys (depth-like), bits [0,32): 4 passes + histogram")
for n in (65536, 262144, 626000, 2_000_000, 8_000_000):
    d = (torch.rand(n, generator=g) * 90 + 0.5).view(torch.int32).to(dev)
    t = run(lib.sc_radix_sort_pairs_u32, n, d, 0, 32)
    print(f"  n={n:9d}: {t:8.1f} us  ({t/4:6.1f} us/pass, {n*16*4/t/1e3:7.1f} GB/s)")
print("u64 keys (tile<<32|depth), bits [32,46): 2 passes + histogram")
for n in (1_000_000, 4_000_000, 11_400_000, 32_000_000):
    tile = torch.randint(0, 9600, (n,), generator=g, dtype=torch.int64)
    k = ((tile << 32) | torch.randint(0, 2**31, (n,), generator=g, dtype=torch.int64)).to(dev)
    t = run(lib.sc_radix_sort_pairs, n, k, 32, 46)
    print(f"  n={n:9d}: {t:8.1f} us  ({t/2:6.1f} us/pass, {n*24*2/t/1e3:7.1f} GB/s)")
print("u32 keys (tile id), bits [0,14): 2 passes + histogram  (what a 32-bit tile sort would cost)")
for n in (4_000_000, 11_400_000, 32_000_000):
    k = torch.randint(0, 9600, (n,), generator=g, dtype=torch.int32).to(dev)
    t = run(lib.sc_radix_sort_pairs_u32, n, k, 0, 14)
    print(f"  n={n:9d}: {t:8.1f} us  ({t/2:6.1f} us/pass, {n*16*2/t/1e3:7.1f} GB/s)")
print("REAL binning keys of the bench scene (entry order): u64 [32,46) vs the same tile ids as u32 [0,14)")
from street_crafter_b200 import rendering as R
from tools.scene import make_camera, make_gaussians
import math
sc = {k: v.to(dev) for k, v in make_gaussians(2_000_000, 3, seed=0).items()}
V, K = make_camera(1920, 1280)
radii, m2, dep, con, _ = R.fully_fused_projection(sc["means"], None, sc["quats"], sc["scales"], V[None].to(dev),
                                                  K[None].to(dev), 1920, 1280)
_, ids, flat = R.isect_tiles(m2, radii, dep, 16, 120, 80, sort=False, n_cameras=1)
n = ids.numel()
t = run(lib.sc_radix_sort_pairs, n, ids, 32, 46)
print(f"  u64 n={n:9d}: {t:8.1f} us  ({t/2:6.1f} us/pass)")
k32 = (ids >> 32).to(torch.int32)
t = run(lib.sc_radix_sort_pairs_u32, n, k32, 0, 14)
print(f"  u32 n={n:9d}: {t:8.1f} us  ({t/2:6.1f} us/pass)")
