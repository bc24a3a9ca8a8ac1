"""Times the SH kernels in the renderer's view mode on 2M Gaussians with ~31% random visibility
(accumulating backward, as the training step runs them) and the stand-alone sc_sh_bwd variants."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from street_crafter_b200._lib import lib, check
dev = torch.device("cuda:0")
M, K = 2_000_000, 16
g = torch.Generator().manual_seed(0)
means = (torch.randn(M, 3, generator=g) * 20).to(dev)
campos = torch.zeros(3, device=dev)
coeffs = torch.randn(M, K, 3, generator=g).to(dev)
vcol4 = torch.randn(M, 4, generator=g).to(dev)
depths = torch.rand(M, generator=g).to(dev)
st = torch.cuda.current_stream().cuda_stream
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

def best(fn, reps=6):
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return min(ts)

for frac in (0.31, 1.0):
    radii = (torch.rand(M, generator=g) < frac).to(torch.int32).to(dev) * 7
    col4 = torch.empty(M, 4, device=dev)
    g_shs = torch.zeros(M, K, 3, device=dev)
    g_means = torch.zeros(M, 3, device=dev)
    tf = best(lambda: check(lib.sc_sh_fwd_view(M, K, 3, means.data_ptr(), campos.data_ptr(), coeffs.data_ptr(),
                                               radii.data_ptr(), depths.data_ptr(), col4.data_ptr(), st), "f"))
    tb = best(lambda: check(lib.sc_sh_bwd_view(M, K, 3, means.data_ptr(), campos.data_ptr(), coeffs.data_ptr(),
                                               radii.data_ptr(), col4.data_ptr(), vcol4.data_ptr(), g_shs.data_ptr(),
                                               g_means.data_ptr(), 1, st), "b"))
    print(f"view mode vis {frac:.2f}: fwd {tf:7.1f} us   bwd(accumulate) {tb:7.1f} us")
