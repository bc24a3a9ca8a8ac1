"""Times sc_sh_bwd variants on 2M Gaussians with ~31% random visibility."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from street_crafter_b200._lib import lib, check
dev = torch.device("cuda:0")
M, K = 2_000_000, 16
g = torch.Generator().manual_seed(0)
dirs = torch.randn(M, 3, generator=g).to(dev)
coeffs = torch.randn(M, K, 3, generator=g).to(dev)
vcol = torch.randn(M, 3, generator=g).to(dev)
st = torch.cuda.current_stream().cuda_stream
for frac in (0.31, 1.0):
    mask = (torch.rand(M, generator=g) < frac).to(torch.uint8).to(dev)
    for acc in (0, 1):
        for with_dirs in (0, 1):
            vco = torch.zeros(M, K, 3, device=dev)
            vd = torch.zeros(M, 3, device=dev)
            ts = []
            for it in range(6):
                e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
                e0.record()
                check(lib.sc_sh_bwd(M, K, 3, dirs.data_ptr(), coeffs.data_ptr(), mask.data_ptr(), vcol.data_ptr(),
                                    vco.data_ptr(), vd.data_ptr() if with_dirs else None, acc, st), "sh")
                e1.record(); torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            print(f"vis {frac:.2f} accumulate {acc} v_dirs {with_dirs}: {min(ts)*1e3:.1f} us")
