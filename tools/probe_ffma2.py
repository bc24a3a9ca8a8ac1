"""FFMA vs FFMA2 (Blackwell packed FP32) issue-rate probe: python tools/probe_ffma2.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from street_crafter_b200 import _lib
dev = torch.device("cuda:0")
sink = torch.empty(148 * 8 * 256, device=dev)
def run(kind, iters=8192):
    st = torch.cuda.current_stream().cuda_stream
    _lib.check(_lib.lib.sc_probe_fp32(kind, 148 * 8, 256, 64, sink.data_ptr(), st), "probe")
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    _lib.check(_lib.lib.sc_probe_fp32(kind, 148 * 8, 256, iters, sink.data_ptr(), st), "probe")
    e1.record(); torch.cuda.synchronize()
    flop = 148 * 8 * 256 * iters * (16 if kind == 0 else 32 if kind == 2 else 8)
    return flop / (e0.elapsed_time(e1) * 1e-3) / 1e12
for k, n in ((0, "FFMA  (scalar)"), (2, "FFMA2 (fma.rn.f32x2)")):
    print(n, "%.1f TFLOP/s" % max(run(k) for _ in range(3)))
