"""Phase timeline of the cooperative binning-prepare kernel (globaltimer stamps of CTA 0).
    python tools/bin_phases.py [n_gaussians]"""
import math, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from street_crafter_b200 import rendering as R
from street_crafter_b200._lib import lib, check
from tools.scene import make_camera, make_gaussians

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
W, H = 1920, 1280
dev = torch.device("cuda:0")
s = {k: v.to(dev) for k, v in make_gaussians(n, seed=0).items()}
V, K = make_camera(W, H)
radii, m2, dep, con, _ = R.fully_fused_projection(s["means"], None, s["quats"], s["scales"], V[None].to(dev), K[None].to(dev), W, H,
                                                  near_plane=0.001, far_plane=1000.0)
tw, th = math.ceil(W / 16), math.ceil(H / 16)
ws_bytes = int(lib.sc_isect_workspace_bytes(n, tw, th))
ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
tpg = torch.empty((1, n), dtype=torch.int32, device=dev)
st = torch.cuda.current_stream().cuda_stream
off = int(lib.sc_bin_phase_stamps_offset(n, tw, th))
names = ["start", "phase0 done", "sync1", "pass0 scatter done", "sync2", "pass1 count done", "sync3", "pass1 bases",
         "pass1 scatter done", "sync4", "pass3 done", "scan pass0 done", "sync9", "scan pass1 done"]
for it in range(4):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    check(lib.sc_bin_prepare(n, m2.data_ptr(), radii.data_ptr(), dep.data_ptr(), 16, tw, th, tpg.data_ptr(), ws.data_ptr(), ws_bytes, st), "prepare")
    e1.record()
    torch.cuda.synchronize()
    t = ws[off:off + 14 * 8].view(torch.int64).cpu().tolist()
    tot = ws[:24].view(torch.int64).cpu().tolist()
print("totals", tot, "event ms", e0.elapsed_time(e1))
for i in range(1, 14):
    print(f"{names[i]:22s} +{(t[i] - t[i-1]) / 1e3:7.1f} us   (t = {(t[i] - t[0]) / 1e3:7.1f})")
