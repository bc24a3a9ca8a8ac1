"""Times the fused photometric loss (fwd + bwd) on one 1920x1280 view.   python tools/loss_bench.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from street_crafter_b200.losses import photometric_loss_render
dev = torch.device("cuda:0")
H, W = 1280, 1920
rc = torch.rand(1, H, W, 4, device=dev).requires_grad_(True)
gt = torch.rand(3, H, W, device=dev)
def run():
    rc.grad = None
    loss = photometric_loss_render(rc, gt, None, 0.2, 1.0)[0]
    loss.backward()
for _ in range(3): run()
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(5): run()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=8, max_name_column_width=60))
