"""Times the prune compaction on the cfg2 Gaussian set: 2 M rows of (xyz 3 | f_dc 1x3 | f_rest 15x3 |
opacity 1 | scaling 3 | rotation 4) floats — the reference's own shapes, ragged widths included — for the
parameters and both Adam moments (18 arrays, 708 B / Gaussian), 90 % kept, against the
same prune done with torch boolean indexing (what the reference does).   python tools/compact_bench.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from street_crafter_b200.compaction import compact_rows
dev = torch.device("cuda:0")
N = 2_000_000
g = torch.Generator().manual_seed(0)
keep = (torch.rand(N, generator=g) < 0.9).to(dev)
ts = [torch.randn((N,) + s, device=dev) for _ in range(3) for s in ((3,), (1, 3), (15, 3), (1,), (3,), (4,))]
def t_ms(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
ours = t_ms(lambda: compact_rows(keep, ts))
ref = t_ms(lambda: [t[keep] for t in ts])
moved = 2 * int(keep.sum()) * sum(t[0].numel() for t in ts) * 4
print(f"compact_rows: {ours:.3f} ms ({moved/ours/1e6:.0f} GB/s algorithmic, incl. scan + host read of the count); "
      f"torch boolean indexing x{len(ts)}: {ref:.3f} ms")
