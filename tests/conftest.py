import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    pytest.parity_stats = []   # filled by the assertion helpers of test_gpu_parity.py


def pytest_terminal_summary(terminalreporter, exitstatus, config):
    """what the image / gradient tolerances actually allowed in this session"""
    stats = getattr(pytest, "parity_stats", [])
    if not stats:
        return
    img = [s for s in stats if s["kind"] == "image"]
    grd = [s for s in stats if s["kind"] == "grad"]
    tr = terminalreporter
    tr.write_line("parity statistics: %d image comparisons, %d with values off by > 1e-4 (total %d values, worst %.2e, "
                  "%d not on a threshold pair); %d gradient comparisons, worst max-rel %.2e, %d entries (of %d) miss the "
                  "per-element bar" % (
                      len(img), sum(1 for s in img if s["n_bad"]), sum(s["n_bad"] for s in img),
                      max([s["max_diff"] for s in img] or [0.0]), sum(s["off_margin"] or 0 for s in img),
                      len(grd), max([s["max_rel"] for s in grd] or [0.0]), sum(s["elem_bad"] for s in grd),
                      sum(s["numel"] for s in grd)))
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        import json
        with open(os.path.join(out, "parity_stats.json"), "w") as f:
            json.dump(stats, f, indent=0)


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)
