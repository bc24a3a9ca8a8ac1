"""Host logic of the step schedules that needs no GPU: the product refuses CPU devices, the capacity check reads
the pinned counts the way the kernels write them, and the schedule modules never import the oracle."""
import ast
import os

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_view_pipeline_refuses_cpu():
    from street_crafter_b200.pipeline import ViewPipeline
    with pytest.raises(RuntimeError, match="no CPU path"):
        ViewPipeline("cpu")


def test_capacity_exceeded_reads_the_intersection_count():
    from street_crafter_b200.rendering import capacity_exceeded
    counts = torch.tensor([1000, 50_000, 7000], dtype=torch.int64)   # {n_visible, n_isects, n_items}
    assert not capacity_exceeded(counts, 50_000)
    assert capacity_exceeded(counts, 49_999)


def test_schedule_modules_do_not_import_the_oracle():
    for mod in ("pipeline.py", "rendering.py", "parallel.py", "model_arena.py", "lidar_condition.py"):
        tree = ast.parse(open(os.path.join(ROOT, "street_crafter_b200", mod)).read())
        for node in ast.walk(tree):
            names = []
            if isinstance(node, ast.Import):
                names = [a.name for a in node.names]
            elif isinstance(node, ast.ImportFrom):
                names = [node.module or ""]
            assert not any(n.split(".")[0] == "oracle" for n in names), (mod, names)


def test_header_declares_the_sync_free_and_noalloc_entries():
    h = open(os.path.join(ROOT, "include", "scb200.h")).read()
    for sym in ("sc_bin_finish_dev", "sc_rasterize_fwd_depth_dev", "sc_rasterize_limit_warps",
                "sc_exclusive_scan_i32_noalloc", "sc_exclusive_scan_i32_noalloc_max_n"):
        assert sym + "(" in h, sym
    assert "no counterpart in the reference" in h.lower() or "NO COUNTERPART IN THE REFERENCE" in h
