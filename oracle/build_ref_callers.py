"""TEST INFRASTRUCTURE (checker, not product).  Compiles the reference's own CALLERS of the hot path
— the functions that sit directly above the drop-in boundary — from the sources where they lie under
/root/reference into oracle/_ref/ref_callers.bin, a marshalled table of Python code objects (git-ignored,
travels to the GPU box with gpurun exactly like oracle/_ref/libsimpleknn_ref.so).  Nothing is copied into
this repository: the function bodies are cut out of the reference files with `ast`, compiled unmodified
(only decorators that need the original module are dropped, listed per entry) and stored as bytecode.

    python oracle/build_ref_callers.py          (run by __graft_entry__.build() when /root/reference exists)

tests/test_gpu_boundary.py loads the table, binds the functions to stand-in `cfg` / camera / model
objects, points `gsplat`, `simple_knn` and `diff_point_rasterization` at this repository's shim packages
and executes them on the GPU: the drop-in claim is checked through the reference's own code."""
from __future__ import annotations

import ast
import hashlib
import json
import marshal
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
OUT = os.path.join(HERE, "_ref", "ref_callers.bin")
META = os.path.join(HERE, "_ref", "ref_callers.json")

# (alias, file, class or None, function, decorators to keep)
ENTRIES = [
    # the renderer's gsplat call sequence (rows A1-A8, A11)
    ("render_kernel_gsplat", "street_gaussian/models/street_gaussian_renderer.py", "StreetGaussianRenderer",
     "render_kernel_gsplat", True),
    # LiDAR condition render through diff_point_rasterization (row A9)
    ("get_device", "data_processor/utils/render_utils.py", None, "get_device", True),
    ("render_pointcloud_diff_point_rasterization", "data_processor/utils/render_utils.py", None,
     "render_pointcloud_diff_point_rasterization", True),
    # scale initialisation through simple_knn (row A10)
    ("create_from_pcd", "street_gaussian/models/gaussian_model.py", "GaussianModel", "create_from_pcd", True),
    # the consumers of means2d.grad / .absgrad / radii after backward (row A8 side effect)
    ("set_max_radii2D", "street_gaussian/models/street_gaussian_model.py", "StreetGaussianModel",
     "set_max_radii2D", True),
    ("add_densification_stats", "street_gaussian/models/street_gaussian_model.py", "StreetGaussianModel",
     "add_densification_stats", True),
    # camera helpers the stand-ins are built with (the reference's own conventions)
    ("getProjectionMatrixK", "street_gaussian/utils/graphics_utils.py", None, "getProjectionMatrixK", True),
    ("getWorld2View2", "street_gaussian/utils/graphics_utils.py", None, "getWorld2View2", True),
    ("focal2fov", "street_gaussian/utils/graphics_utils.py", None, "focal2fov", True),
    ("RGB2SH", "street_gaussian/utils/sh_utils.py", None, "RGB2SH", True),
    ("inverse_sigmoid", "street_gaussian/utils/general_utils.py", None, "inverse_sigmoid", True),
    # checkpoint / PLY formats and the optimizer set-up (row N4)
    ("gm_state_dict", "street_gaussian/models/gaussian_model.py", "GaussianModel", "state_dict", True),
    ("gm_load_state_dict", "street_gaussian/models/gaussian_model.py", "GaussianModel", "load_state_dict", True),
    ("gm_training_setup", "street_gaussian/models/gaussian_model.py", "GaussianModel", "training_setup", True),
    ("gm_make_ply", "street_gaussian/models/gaussian_model.py", "GaussianModel", "make_ply", True),
    ("gm_load_ply", "street_gaussian/models/gaussian_model.py", "GaussianModel", "load_ply", True),
    ("gm_construct_list_of_attributes", "street_gaussian/models/gaussian_model.py", "GaussianModel",
     "construct_list_of_attributes", True),
    # LiDAR condition pipeline (row N1): aggregation of +-delta frames, actor posing
    ("make_lidar_ply", "street_gaussian/pointcloud_processor/base_processor.py", "BasePointCloudProcessor",
     "make_lidar_ply", True),
    ("transform_lidar_ply", "street_gaussian/pointcloud_processor/base_processor.py", "BasePointCloudProcessor",
     "transform_lidar_ply", True),
    ("render_condition", "street_gaussian/pointcloud_processor/waymo_processor.py", "WaymoPointCloudProcessor",
     "render_condition", True),
]


def _find(tree: ast.Module, cls: str | None, fn: str) -> ast.FunctionDef:
    body = tree.body
    if cls is not None:
        body = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == cls).body
    hits = [n for n in body if isinstance(n, ast.FunctionDef) and n.name == fn]
    if len(hits) != 1:
        raise KeyError(f"{cls}.{fn}: {len(hits)} definitions")
    return hits[0]


def extract(ref_root: str = REF):
    """-> ({alias: code object of a module holding that one def}, meta)"""
    table, meta = {}, {"python": sys.version.split()[0], "entries": {}}
    cache = {}
    for alias, rel, cls, fn, keep_deco in ENTRIES:
        path = os.path.join(ref_root, rel)
        if not os.path.exists(path):
            continue
        if path not in cache:
            src = open(path).read()
            cache[path] = (src, ast.parse(src))
        src, tree = cache[path]
        try:
            node = _find(tree, cls, fn)
        except (KeyError, StopIteration):
            continue
        seg = ast.get_source_segment(src, node) or ""
        if not keep_deco:
            node.decorator_list = []
        node.name = alias
        mod = ast.Module(body=[node], type_ignores=[])
        table[alias] = compile(mod, path, "exec")
        meta["entries"][alias] = {"file": rel, "class": cls, "function": fn, "lines": [node.lineno, node.end_lineno],
                                  "sha256": hashlib.sha256(seg.encode()).hexdigest()}
    return table, meta


def build(force: bool = False) -> str | None:
    if not os.path.isdir(REF):
        return OUT if os.path.exists(OUT) else None
    table, meta = extract()
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    blob = marshal.dumps(table)
    if force or not os.path.exists(OUT) or open(OUT, "rb").read() != blob:
        with open(OUT, "wb") as f:
            f.write(blob)
    with open(META, "w") as f:
        json.dump(meta, f, indent=1)
    return OUT


def load():
    """-> ({alias: code}, meta) from oracle/_ref, or (None, reason)"""
    if not (os.path.exists(OUT) and os.path.exists(META)):
        return None, "oracle/_ref/ref_callers.bin not built (run __graft_entry__.build() where /root/reference exists)"
    meta = json.load(open(META))
    if meta.get("python") != sys.version.split()[0]:
        return None, f"ref_callers.bin was compiled by python {meta.get('python')}, this is {sys.version.split()[0]}"
    return marshal.loads(open(OUT, "rb").read()), meta


if __name__ == "__main__":
    p = build(force="--force" in sys.argv)
    print(p)
