"""ctypes binding of libscb200.so (the C ABI declared in include/scb200.h).

There is NO fallback: if the shared library cannot be loaded the import fails loudly, and
every call raises RuntimeError on a non-zero status.  The library is built in-tree by
`python -m street_crafter_b200.build` (nvcc, sm_100a)."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# SCB200_LIB (tuning sweeps only): load a pre-built variant of the library instead
LIB_PATH = os.environ.get("SCB200_LIB") or os.path.join(_HERE, "libscb200.so")

_i64, _i32, _f32, _ptr = C.c_int64, C.c_int, C.c_float, C.c_void_p

# name -> (restype, argtypes); mirrors include/scb200.h one to one
SIGNATURES = {
    "sc_last_error": (C.c_char_p, []),
    "sc_version": (_i32, []),
    "sc_launch_count": (_i64, []),
    "sc_projection_fwd": (_i32, [_i64, _i32, _ptr, _ptr, _ptr, _ptr, _ptr, _i32, _i32, _f32, _f32, _f32,
                                 _f32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "sc_projection_bwd": (_i32, [_i64, _i32, _ptr, _ptr, _ptr, _ptr, _ptr, _i32, _i32, _f32, _ptr, _ptr,
                                 _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _i32, _ptr]),
    "sc_sh_fwd": (_i32, [_i64, _i32, _i32, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "sc_sh_bwd": (_i32, [_i64, _i32, _i32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _i32, _ptr]),
    "sc_projection_fwd_opacity": (_i32, [_i64, _i32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _i32, _i32, _f32, _f32,
                                         _f32, _f32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "sc_compact_rows": (_i32, [_i64, _ptr, _ptr, _i32, _ptr, _ptr, _ptr, _ptr]),
    "sc_densify_classify": (_i32, [_i64, _ptr, _ptr, _ptr, _ptr, _f32, _f32, _f32, _f32, _i32, _f32, _f32, _i32, _ptr, _ptr,
                                   _ptr]),
    "sc_densify_move": (_i32, [_i64, _ptr, _ptr, _i64, _i64, _i64, _i64, _i32, _ptr, _ptr, _ptr, _f32, _ptr, _i32, _ptr, _ptr,
                               _ptr, _ptr, _ptr]),
    "sc_assemble_fwd": (_i32, [_i32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _i32, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "sc_assemble_bwd": (_i32, [_i32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _i32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr,
                               _ptr]),
    "sc_assemble_act_fwd": (_i32, [_i32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "sc_assemble_act_bwd": (_i32, [_i32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "sc_assemble_features_fwd": (_i32, [_i32, _ptr, _i32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "sc_assemble_features_bwd": (_i32, [_i32, _ptr, _i32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "sc_densification_stats": (_i32, [_i64, _ptr, _ptr, _ptr, _ptr, _i32, _i32, _ptr, _ptr, _ptr, _ptr]),
    "sc_rasterize_fwd_depth": (_i32, [_i32, _i64, _i64, _i32, _i32, _i32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr,
                                      _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "sc_rasterize_limit_warps": (_i32, [_i32]),
    "sc_rasterize_fwd_depth_dev": (_i32, [_i32, _i64, _i64, _ptr, _i32, _i32, _i32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr,
                                          _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "sc_projection_bwd_opacity": (_i32, [_i64, _i32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _i32, _i32, _f32, _ptr,
                                         _ptr, _ptr, _i32, _ptr, _ptr, _i32, _ptr, _ptr, _ptr, _ptr, _i32, _ptr]),
    "sc_sh_fwd_view": (_i32, [_i64, _i32, _i32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "sc_sh_bwd_view": (_i32, [_i64, _i32, _i32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _i32, _ptr]),
    "sc_isect_count": (_i32, [_i64, _ptr, _ptr, _ptr, _i32, _i32, _i32, _ptr, _ptr, _ptr, _ptr]),
    "sc_isect_gather_counts": (_i32, [_i64, _ptr, _ptr, _ptr, _ptr]),
    "sc_scan_workspace_bytes": (_i64, [_i64]),
    "sc_exclusive_scan_i32": (_i32, [_i64, _ptr, _ptr, _ptr, _ptr, _i64, _ptr]),
    "sc_exclusive_scan_i32_noalloc_max_n": (_i64, []),
    "sc_exclusive_scan_i32_noalloc": (_i32, [_i64, _ptr, _ptr, _ptr, _ptr]),
    "sc_isect_emit": (_i32, [_i64, _i32, _ptr, _ptr, _ptr, _ptr, _ptr, _i32, _i32, _i32, _i32, _i64, _ptr,
                             _ptr, _ptr]),
    "sc_sort_workspace_bytes": (_i64, [_i64]),
    "sc_radix_sort_pairs": (_i32, [_i64, _ptr, _ptr, _ptr, _ptr, _i32, _i32, _ptr, _i64, _ptr]),
    "sc_radix_sort_pairs_u32": (_i32, [_i64, _ptr, _ptr, _ptr, _ptr, _i32, _i32, _ptr, _i64, _ptr]),
    "sc_isect_workspace_bytes": (_i64, [_i64, _i32, _i32]),
    "sc_isect_prepare": (_i32, [_i64, _ptr, _ptr, _ptr, _i32, _i32, _i32, _ptr, _ptr, _i64, _ptr]),
    "sc_isect_emit_sorted": (_i32, [_i64, _i32, _ptr, _ptr, _ptr, _i32, _i32, _i32, _i32, _i32, _i64, _i64,
                                    _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _i64, _ptr]),
    "sc_isect_offset_encode": (_i32, [_i64, _ptr, _i32, _i32, _i32, _i32, _ptr, _ptr]),
    "sc_bin_supported": (_i32, [_i32, _i32, _i32]),
    "sc_bin_reserve_sms": (_i32, [_i32]),
    "sc_bin_prepare": (_i32, [_i64, _ptr, _ptr, _ptr, _i32, _i32, _i32, _ptr, _ptr, _i64, _ptr]),
    "sc_bin_sort_workspace_bytes": (_i64, [_i64]),
    "sc_bin_phase_stamps_offset": (_i64, [_i64, _i32, _i32]),
    "sc_bin_finish": (_i32, [_i64, _ptr, _i32, _i32, _i64, _i64, _i64, _ptr, _ptr, _ptr, _ptr, _ptr, _i64, _ptr]),
    "sc_bin_finish_dev": (_i32, [_i64, _ptr, _i32, _i32, _i64, _ptr, _ptr, _ptr, _ptr, _ptr, _i64, _ptr]),
    "sc_rasterize_fwd": (_i32, [_i32, _i64, _i64, _i32, _i32, _i32, _i32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr,
                                _ptr, _i32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "sc_rasterize_bwd": (_i32, [_i32, _i64, _i64, _i32, _i32, _i32, _i32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr,
                                _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "sc_photometric_workspace_bytes": (_i64, [_i32, _i32]),
    "sc_photometric_fwd": (_i32, [_i32, _i32, _ptr, _i64, _i64, _ptr, _ptr, _f32, _f32, _ptr, _i64, _ptr, _ptr]),
    "sc_photometric_bwd": (_i32, [_i32, _i32, _ptr, _i64, _i64, _ptr, _ptr, _f32, _f32, _ptr, _ptr, _ptr, _i64,
                                  _i64, _i32, _ptr]),
    "sc_knn_workspace_bytes": (_i64, [_i64]),
    "sc_knn3_mean_dist2": (_i32, [_i64, _ptr, _ptr, _ptr, _i64, _ptr]),
    "sc_lidar_segment_bytes": (_i64, []),
    "sc_lidar_gather": (_i32, [_i64, _i32, _ptr, _ptr, _f32, _f32, _i32, _i32, _i32, _ptr, _ptr, _ptr, _ptr]),
    "sc_lidar_pack": (_i32, [_i64, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "sc_adam_step": (_i32, [_i32, _ptr, _ptr, _i64, _i64, _ptr, _ptr, _ptr, _ptr, C.c_double, C.c_double, C.c_double,
                            _i64, _ptr]),
    "sc_ply_row_width": (_i32, [_i32, _i32]),
    "sc_ply_pack": (_i32, [_i64, _i32, _i32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "sc_ply_unpack": (_i32, [_i64, _i32, _i32, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "sc_probe_fp32": (_i32, [_i32, _i32, _i32, _i32, _ptr, _ptr]),
}


def _bind(path: str):
    lib = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError when the build is stale
        fn.restype = res
        fn.argtypes = args
    return lib


def _sources_newer_than(lib_path: str) -> bool:
    t = os.path.getmtime(lib_path)
    csrc = os.path.join(_HERE, "csrc")
    deps = [os.path.join(csrc, f) for f in os.listdir(csrc)] + [os.path.join(_HERE, "..", "include", "scb200.h")]
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def _load() -> C.CDLL:
    """dlopen the in-tree library, (re)building it first with nvcc when it is missing or older
    than its sources (nvcc cross-compiles sm_100a without a GPU).  Never a CPU fallback."""
    from .build import build
    if not os.environ.get("SCB200_LIB"):
        try:
            build()  # incremental: a no-op when libscb200.so is newer than every source
        except Exception as e:
            # a box without nvcc may still carry an up-to-date library; a library OLDER than its sources
            # must never be bound silently (tests / bench would exercise stale kernels)
            if not os.path.exists(LIB_PATH) or _sources_newer_than(LIB_PATH):
                raise
            import warnings
            warnings.warn(f"street_crafter_b200: rebuild skipped ({type(e).__name__}: {e}); "
                          f"{LIB_PATH} is up to date with its sources")
    try:
        return _bind(LIB_PATH)
    except AttributeError as e:
        raise ImportError(f"{LIB_PATH} does not export every symbol of include/scb200.h; rebuild "
                          f"with `python -m street_crafter_b200.build --force`") from e


lib = _load()


def check(status: int, what: str) -> int:
    if status < 0:
        msg = lib.sc_last_error()
        raise RuntimeError(f"{what} failed (status {status}): {msg.decode() if msg else ''}")
    return status


def launch_count() -> int:
    return int(lib.sc_launch_count())
