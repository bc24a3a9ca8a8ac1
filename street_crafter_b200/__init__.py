"""street_crafter_b200 — B200-native (sm_100a) Gaussian-splatting hot path behind the call
surface StreetCrafter uses (gsplat.rendering ops, simple_knn distCUDA2, PointRasterizer).

Importing this package loads libscb200.so; there is no CPU or library fallback."""
from . import _lib  # noqa: F401  (fails loudly when the CUDA library is missing)
from .knn import distCUDA2
from .point_rasterizer import PointRasterizationSettings, PointRasterizer
from .rendering import (fully_fused_projection, isect_offset_encode, isect_tiles, rasterization,
                        rasterize_to_pixels, spherical_harmonics)

__all__ = ["rasterization", "fully_fused_projection", "isect_tiles", "isect_offset_encode",
           "rasterize_to_pixels", "spherical_harmonics", "distCUDA2", "PointRasterizationSettings",
           "PointRasterizer"]
