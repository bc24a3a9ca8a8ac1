"""`gsplat.rendering` shim -> street_crafter_b200.rendering (see gsplat/__init__.py)."""
from street_crafter_b200.rendering import (fully_fused_projection, isect_offset_encode, isect_tiles,  # noqa
                                           rasterization, rasterize_to_pixels, spherical_harmonics)
