"""Import shim: lets the UNMODIFIED reference renderer's
`from gsplat.rendering import ...` (street_gaussian_renderer.py:204) resolve to
street_crafter_b200.  Put the repo root on PYTHONPATH ahead of any real gsplat."""
from street_crafter_b200.rendering import (fully_fused_projection, isect_offset_encode, isect_tiles,  # noqa
                                           rasterization, rasterize_to_pixels, spherical_harmonics)
