"""Import shim for `from diff_point_rasterization import PointRasterizationSettings, PointRasterizer`
(/root/reference/data_processor/utils/render_utils.py:129)."""
from street_crafter_b200.point_rasterizer import PointRasterizationSettings, PointRasterizer  # noqa: F401
