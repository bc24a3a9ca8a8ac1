"""CPU-side checks of the drop-in boundary (no GPU needed): the bytecode table of the reference's callers is
current, and PointRasterizer decodes the camera the reference caller actually packs
(data_processor/utils/render_utils.py:131-163) — principal point, focal lengths and near/far from
`projmatrix` / `viewmatrix`, not from the image centre."""
import os

import numpy as np
import pytest
import torch

from oracle import build_ref_callers as BRC
from oracle.ref_standins import easyvolcap_stub
from tools.scene import make_camera


def _table():
    t, meta = BRC.load()
    if t is None:
        pytest.skip(meta)
    return t, meta


def test_ref_callers_table_is_current():
    """where /root/reference exists (the build container) the stored bytecode must be what the sources
    compile to today, and every entry the boundary tests bind must be present"""
    t, meta = _table()
    need = {"render_kernel_gsplat", "render_pointcloud_diff_point_rasterization", "create_from_pcd", "set_max_radii2D",
            "add_densification_stats", "getProjectionMatrixK", "focal2fov", "make_lidar_ply", "transform_lidar_ply",
            "render_condition"}
    assert need <= set(t), need - set(t)
    if os.path.isdir(BRC.REF):
        _, fresh = BRC.extract()
        assert fresh["entries"] == meta["entries"], "oracle/_ref/ref_callers.bin is stale: python oracle/build_ref_callers.py"
    e = meta["entries"]["render_kernel_gsplat"]
    assert e["file"].endswith("street_gaussian_renderer.py") and e["lines"] == [172, 302]


@pytest.mark.parametrize("cxf,cyf,near,far", [(0.5, 0.5, 1.0, 100.0), (0.47, 0.55, 1.0, 100.0), (0.52, 0.44, 0.5, 250.0)])
def test_point_rasterizer_decodes_the_callers_camera(cxf, cyf, near, far):
    from street_crafter_b200.point_rasterizer import PointRasterizationSettings, decode_camera
    t, _ = _table()
    conv = easyvolcap_stub(t)["easyvolcap.utils.gaussian_utils"].convert_to_gaussian_camera
    W, H = 1024, 576
    V, K = make_camera(W, H, yaw_deg=17.0, z_shift=35.0, x_shift=-4.0)
    K = K.clone()
    K[0, 2], K[1, 2], K[1, 1] = cxf * W, cyf * H, K[0, 0] * 1.02
    as_t = torch.as_tensor
    cam = conv(K, V[:3, :3], V[:3, 3:], as_t(H), as_t(W), as_t(near), as_t(far), K.numpy(), None, None, as_t(H), as_t(W),
               as_t(near), as_t(far))
    s = PointRasterizationSettings(
        image_height=cam.image_height, image_width=cam.image_width, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
        bg=torch.zeros(3), scale_modifier=1.0, viewmatrix=cam.world_view_transform, projmatrix=cam.full_proj_transform,
        sh_degree=0, max_hit=10, campos=cam.camera_center, prefiltered=False, debug=False)
    fx, fy, cx, cy, n, f = decode_camera(s)
    assert abs(fx - float(K[0, 0])) < 2e-3 and abs(fy - float(K[1, 1])) < 2e-3, (fx, fy)
    assert abs(cx - cxf * W) < 2e-3 and abs(cy - cyf * H) < 2e-3, (cx, cy)
    assert n == pytest.approx(near, rel=1e-5) and f == pytest.approx(far, rel=1e-4), (n, f)
    # overrides still win; a projmatrix of another form falls back to tanfov* and the image centre
    s2 = s._replace(principal=(10.0, 20.0), znear=2.0, zfar=50.0)
    assert decode_camera(s2)[2:] == (10.0, 20.0, 2.0, 50.0)
    s3 = s._replace(projmatrix=torch.eye(4))
    fx3, fy3, cx3, cy3, n3, f3 = decode_camera(s3)
    assert (cx3, cy3, n3, f3) == (W / 2.0, H / 2.0, 1.0, 100.0) and abs(fx3 - float(K[0, 0])) < 1e-2
