"""TEST INFRASTRUCTURE (checker, not product): stand-ins for the objects and absent third-party glue the
reference's callers touch around the drop-in boundary (easyvolcap's 4-line affine helpers and its
`convert_to_gaussian_camera`, the reference `Camera` fields), used by tests/test_gpu_boundary.py and
tests/test_boundary_cpu.py together with the bytecode table of oracle/build_ref_callers.py."""
import math
import types

import numpy as np
import torch


def bind(table, names, ns):
    """exec the reference's function definitions into `ns` (their module globals) -> ns"""
    for n in names:
        exec(table[n], ns)
    return ns


def affine_padding(c2w):
    """easyvolcap.utils.math_utils.affine_padding: [..,3,4] -> [..,4,4] with a (0,0,0,1) row"""
    row = torch.zeros_like(c2w[..., :1, :])
    row[..., 0, 3] = 1.0
    return torch.cat([c2w, row], dim=-2)


def affine_inverse(m):
    """easyvolcap.utils.math_utils.affine_inverse of a rigid [4,4] transform"""
    R, t = m[..., :3, :3], m[..., :3, 3:]
    Rt = R.mT
    out = torch.zeros_like(m)
    out[..., :3, :3] = Rt
    out[..., :3, 3:] = -Rt @ t
    out[..., 3, 3] = 1.0
    return out


def make_ref_camera(V, K, W, H, dev):
    """stand-in for street_gaussian.utils.camera_utils.Camera (fields as set at camera_utils.py:30-58)"""
    c2w = torch.linalg.inv(V.double()).float()
    return types.SimpleNamespace(
        image_width=W, image_height=H, R=c2w[:3, :3].numpy(), T=V[:3, 3].numpy(),
        world_view_transform=V.T.contiguous().to(dev), K=K.to(dev), znear=0.001, zfar=1000.0,
        camera_center=torch.linalg.inv(V.T.double()).float()[3, :3].to(dev))


def easyvolcap_stub(table):
    """`easyvolcap.utils.gaussian_utils.convert_to_gaussian_camera` (absent third-party glue), restated
    on the reference tree's own `getProjectionMatrixK` / `focal2fov` (graphics_utils.py:79-110)"""
    ns = bind(table, ["getProjectionMatrixK", "focal2fov"], {"torch": torch, "math": math, "np": np})

    def convert_to_gaussian_camera(K, R, T, H, W, n, f, cpu_K, cpu_R, cpu_T, cpu_H, cpu_W, cpu_n=0.01, cpu_f=100.0):
        o = types.SimpleNamespace()
        o.image_height, o.image_width = cpu_H, cpu_W
        o.FoVx = ns["focal2fov"](float(cpu_K[0, 0]), float(cpu_W))
        o.FoVy = ns["focal2fov"](float(cpu_K[1, 1]), float(cpu_H))
        w2c = torch.eye(4, device=K.device)
        w2c[:3, :3], w2c[:3, 3:] = R, T
        o.world_view_transform = w2c.transpose(0, 1).float()
        P = ns["getProjectionMatrixK"](K, float(cpu_H), float(cpu_W), float(cpu_n), float(cpu_f))
        o.projection_matrix = P.transpose(0, 1).float().to(K.device)
        o.full_proj_transform = torch.matmul(o.world_view_transform, o.projection_matrix).float()
        o.camera_center = (-R.mT @ T)[..., 0].float()
        o.tanfovx, o.tanfovy = np.tan(o.FoVx * 0.5), np.tan(o.FoVy * 0.5)
        return o

    pkg = types.ModuleType("easyvolcap")
    utils = types.ModuleType("easyvolcap.utils")
    gu = types.ModuleType("easyvolcap.utils.gaussian_utils")
    gu.convert_to_gaussian_camera = convert_to_gaussian_camera
    pkg.utils, utils.gaussian_utils = utils, gu
    return {"easyvolcap": pkg, "easyvolcap.utils": utils, "easyvolcap.utils.gaussian_utils": gu}
