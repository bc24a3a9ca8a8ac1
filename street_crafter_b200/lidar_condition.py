"""LiDAR condition pipeline on the device: the reference's
`WaymoPointCloudProcessor.render_condition`
(/root/reference/street_gaussian/pointcloud_processor/waymo_processor.py:179-242, aggregation in
base_processor.py:32-63, splat in data_processor/utils/render_utils.py:83-183) without the NumPy
concatenation, the host<->device copies and the PNG round trip that the diffusion runner reads back
(street_gaussian/utils/diffusion_utils.py:65-97, :268-281).

The per-frame clouds are uploaded ONCE; a condition frame is then: one gather kernel (aggregate
+-delta_frames of background, pose every active actor's clouds in float64 like the reference, splat
scales) -> projection -> binning -> compositing with max_hit -> one pack kernel (uint8 [H,W,4] = the
content of the reference's rgb + mask PNG pair, and / or float [4,H,W]).  No CPU path."""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
from torch import Tensor

from . import rendering as R
from ._lib import check, lib

_SEG_DTYPE = np.dtype([("dst0", "<i8"), ("count", "<i8"), ("src", "<u8"), ("is_f64", "<i4"), ("has_pose", "<i4"),
                       ("pose", "<f8", (12,))])
MAX_SEGS = 512


class LidarClouds:
    """The reference's `ply_dict` ({'background': {frame: [n,6]}, actor_id: {frame: [n,6]}}, rows = xyz + rgb;
    waymo_processor.py:39-97) on the device: world-space background clouds as float32 (what
    `torch.as_tensor(points).float()` makes of them, render_utils.py:108), actor clouds in their box frame as
    float64 (they are posed in float64 before that cast, base_processor.py:58-63), frames stored consecutively
    so that a +-delta_frames window is ONE contiguous run per cloud."""

    def __init__(self, ply_dict: Dict, device):
        assert int(lib.sc_lidar_segment_bytes()) == _SEG_DTYPE.itemsize
        self.device = torch.device(device)
        self.bkgd, self.bkgd_range = self._pack(ply_dict["background"], torch.float32)
        self.actors: Dict = {}
        for key, frames in ply_dict.items():
            if key in ("background", "background_visible"):
                continue
            self.actors[key] = self._pack(frames, torch.float64)

    def _pack(self, frames: Dict[int, np.ndarray], dtype) -> Tuple[Tensor, Dict[int, Tuple[int, int]]]:
        rows, ranges, off = [], {}, 0
        for f in sorted(frames):
            a = np.asarray(frames[f])[:, :6]
            ranges[f] = (off, off + a.shape[0])
            off += a.shape[0]
            rows.append(torch.from_numpy(np.ascontiguousarray(a)).to(dtype))
        t = torch.cat(rows, 0) if rows else torch.zeros(0, 6, dtype=dtype)
        return t.contiguous().to(self.device), ranges

    @staticmethod
    def window(ranges: Dict[int, Tuple[int, int]], start_frame: int, end_frame: int) -> Optional[Tuple[int, int]]:
        """rows of the frames in [start_frame, end_frame] that exist (make_lidar_ply skips missing ones)"""
        fr = [f for f in ranges if start_frame <= f <= end_frame]
        if not fr:
            return None
        return ranges[min(fr)][0], ranges[max(fr)][1]


_IDENTITY_QUATS: Dict[int, Tensor] = {}


def _identity_quats(n: int, dev) -> Tensor:
    key = dev.index if dev.index is not None else torch.cuda.current_device()
    q = _IDENTITY_QUATS.get(key)
    if q is None or q.shape[0] < n:
        q = torch.zeros(max(n, 1 << 20), 4, dtype=torch.float32, device=dev)
        q[:, 0] = 1.0
        _IDENTITY_QUATS[key] = q
    return q[:n]


@torch.no_grad()
def render_condition(clouds: LidarClouds, c2w, ixt, H: int, W: int, start_frame: int, end_frame: int,
                     actor_poses: Sequence[Tuple[object, np.ndarray]], scale: float = 0.01, use_ndc_scale: bool = True,
                     occ: float = 1.0, max_hit: int = 10, near: float = 1.0, far: float = 100.0,
                     out_uint8: bool = True, out_float: bool = False):
    """One condition frame.  c2w [4,4], ixt [3,3] (array-likes); frames [start_frame, end_frame] of the background
    and of every actor in `actor_poses` = [(actor_id, pose[4,4] float64 object->world)] in the reference's order
    (waymo_processor.py:205-224; actors without clouds in the window are skipped like make_lidar_ply does).
    -> (uint8 [H,W,4] | None, float32 [4,H,W] | None): rgb + mask, quantised like the reference's PNG pair."""
    dev = clouds.device
    segs, n = [], 0
    w = LidarClouds.window(clouds.bkgd_range, start_frame, end_frame)
    if w is not None and w[1] > w[0]:
        segs.append((n, w[1] - w[0], clouds.bkgd.data_ptr() + w[0] * 24, 0, 0, np.zeros(12)))
        n += w[1] - w[0]
    for actor_id, pose in actor_poses:
        if actor_id not in clouds.actors:
            continue
        t, ranges = clouds.actors[actor_id]
        w = LidarClouds.window(ranges, start_frame, end_frame)
        if w is None or w[1] <= w[0]:
            continue
        p = np.asarray(pose, dtype=np.float64)
        segs.append((n, w[1] - w[0], t.data_ptr() + w[0] * 48, 1, 1, p[:3, :4].reshape(-1)))
        n += w[1] - w[0]
    if len(segs) > MAX_SEGS:
        raise RuntimeError(f"{len(segs)} point-cloud segments in one frame (limit {MAX_SEGS})")
    c2w_t = torch.as_tensor(np.asarray(c2w), dtype=torch.float32)
    # affine_inverse of the float32 camera pose, as the reference does on the device (render_utils.py:105-106)
    Rm, tv = c2w_t[:3, :3], c2w_t[:3, 3:]
    w2c = torch.eye(4)
    w2c[:3, :3] = Rm.T
    w2c[:3, 3:] = -Rm.T @ tv
    K = torch.as_tensor(np.asarray(ixt), dtype=torch.float32)
    st = torch.cuda.current_stream(dev).cuda_stream
    with torch.cuda.device(dev):
        means = torch.empty(n, 3, dtype=torch.float32, device=dev)
        rgb = torch.empty(n, 3, dtype=torch.float32, device=dev)
        scales = torch.empty(n, 3, dtype=torch.float32, device=dev)
        cam = torch.cat([w2c.reshape(-1), K.reshape(-1)]).to(dev, non_blocking=True)
        if n:
            table = np.zeros(len(segs), dtype=_SEG_DTYPE)
            for i, (d0, cnt, src, f64, hp, pose) in enumerate(segs):
                table[i] = (d0, cnt, src, f64, hp, pose)
            segs_dev = torch.from_numpy(table.view(np.uint8)).to(dev, non_blocking=True)
            check(lib.sc_lidar_gather(n, len(segs), segs_dev.data_ptr(), cam.data_ptr() + 8 * 4, float(scale),
                                      float(K[0, 0]), int(H), int(W), 1 if use_ndc_scale else 0, means.data_ptr(),
                                      rgb.data_ptr(), scales.data_ptr(), st), "sc_lidar_gather")
        viewmat = cam[:16].view(1, 4, 4)
        Kd = cam[16:25].view(1, 3, 3)
        radii, means2d, depths, conics, _ = R.fully_fused_projection(
            means, None, _identity_quats(n, dev), scales, viewmat, Kd, W, H, near_plane=near, far_plane=far,
            calc_compensations=False)
        tw, th = math.ceil(W / 16.0), math.ceil(H / 16.0)
        state = R._isect_begin(means2d, radii, depths, 16, tw, th, torch.empty((1, n), dtype=torch.int32, device=dev), st)
        _, ids, flat, offsets = R._isect_finish(state, need_ids=False)
        if offsets is None:
            offsets = R.isect_offset_encode(ids, 1, tw, th)
        op = torch.full((1, n), float(occ), dtype=torch.float32, device=dev)
        rc, ra = R.rasterize_to_pixels(means2d, conics, rgb[None], op, W, H, 16, offsets, flat, backgrounds=None,
                                       max_hit=int(max_hit))
        out_u8 = torch.empty((H, W, 4), dtype=torch.uint8, device=dev) if out_uint8 else None
        out_f = torch.empty((4, H, W), dtype=torch.float32, device=dev) if out_float else None
        check(lib.sc_lidar_pack(H * W, rc.data_ptr(), ra.data_ptr(), None if out_u8 is None else out_u8.data_ptr(),
                                None if out_f is None else out_f.data_ptr(), st), "sc_lidar_pack")
    return out_u8, out_f


def box_pose(ego_pose: np.ndarray, box: Dict) -> np.ndarray:
    """ego_pose @ pose_vehicle of a tracked box (waymo_processor.py:214-220), float64"""
    h = box["heading"]
    pv = np.array([[np.cos(h), -np.sin(h), 0, box["center_x"]], [np.sin(h), np.cos(h), 0, box["center_y"]],
                   [0, 0, 1, box["center_z"]], [0, 0, 0, 1]])
    return np.asarray(ego_pose) @ pv


@torch.no_grad()
def render_conditions(clouds: LidarClouds, cameras: List[Dict], first_frame: int, last_frame: int,
                      delta_frames: int = 10, **kw) -> Tensor:
    """A trajectory of condition frames -> float32 [F,4,H,W] on the device (the tensor the diffusion runner
    stacks from the PNG files, diffusion_utils.py:268-281, before its own crop / resize).  cameras: dicts with
    c2w, ixt, H, W, frame, actor_poses (see render_condition)."""
    outs = []
    for cam in cameras:
        f = int(cam["frame"])
        s, e = max(first_frame, f - delta_frames), min(last_frame, f + delta_frames)
        _, o = render_condition(clouds, cam["c2w"], cam["ixt"], cam["H"], cam["W"], s, e, cam.get("actor_poses", ()),
                                out_uint8=False, out_float=True, **kw)
        outs.append(o)
    return torch.stack(outs, 0)
