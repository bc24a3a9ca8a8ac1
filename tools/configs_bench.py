"""Timings of the BASELINE.json configurations that are NOT the bench line (bench.py measures
configs[1]); they are parity-test cases, this tool only records how long they take on a B200.

    python tools/configs_bench.py lidar        # configs[2]: 5 M points -> 25 frames 1024x576 (PointRasterizer)
    python tools/configs_bench.py sweep        # configs[4]: render-only 1/2/4/8 M Gaussians at 1920x1280
    python tools/configs_bench.py knn          # configs[4]: distCUDA2 on 5 M points
    torchrun --nproc-per-node N tools/configs_bench.py multiview   # configs[3]: 6 M Gaussians, 40 views

Prints one JSON line per measurement (CUDA events, 1 warm-up pass, max over ranks where N > 1)."""
from __future__ import annotations

import json
import time
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from tools.scene import make_camera, make_gaussians, make_lidar_points  # noqa: E402


def ev_ms(fn, reps=1):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def lidar(dev):
    from street_crafter_b200.knn import distCUDA2
    from street_crafter_b200.point_rasterizer import PointRasterizationSettings, PointRasterizer
    P, W, H, frames = 5_000_000, 1024, 576, 25
    xyz, rgb = make_lidar_points(P, seed=3)
    xyz, rgb = xyz.to(dev), rgb.to(dev)
    fx = 2055.6 * W / 1920.0
    tanfovx, tanfovy = W / (2 * fx), H / (2 * fx)
    # render_utils.py:125-127: radius from the KNN distance, clamped
    radius = torch.sqrt(torch.clamp_min(distCUDA2(xyz), 1e-7)).clamp(max=0.05)
    opac = torch.ones(P, 1, device=dev)

    def trajectory():
        out = []
        for f in range(frames):
            V, _ = make_camera(W, H, yaw_deg=-10.0 + 20.0 * f / (frames - 1), z_shift=0.4 * f, x_shift=0.05 * f)
            s = PointRasterizationSettings(H, W, tanfovx, tanfovy, torch.zeros(3, device=dev), 1.0, V.T.to(dev),
                                           torch.eye(4, device=dev), 0, 10, torch.zeros(3, device=dev), False, False)
            with torch.no_grad():
                img, depth, alpha, radii = PointRasterizer(s)(xyz, None, rgb, opac, radius)
            out.append(img)
        return out

    trajectory()
    ms = ev_ms(trajectory)
    cov = float((trajectory()[frames // 2].sum(0) > 0).float().mean())
    print(json.dumps({"config": "configs[2] lidar condition render", "points": P, "frames": frames, "size": [W, H],
                      "ms_total": ms, "frames_per_s": frames / ms * 1e3, "Mpix_per_s": frames * W * H / ms / 1e3,
                      "mid_frame_coverage": cov}))


def sweep(dev):
    from street_crafter_b200.renderer import render_kernel_gsplat
    W, H = 1920, 1280
    for n_m in (1, 2, 4, 8):
        N = n_m * 1_000_000
        sc = {k: v.to(dev) for k, v in make_gaussians(N, 3, seed=11).items()}
        cams = [make_camera(W, H, yaw_deg=y, z_shift=z) for y, z in ((0.0, 0.0), (-35.0, 2.0), (35.0, 4.0))]
        cams = [(V[None].to(dev), K[None].to(dev), (-(V[:3, :3].T @ V[:3, 3])).to(dev)) for V, K in cams]
        occ1 = sc["opacities"].reshape(-1, 1)

        def views():
            with torch.no_grad():
                for V, K, c in cams:
                    render_kernel_gsplat(sc["means"], sc["shs"], sc["scales"], sc["quats"], occ1,
                                         V, K, c, W, H, 3, train=False)

        views()
        ms = ev_ms(views, 3) / len(cams)
        # the same three views as ONE render-only CUDA graph (sync-free binning): what a trajectory render replays
        from street_crafter_b200.pipeline import GraphedStep, ViewPipeline
        gs = GraphedStep(ViewPipeline(dev), None, lambda: (sc["means"], sc["shs"], sc["scales"], sc["quats"], occ1), cams,
                         W, H, 3, None, None)
        gs.replay()
        ms_g = ev_ms(gs.replay, 5) / len(cams)
        torch.cuda.synchronize()
        assert gs.overflowed() == []
        print(json.dumps({"config": "configs[4] render-only sweep", "gaussians": N, "size": [W, H],
                          "ms_per_view": ms, "views_per_s": 1e3 / ms, "Mpix_per_s": W * H / ms / 1e3,
                          "graph_ms_per_view": ms_g, "graph_views_per_s": 1e3 / ms_g}))
        del sc, gs
        torch.cuda.empty_cache()


def knn(dev):
    """distCUDA2 beside the REFERENCE's own simple_knn.cu (oracle/_ref/libsimpleknn_ref.so, compiled from
    /root/reference/submodules/simple-knn where it lies; simple_knn.cu:186-222) on the same device and points"""
    import ctypes
    from street_crafter_b200.knn import distCUDA2
    ref = None
    path = os.path.join(ROOT, "oracle", "_ref", "libsimpleknn_ref.so")
    if os.path.exists(path):
        ref = ctypes.CDLL(path)
        ref.ref_knn.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    for P in (100_000, 1_000_000, 5_000_000):
        xyz, _ = make_lidar_points(P, seed=5)
        xyz = xyz.to(dev).contiguous()
        distCUDA2(xyz)
        ms = ev_ms(lambda: distCUDA2(xyz), 3)
        row = {"config": "configs[4] distCUDA2", "points": P, "ms": ms, "Mpoints_per_s": P / ms / 1e3,
               "frac_of_64B_per_point_hbm_floor": 64.0 * P / (ms * 1e-3) / 6535.4e9}
        if ref is not None:
            out = torch.zeros(P, device=dev)
            ref.ref_knn(P, xyz.data_ptr(), out.data_ptr())       # warm-up (the reference syncs inside)
            t0 = time.perf_counter()
            ref.ref_knn(P, xyz.data_ptr(), out.data_ptr())
            torch.cuda.synchronize()
            ms_ref = (time.perf_counter() - t0) * 1e3              # its two blocking cudaMemcpy make wall clock the honest timer
            got = distCUDA2(xyz)
            row.update({"reference_simple_knn_ms": ms_ref, "speedup_vs_reference_kernel": ms_ref / ms,
                        "bit_identical_frac": float((got == out).float().mean())})
        print(json.dumps(row))


def lidar_pipeline(dev):
    """configs[2] through the device pipeline of row N1 (street_crafter_b200/lidar_condition.py): 25 condition frames of
    a 5 M-point aggregate (21 background frames + 2 actors), gather -> projection -> binning -> max_hit compositing ->
    uint8 pack, nothing leaves the device"""
    import numpy as np
    from street_crafter_b200.lidar_condition import LidarClouds, render_condition
    W, H, frames, n_frames = 1024, 576, 25, 45
    per = 5_000_000 // 21
    rng = np.random.default_rng(0)
    ply = {"background": {}}
    for f in range(n_frames):
        xyz = np.stack([rng.uniform(-40, 40, per), rng.uniform(-3, 6, per), rng.uniform(1, 80, per) + 0.4 * f], -1)
        ply["background"][f] = np.concatenate([xyz, rng.uniform(0, 1, (per, 3))], -1).astype(np.float32)
    for a in range(2):
        ply[f"car{a}"] = {f: np.concatenate([rng.normal(0, 1, (2000, 3)), rng.uniform(0, 1, (2000, 3))], -1) for f in range(n_frames)}
    clouds = LidarClouds(ply, dev)
    K = make_camera(W, H)[1].numpy()

    def trajectory():
        outs = []
        for i in range(frames):
            f = 10 + i
            V, _ = make_camera(W, H, yaw_deg=-10.0 + 20.0 * i / (frames - 1), z_shift=0.4 * f, x_shift=0.05 * i)
            c2w = torch.linalg.inv(V.double()).float().numpy()
            poses = [(f"car{a}", np.array([[1, 0, 0, 3.0 * a], [0, 1, 0, 0.5], [0, 0, 1, 15.0 + 0.4 * f], [0, 0, 0, 1.0]])) for a in range(2)]
            u8, _ = render_condition(clouds, c2w, K, H, W, f - 10, f + 10, poses, scale=0.01, use_ndc_scale=True)
            outs.append(u8)
        return outs

    trajectory()
    ms = ev_ms(trajectory)
    cov = float((trajectory()[frames // 2][..., 3] > 0).float().mean())
    print(json.dumps({"config": "configs[2] LiDAR condition pipeline on the device (N1)", "points_per_frame": 21 * per + 2 * 21 * 2000,
                      "frames": frames, "size": [W, H], "ms_total": ms, "frames_per_s": frames / ms * 1e3,
                      "mid_frame_coverage": cov}))


def multiview(dev):
    import torch.distributed as dist
    from street_crafter_b200.losses import photometric_loss_render
    from street_crafter_b200.parallel import GaussianArena, allreduce_gradients, shard_views
    from street_crafter_b200.renderer import render_kernel_gsplat
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    N, W, H, n_views = 6_000_000, 1920, 1280, 40
    arena = GaussianArena(make_gaussians(N, 3, seed=21), dev)
    g = torch.Generator().manual_seed(5)
    gt = torch.rand(3, H, W, generator=g).to(dev)
    cams = []
    for v in range(n_views):
        cam, frame = v % 5, v // 5
        V, K = make_camera(W, H, yaw_deg=-60.0 + 30.0 * cam, z_shift=1.5 * frame)
        cams.append((V[None].to(dev), K[None].to(dev), (-(V[:3, :3].T @ V[:3, 3])).to(dev)))
    mine = list(shard_views(n_views, rank, world))
    sinks = arena.sinks()

    def step():
        arena.zero_grad()
        for v in mine:
            V, K, c = cams[v]
            out = render_kernel_gsplat(arena.means, arena.shs, arena.scales, arena.quats, arena.opacities,
                                       V, K, c, W, H, 3, grad_sinks=sinks)
            photometric_loss_render(out["render_colors"], gt, None, 0.2, 1.0)[0].backward()
        if world > 1:
            allreduce_gradients(arena)

    step()
    if world > 1:
        dist.barrier()
    ms = ev_ms(step, 2)
    t = torch.tensor([ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        ms = float(t)
        print(json.dumps({"config": "configs[3] multi-camera batch", "gaussians": N, "views": n_views, "n_gpus": world,
                          "ms_per_step": ms, "views_per_s": n_views / ms * 1e3,
                          "Mpix_per_s": n_views * W * H / ms / 1e3, "scaling": "strong"}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    for what in sys.argv[1:] or ["lidar", "sweep", "knn"]:
        {"lidar": lidar, "lidar_pipeline": lidar_pipeline, "sweep": sweep, "knn": knn, "multiview": multiview}[what](dev)
