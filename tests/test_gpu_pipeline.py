"""pipeline.ViewPipeline (compositing on the compute stream, the HBM-bound stages of the neighbouring views on a
memory stream) against the sequential schedule: the same kernels on the same inputs, so images are bit-equal and
gradients agree to the rounding of the compositing backward's floating-point atomics."""
import pytest
import torch

from tools.scene import make_camera, make_gaussians

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("ahead,pipe_sh,pipe_tail", [(1, True, True), (3, True, True), (1, False, True), (2, True, False)])
def test_view_pipeline_equals_sequential(ahead, pipe_sh, pipe_tail):
    from street_crafter_b200.parallel import GaussianArena
    from street_crafter_b200.pipeline import ViewPipeline
    from street_crafter_b200.renderer import render_kernel_gsplat
    dev = torch.device("cuda:0")
    n, W, H, deg = 60000, 640, 400, 3
    scene = make_gaussians(n, sh_degree=deg, seed=21)
    views = []
    for yaw in (-30.0, 0.0, 25.0, 50.0):
        V, K = make_camera(W, H, yaw_deg=yaw)
        views.append((V[None].to(dev), K[None].to(dev), (-(V[:3, :3].T @ V[:3, 3])).to(dev)))
    g = torch.Generator().manual_seed(5)
    v_rc = [torch.randn(1, H, W, 4, generator=g).to(dev) for _ in views]
    v_ra = [torch.randn(1, H, W, 1, generator=g).to(dev) for _ in views]

    def bwd(v, out):
        torch.autograd.backward([out["render_colors"], out["render_alphas"]], [v_rc[v], v_ra[v]])

    res = []
    for mode in ("sequential", "pipeline", "pipeline"):   # the pipeline twice: a second step reuses freed blocks
        arena = GaussianArena(scene, dev)
        sinks = arena.sinks()
        arena.zero_grad()
        if mode == "sequential":
            outs = []
            for v, (V, K, C) in enumerate(views):
                out = render_kernel_gsplat(arena.means, arena.shs, arena.scales, arena.quats, arena.opacities, V, K, C, W, H,
                                           deg, grad_sinks=sinks, aux_ids=False)
                bwd(v, out)
                outs.append(out)
        else:
            pipe = ViewPipeline(dev, ahead=ahead, pipeline_sh=pipe_sh, pipeline_tail=pipe_tail)
            outs = pipe.step(arena.means, arena.shs, arena.scales, arena.quats, arena.opacities, views, W, H, deg, sinks, bwd)
        torch.cuda.synchronize()
        res.append((arena, outs))
    (a0, o0) = res[0]
    for a1, o1 in res[1:]:
        for x, y in zip(o0, o1):
            for k in ("render_colors", "render_alphas", "depth", "radii", "visibility_filter"):
                assert torch.equal(x[k], y[k]), k
            assert torch.equal(x["aux"]["flatten_ids"], y["aux"]["flatten_ids"])
            assert torch.equal(x["aux"]["isect_offsets"], y["aux"]["isect_offsets"])
            gx, gy = x["viewspace_points"].grad, y["viewspace_points"].grad
            assert (gx - gy).abs().max().item() <= 1e-4 * gx.abs().max().item()
        for k, gs in a0.sinks().items():
            g1 = a1.sinks()[k]
            assert gs.abs().max().item() > 0, k
            # same kernels, atomics in another order (see test_prepare_view_two_phase_equals_single_pass)
            assert (gs - g1).abs().max().item() <= 1e-4 * gs.abs().max().item(), k


def _scene_views(dev, n=60000, W=640, H=400, deg=3, yaws=(-30.0, 0.0, 25.0)):
    scene = make_gaussians(n, sh_degree=deg, seed=33)
    views = []
    for yaw in yaws:
        V, K = make_camera(W, H, yaw_deg=yaw)
        views.append((V[None].to(dev), K[None].to(dev), (-(V[:3, :3].T @ V[:3, 3])).to(dev)))
    return scene, views, W, H, deg


def test_sync_free_binning_equals_read_back_path():
    """sc_bin_finish_dev + sc_rasterize_fwd_depth_dev (counts on the device, outputs sized by a capacity) against the
    read-back path: the valid prefix of the list, the offsets and the image are bit-equal; a capacity that is too
    small is reported and stays in bounds"""
    from street_crafter_b200.rendering import capacity_exceeded, fused_view, prepare_view, prepare_view_begin, prepare_view_finish
    dev = torch.device("cuda:0")
    scene, views, W, H, deg = _scene_views(dev)
    c = {k: v.to(dev) for k, v in scene.items()}
    V, K, C = views[1]
    ref = prepare_view(c["means"], c["quats"], c["scales"], c["opacities"], V, K, W, H, near_plane=0.001, far_plane=1000.0,
                       need_isect_ids=True)
    n_is = ref["flatten_ids"].numel()
    assert n_is > 10000
    host = torch.zeros(3, dtype=torch.int64).pin_memory()
    for cap in (n_is, n_is + 1, int(1.25 * n_is) + 4096):
        pb = prepare_view_finish(prepare_view_begin(c["means"], c["quats"], c["scales"], c["opacities"], V, K, W, H,
                                                    near_plane=0.001, far_plane=1000.0, capacity=cap, counts_host=host), True)
        torch.cuda.synchronize()
        assert int(host[1]) == n_is and not capacity_exceeded(host, cap)
        assert pb["flatten_ids"].numel() == cap
        assert torch.equal(pb["flatten_ids"][:n_is], ref["flatten_ids"])
        assert torch.equal(pb["isect_ids"][:n_is], ref["isect_ids"])
        assert torch.equal(pb["isect_offsets"], ref["isect_offsets"])
        assert torch.equal(pb["tpg"], ref["tpg"])
        rc0, ra0, _ = fused_view(c["means"], c["quats"], c["scales"], c["opacities"], c["shs"], V, K, C, W, H, deg,
                                 near_plane=0.001, far_plane=1000.0, prepared=ref)
        rc1, ra1, _ = fused_view(c["means"], c["quats"], c["scales"], c["opacities"], c["shs"], V, K, C, W, H, deg,
                                 near_plane=0.001, far_plane=1000.0, prepared=pb)
        assert torch.equal(rc0, rc1) and torch.equal(ra0, ra1)
    # too small: flagged, the view renders EMPTY (all offsets 0), nothing out of bounds
    cap = n_is // 2
    pb = prepare_view_finish(prepare_view_begin(c["means"], c["quats"], c["scales"], c["opacities"], V, K, W, H,
                                                near_plane=0.001, far_plane=1000.0, capacity=cap, counts_host=host), False)
    rc1, _, _ = fused_view(c["means"], c["quats"], c["scales"], c["opacities"], c["shs"], V, K, C, W, H, deg,
                           near_plane=0.001, far_plane=1000.0, prepared=pb)
    torch.cuda.synchronize()
    assert capacity_exceeded(host, cap) and int(host[1]) == n_is
    assert int(pb["isect_offsets"].max()) == 0 and float(rc1.abs().max()) == 0.0


def test_graphed_step_equals_eager_and_reports_overflow():
    from street_crafter_b200.parallel import GaussianArena
    from street_crafter_b200.pipeline import GraphedStep, ViewPipeline
    from street_crafter_b200.renderer import render_kernel_gsplat
    dev = torch.device("cuda:0")
    scene, views, W, H, deg = _scene_views(dev)
    arena = GaussianArena(scene, dev)
    sinks = arena.sinks()
    g = torch.Generator().manual_seed(5)
    v_rc = [torch.randn(1, H, W, 4, generator=g).to(dev) for _ in views]
    v_ra = [torch.randn(1, H, W, 1, generator=g).to(dev) for _ in views]

    def bwd(v, out):
        torch.autograd.backward([out["render_colors"], out["render_alphas"]], [v_rc[v], v_ra[v]])

    arena.zero_grad()
    ref_out = []
    for v, (V, K, C) in enumerate(views):
        out = render_kernel_gsplat(arena.means, arena.shs, arena.scales, arena.quats, arena.opacities, V, K, C, W, H, deg,
                                   grad_sinks=sinks, aux_ids=False)
        bwd(v, out)
        ref_out.append(out["render_colors"].clone())
    torch.cuda.synchronize()
    ref = {k: t.clone() for k, t in sinks.items()}
    gs = GraphedStep(ViewPipeline(dev), arena.zero_grad,
                     lambda: (arena.means, arena.shs, arena.scales, arena.quats, arena.opacities), views, W, H, deg,
                     lambda: sinks, bwd)
    for _ in range(3):
        outs = gs.replay()
        torch.cuda.synchronize()
        assert gs.overflowed() == []
        for o, r in zip(outs, ref_out):
            assert torch.equal(o["render_colors"], r)
        for k in ref:
            assert (sinks[k] - ref[k]).abs().max().item() <= 1e-4 * ref[k].abs().max().item(), k
    # two graphs (what the multi-GPU step replays around the start of its SH-gradient exchange)
    gs2 = GraphedStep(ViewPipeline(dev), arena.zero_grad,
                      lambda: (arena.means, arena.shs, arena.scales, arena.quats, arena.opacities), views, W, H, deg,
                      lambda: sinks, bwd, split_projection_backward=True)
    for _ in range(2):
        gs2.replay_main()
        torch.cuda.synchronize()
        assert (sinks["shs"] - ref["shs"]).abs().max().item() <= 1e-4 * ref["shs"].abs().max().item()   # final already
        assert float(sinks["quats"].abs().max()) == 0.0   # the projection backward has not run yet
        gs2.replay_tail()
        torch.cuda.synchronize()
        for k in ref:
            assert (sinks[k] - ref[k]).abs().max().item() <= 1e-4 * ref[k].abs().max().item(), k
    del gs2
    # the camera moves (static tensors rewritten in place): same graph, new image; a view that sees much more than
    # the capacity allows is reported
    V2, K2 = make_camera(W, H, yaw_deg=10.0)
    views[0][0].copy_(V2[None])
    views[0][2].copy_(-(V2[:3, :3].T @ V2[:3, 3]))
    outs = gs.replay()
    torch.cuda.synchronize()
    arena.zero_grad()
    out = render_kernel_gsplat(arena.means, arena.shs, arena.scales, arena.quats, arena.opacities, views[0][0], views[0][1],
                               views[0][2], W, H, deg, grad_sinks=sinks, aux_ids=False)
    torch.cuda.synchronize()
    if gs.overflowed() == []:
        assert torch.equal(outs[0]["render_colors"], out["render_colors"])
    gs.capacities[1] = 4096   # pretend: capacity far too small (takes effect at the next capture)
    gs.margin = 0.1
    gs.recapture()
    gs.replay()
    torch.cuda.synchronize()
    assert gs.overflowed() == [0, 1, 2]


def test_compositing_occupancy_limit_changes_nothing_but_residency():
    """sc_rasterize_limit_warps (a dynamic shared-memory request that caps the resident one-warp CTAs per SM):
    same images, same gradients to atomic rounding"""
    from street_crafter_b200._lib import check, lib
    from street_crafter_b200.renderer import render_kernel_gsplat
    dev = torch.device("cuda:0")
    scene, views, W, H, deg = _scene_views(dev, n=30000)
    V, K, C = views[1]
    res = []
    try:
        for warps in (0, 6, 12):
            check(lib.sc_rasterize_limit_warps(warps), "sc_rasterize_limit_warps")
            c = {k: v.to(dev).requires_grad_(True) for k, v in scene.items()}
            out = render_kernel_gsplat(c["means"], c["shs"], c["scales"], c["quats"], c["opacities"][:, None], V, K, C, W, H, deg)
            (out["render_colors"].sum() + out["render_alphas"].sum()).backward()
            res.append((out["render_colors"].detach().clone(), c["means"].grad.clone()))
    finally:
        lib.sc_rasterize_limit_warps(0)
    assert lib.sc_rasterize_limit_warps(65) == -1
    for img, g in res[1:]:
        assert torch.equal(img, res[0][0])
        assert (g - res[0][1]).abs().max().item() <= 1e-4 * res[0][1].abs().max().item()


def test_render_only_graph_equals_eager_eval():
    """GraphedStep without a backward: a trajectory render replays ONE graph per batch of views; the camera tensors
    are rewritten in place"""
    from street_crafter_b200.pipeline import GraphedStep, ViewPipeline
    from street_crafter_b200.renderer import render_kernel_gsplat
    dev = torch.device("cuda:0")
    scene, views, W, H, deg = _scene_views(dev)
    c = {k: v.to(dev) for k, v in scene.items()}
    occ1 = c["opacities"][:, None]
    gs = GraphedStep(ViewPipeline(dev), None, lambda: (c["means"], c["shs"], c["scales"], c["quats"], occ1), views, W, H,
                     deg, None, None)
    for shift in (0.0, 4.0):
        if shift:
            V2, K2 = make_camera(W, H, yaw_deg=5.0, z_shift=shift)
            views[2][0].copy_(V2[None]); views[2][2].copy_(-(V2[:3, :3].T @ V2[:3, 3]))
        outs = gs.replay()
        torch.cuda.synchronize()
        if gs.overflowed():
            gs.recapture()
            outs = gs.replay()
            torch.cuda.synchronize()
        for (V, K, C), o in zip(views, outs):
            with torch.no_grad():
                ref = render_kernel_gsplat(c["means"], c["shs"], c["scales"], c["quats"], occ1, V, K, C, W, H, deg, train=False)
            for k in ("rgb", "acc", "depth", "radii", "visibility_filter"):
                assert torch.equal(o[k], ref[k]), k
            assert float(o["rgb"].max()) <= 1.0
