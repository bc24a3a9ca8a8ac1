"""GPU parity tests (-m gpu): the sm_100a kernels, called through the C ABI, against the CPU
oracle on the same seeded inputs.

Bars (BASELINE.json north_star):
  * radii, tiles_per_gauss, sorted isect_ids, flatten_ids, isect_offsets  -> BIT-EXACT
  * rendered RGB / depth / alpha                                           -> 1e-4 max-abs (fp32)
  * gradients                                                              -> 1e-3 relative
The alpha >= 1/255 accept test is a hard discontinuity of the algorithm itself: two fp32
implementations whose exp() differ in the last bits can disagree on a (pixel, Gaussian) pair
that sits exactly on the threshold, which moves that pixel by up to alpha*T*colour ~ 4e-3.
IMAGE_OUTLIER_FRAC bounds how many pixels may do that (and by how much); everything else must
meet 1e-4."""
import math
import os

import numpy as np
import pytest
import torch

from oracle import gsplat_torch as G
from oracle import knn as KO
from tools.scene import make_camera, make_gaussians, make_lidar_points

pytestmark = pytest.mark.gpu

IMG_TOL = 1e-4
GRAD_RTOL = 1e-3
IMAGE_OUTLIER_FRAC = 2e-6   # measured: 8 of 3.5e7 compared values, every one on a threshold pair
IMAGE_OUTLIER_MAX = 8e-3  # ~ 2/255 of a unit colour


@pytest.fixture(scope="module")
def S():
    import street_crafter_b200 as S
    return S


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


def cuda(d, dev):
    return {k: v.to(dev) for k, v in d.items()}


MARGIN_TOL = 2e-3   # an outlier pixel must hold a pair within this RELATIVE distance of alpha = 1/255 or T = 1e-4
GRAD_ELEM_ATOL = 1e-6   # x max|reference|: the absolute floor of the per-element relative gradient check
GRAD_ELEM_BAD_FRAC = 1e-4


def _record(kind, what, **kw):
    """parity statistics of this session: printed by conftest's terminal summary, written to
    gpurun_out/parity_stats.json (DESIGN.md section 2 quotes them)"""
    test = os.environ.get("PYTEST_CURRENT_TEST", "").split(" ")[0].split("::")[-1]
    getattr(pytest, "parity_stats", []).append(dict(kind=kind, test=test, what=what, **kw))


def assert_image_close(got, want, what, scale=1.0, margin=None):
    """1e-4 max-abs (north_star), except pixels that sit on a hard discontinuity of the algorithm: at most
    IMAGE_OUTLIER_FRAC of the values, none above IMAGE_OUTLIER_MAX, and — when the oracle's `margin` map
    ([C,H,W], oracle/gsplat_torch.py rasterize_to_pixels(return_margin=True)) is given — EVERY one of them on a
    pixel that holds a (pixel, Gaussian) pair within MARGIN_TOL of the alpha = 1/255 or T = 1e-4 test."""
    diff = (got.cpu() - want).abs() / scale
    bad = diff > IMG_TOL
    n_bad = int(bad.sum())
    frac = n_bad / max(1, bad.numel())
    off_margin = None
    if margin is not None and n_bad:
        m = margin
        while m.dim() < bad.dim():
            m = m[..., None]
        off_margin = int((bad & ~(m.expand_as(bad) < MARGIN_TOL)).sum())
    _record("image", what, n_bad=n_bad, numel=bad.numel(), max_diff=float(diff.max().detach()), off_margin=off_margin)
    assert n_bad <= max(2, IMAGE_OUTLIER_FRAC * bad.numel()), f"{what}: {n_bad} of {bad.numel()} values off by > {IMG_TOL}"
    assert diff.max().item() <= IMAGE_OUTLIER_MAX, f"{what}: max diff {diff.max().item()}"
    assert not off_margin, f"{what}: {off_margin} of {n_bad} outliers are NOT on an alpha/T threshold pair"
    return n_bad


def assert_grad_close(got, want, what, rtol=GRAD_RTOL):
    """(1) global: max-abs and L2 error <= 1e-3 of the reference's scale; (2) PER ELEMENT (north_star's
    "1e-3 relative"): |got - want| <= 1e-3 |want| + GRAD_ELEM_ATOL max|want| on all but GRAD_ELEM_BAD_FRAC of
    the entries (fp32 sums of +- contributions in a different order cancel differently: both sides carry
    rounding of ~1e-6 x the LARGEST partial sum, which the floor stands for)."""
    got = got.detach().cpu().double()
    want = want.detach().double()
    assert torch.isfinite(got).all(), what
    denom = want.abs().max().item() + 1e-30
    err = (got - want).abs().max().item() / denom
    l2 = ((got - want).norm() / (want.norm() + 1e-30)).item()
    bad = (got - want).abs() > rtol * want.abs() + GRAD_ELEM_ATOL * denom
    frac_bad = float(bad.double().mean()) if bad.numel() else 0.0
    _record("grad", what, max_rel=err, l2_rel=l2, elem_bad=int(bad.sum()), numel=bad.numel())
    assert err <= rtol and l2 <= rtol, f"{what}: max-rel {err:.3e}, l2-rel {l2:.3e}"
    # measured: 22 of 3.6e7 entries over the whole suite; at most 3 in one comparison (the 0.8 m splat stress case,
    # whose conic gradients sum ~1e4 pixel terms of either sign under ex2.approx)
    assert int(bad.sum()) <= max(4, GRAD_ELEM_BAD_FRAC * bad.numel()), \
        f"{what}: {int(bad.sum())} of {bad.numel()} entries miss the per-element bar"


# ------------------------------------------------------------------------------ projection
@pytest.mark.parametrize("n,W,H,yaw", [(20000, 960, 640, 0.0), (5003, 1920, 886, 45.0), (1, 64, 48, 0.0),
                                       (257, 100, 70, -45.0)])
def test_projection_bit_exact(S, dev, n, W, H, yaw):
    s = make_gaussians(n, seed=n)
    V, K = make_camera(W, H, yaw_deg=yaw)
    want = G.fully_fused_projection(s["means"], None, s["quats"], s["scales"], V[None], K[None], W, H,
                                    near_plane=0.001, far_plane=1000.0, calc_compensations=True)
    c = cuda(s, dev)
    got = S.fully_fused_projection(c["means"], None, c["quats"], c["scales"], V[None].to(dev), K[None].to(dev), W,
                                   H, near_plane=0.001, far_plane=1000.0, calc_compensations=True, packed=False)
    names = ["radii", "means2d", "depths", "conics", "compensations"]
    for nm, g, w in zip(names, got, want):
        assert g.shape == w.shape and g.dtype == w.dtype, nm
        assert torch.equal(g.cpu(), w), f"{nm} not bit-exact: {(g.cpu() != w).sum().item()} mismatches"
    assert got[0].dtype == torch.int32


def test_projection_multi_camera_and_no_comp(S, dev):
    s = make_gaussians(3000, seed=11)
    cams = [make_camera(640, 384, yaw_deg=y, z_shift=z) for y, z in [(0, 0), (45, 1.5), (-45, 3.0)]]
    V = torch.stack([c[0] for c in cams])
    K = torch.stack([c[1] for c in cams])
    want = G.fully_fused_projection(s["means"], None, s["quats"], s["scales"], V, K, 640, 384,
                                    near_plane=0.001, far_plane=1000.0, calc_compensations=False)
    c = cuda(s, dev)
    got = S.fully_fused_projection(c["means"], None, c["quats"], c["scales"], V.to(dev), K.to(dev), 640, 384,
                                   near_plane=0.001, far_plane=1000.0)
    assert got[4] is None and want[4] is None
    for g, w in zip(got[:4], want[:4]):
        assert torch.equal(g.cpu(), w)


def test_projection_backward(S, dev):
    s = make_gaussians(6000, seed=21)
    W, H = 640, 384
    V, K = make_camera(W, H)
    leaves = [s[k].clone().requires_grad_(True) for k in ("means", "quats", "scales")]
    r, m2, dp, cn, cp = G.fully_fused_projection(leaves[0], None, leaves[1], leaves[2], V[None], K[None], W, H,
                                                 near_plane=0.001, far_plane=1000.0, calc_compensations=True)
    g = torch.Generator().manual_seed(0)
    ws = [torch.randn(t.shape, generator=g) for t in (m2, dp, cn, cp)]
    # conic entries span many orders of magnitude: weight them so every path matters
    (m2 * ws[0]).sum().add((dp * ws[1]).sum()).add((cn * ws[2]).sum()).add((cp * ws[3]).sum()).backward()
    cl = [s[k].to(dev).requires_grad_(True) for k in ("means", "quats", "scales")]
    r2, m2g, dpg, cng, cpg = S.fully_fused_projection(cl[0], None, cl[1], cl[2], V[None].to(dev), K[None].to(dev),
                                                      W, H, near_plane=0.001, far_plane=1000.0,
                                                      calc_compensations=True)
    ((m2g * ws[0].to(dev)).sum() + (dpg * ws[1].to(dev)).sum() + (cng * ws[2].to(dev)).sum()
     + (cpg * ws[3].to(dev)).sum()).backward()
    vis = (r[0] > 0)
    for nm, a, b in zip(("means", "quats", "scales"), cl, leaves):
        ref = torch.nan_to_num(b.grad)
        assert_grad_close(a.grad, ref, f"v_{nm}")
        assert (a.grad.cpu()[~vis] == 0).all(), f"culled Gaussians must get zero v_{nm}"


# ------------------------------------------------------------------------------ binning
def _binning_case(n, W, H, seed, C=1):
    s = make_gaussians(n, seed=seed)
    cams = [make_camera(W, H, yaw_deg=20.0 * c, z_shift=1.5 * c) for c in range(C)]
    V = torch.stack([c[0] for c in cams])
    K = torch.stack([c[1] for c in cams])
    proj = G.fully_fused_projection(s["means"], None, s["quats"], s["scales"], V, K, W, H, near_plane=0.001,
                                    far_plane=1000.0, calc_compensations=True)
    return s, V, K, proj


@pytest.mark.parametrize("n,W,H,C", [(30000, 960, 640, 1), (4000, 1920, 886, 1), (2000, 320, 192, 3),
                                     (50, 33, 17, 1), (150000, 960, 640, 1), (40000, 640, 384, 2)])
def test_isect_bit_exact(S, dev, n, W, H, C):
    s, V, K, (radii, m2, dep, con, comp) = _binning_case(n, W, H, seed=n + C, C=C)
    tw, th = math.ceil(W / 16), math.ceil(H / 16)
    tpg_w, ids_w, flat_w = G.isect_tiles(m2, radii, dep, 16, tw, th, n_cameras=C)
    off_w = G.isect_offset_encode(ids_w, C, tw, th)
    tpg, ids, flat = S.isect_tiles(m2.to(dev), radii.to(dev), dep.to(dev), 16, tw, th, packed=False, n_cameras=C)
    off = S.isect_offset_encode(ids, C, tw, th)
    assert ids.dtype == torch.int64 and flat.dtype == torch.int32 and tpg.dtype == torch.int32
    assert off.dtype == torch.int32 and off.shape == (C, th, tw)
    assert torch.equal(tpg.cpu(), tpg_w), "tiles_per_gauss"
    assert ids.numel() == ids_w.numel() > 0
    assert torch.equal(ids.cpu(), ids_w), "sorted isect_ids"
    assert torch.equal(flat.cpu(), flat_w), "flatten_ids (stable tie-break)"
    assert torch.equal(off.cpu(), off_w), "isect_offsets"
    # unsorted emission order too
    _, ids_u_w, flat_u_w = G.isect_tiles(m2, radii, dep, 16, tw, th, sort=False, n_cameras=C)
    _, ids_u, flat_u = S.isect_tiles(m2.to(dev), radii.to(dev), dep.to(dev), 16, tw, th, sort=False, n_cameras=C)
    assert torch.equal(ids_u.cpu(), ids_u_w) and torch.equal(flat_u.cpu(), flat_u_w)


def test_isect_empty(S, dev):
    m2 = torch.zeros(1, 10, 2, device=dev)
    radii = torch.zeros(1, 10, dtype=torch.int32, device=dev)
    dep = torch.zeros(1, 10, device=dev)
    tpg, ids, flat = S.isect_tiles(m2, radii, dep, 16, 4, 3, n_cameras=1)
    assert ids.numel() == 0 and flat.numel() == 0 and int(tpg.sum()) == 0
    off = S.isect_offset_encode(ids, 1, 4, 3)
    assert off.shape == (1, 3, 4) and (off == 0).all()


def test_isect_ties_are_stable(S, dev):
    """many Gaussians with IDENTICAL depth in one tile: order must be emission order."""
    n = 5000
    m2 = torch.full((1, n, 2), 24.0)
    radii = torch.full((1, n), 2, dtype=torch.int32)
    dep = torch.full((1, n), 5.0)
    dep[0, ::7] = 3.0
    _, ids_w, flat_w = G.isect_tiles(m2, radii, dep, 16, 8, 8, n_cameras=1)
    _, ids, flat = S.isect_tiles(m2.to(dev), radii.to(dev), dep.to(dev), 16, 8, 8, n_cameras=1)
    assert torch.equal(ids.cpu(), ids_w) and torch.equal(flat.cpu(), flat_w)


@pytest.mark.parametrize("n", [1, 31, 4095, 4096, 4097, 100003, 1 << 20])
@pytest.mark.parametrize("begin_bit,end_bit", [(0, 8), (0, 30), (0, 44), (0, 46), (0, 64), (32, 46), (32, 44),
                                               (7, 21)])
def test_radix_sort_matches_stable_sort(S, dev, n, begin_bit, end_bit):
    from street_crafter_b200._lib import check, lib
    g = torch.Generator().manual_seed(n * 131 + end_bit)
    hi = min(end_bit, 62)
    keys = torch.randint(0, 2 ** hi, (n,), generator=g, dtype=torch.int64)
    if n > 100:
        keys[n // 3: n // 3 + n // 4] = keys[0]          # long run of duplicates
        keys[::5] &= 0xFFFF                              # and clustered small keys
    if end_bit == 64:
        keys[::3] |= (1 << 62)
    vals = torch.arange(n, dtype=torch.int32)
    ka, va = keys.to(dev), vals.to(dev)
    kb, vb = torch.empty_like(ka), torch.empty_like(va)
    ws_bytes = int(lib.sc_sort_workspace_bytes(n))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    where = check(lib.sc_radix_sort_pairs(n, ka.data_ptr(), va.data_ptr(), kb.data_ptr(), vb.data_ptr(),
                                          begin_bit, end_bit, ws.data_ptr(), ws_bytes,
                                          torch.cuda.current_stream().cuda_stream), "sort")
    ko, vo = (kb, vb) if where == 0 else (ka, va)
    # the digit actually sorted on: bits [begin_bit, end_bit) as an unsigned number
    digit = (keys >> begin_bit) & ((1 << (end_bit - begin_bit)) - 1) if end_bit - begin_bit < 63 else keys
    _, order = torch.sort(digit, stable=True)
    assert torch.equal(ko.cpu(), keys[order])            # whole keys travel, order is by the digit
    assert torch.equal(vo.cpu(), vals[order])


@pytest.mark.parametrize("n", [1, 33, 4097, 300001])
@pytest.mark.parametrize("begin_bit,end_bit", [(0, 32), (0, 30), (5, 17)])
def test_radix_sort_u32_matches_stable_sort(S, dev, n, begin_bit, end_bit):
    from street_crafter_b200._lib import check, lib
    g = torch.Generator().manual_seed(n + end_bit)
    keys = torch.randint(0, 2 ** 32, (n,), generator=g, dtype=torch.int64)
    if n > 100:
        keys[n // 2: n // 2 + n // 4] = 0xFFFFFFFF       # the "emits nothing" sentinel run
        keys[::7] = keys[0]
    vals = torch.arange(n, dtype=torch.int32)
    ka = keys.to(torch.uint32).view(torch.int32).to(dev) if hasattr(torch, "uint32") else None
    va = vals.to(dev)
    kb, vb = torch.empty_like(ka), torch.empty_like(va)
    ws_bytes = int(lib.sc_sort_workspace_bytes(n))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    where = check(lib.sc_radix_sort_pairs_u32(n, ka.data_ptr(), va.data_ptr(), kb.data_ptr(), vb.data_ptr(),
                                              begin_bit, end_bit, ws.data_ptr(), ws_bytes,
                                              torch.cuda.current_stream().cuda_stream), "sort32")
    ko, vo = (kb, vb) if where == 0 else (ka, va)
    digit = (keys >> begin_bit) & ((1 << (end_bit - begin_bit)) - 1)
    _, order = torch.sort(digit, stable=True)
    assert torch.equal(ko.cpu().view(torch.uint32).to(torch.int64), keys[order])
    assert torch.equal(vo.cpu(), vals[order])


# ------------------------------------------------------------------------------ SH
@pytest.mark.parametrize("K,deg", [(1, 0), (4, 1), (4, 0), (9, 2), (16, 3), (16, 1), (25, 4), (25, 3)])
@pytest.mark.parametrize("M", [1, 33, 5000])
def test_sh_forward_backward(S, dev, K, deg, M):
    g = torch.Generator().manual_seed(K * 100 + deg + M)
    dirs = (torch.randn(1, M, 3, generator=g) * 5).requires_grad_(True)
    coeffs = torch.randn(1, M, K, 3, generator=g).requires_grad_(True)
    masks = torch.rand(1, M, generator=g) > 0.3
    w = torch.randn(1, M, 3, generator=g)
    out_w = G.spherical_harmonics(deg, dirs, coeffs, masks)
    (out_w * w).sum().backward()
    d2 = dirs.detach().to(dev).requires_grad_(True)
    c2 = coeffs.detach().to(dev).requires_grad_(True)
    out = S.spherical_harmonics(deg, d2, c2, masks=masks.to(dev))
    (out * w.to(dev)).sum().backward()
    assert torch.allclose(out.cpu(), out_w, atol=2e-5, rtol=1e-5), (out.cpu() - out_w).abs().max()
    assert (out.cpu()[~masks] == 0).all()
    assert_grad_close(c2.grad, coeffs.grad, "v_coeffs", 1e-4)
    if deg > 0:
        assert_grad_close(d2.grad, dirs.grad, "v_dirs")
    else:
        assert d2.grad.abs().max().item() < 1e-6
    # no mask at all
    out_nm = S.spherical_harmonics(deg, d2.detach(), c2.detach())
    assert torch.allclose(out_nm.cpu(), G.spherical_harmonics(deg, dirs.detach(), coeffs.detach()), atol=2e-5)


def test_sh_golden_from_reference(S, dev):
    """same fixture the CPU suite uses: values produced by the reference's own sh_utils.eval_sh"""
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "sh_eval.npz"))
    dirs = torch.from_numpy(z["dirs"]).to(dev)
    coeffs = torch.from_numpy(z["coeffs"]).to(dev)
    for deg in range(5):
        got = S.spherical_harmonics(deg, dirs, coeffs, None)
        want = torch.from_numpy(z[f"colors_deg{deg}"])
        assert torch.allclose(got.cpu(), want, atol=3e-5, rtol=1e-5), deg


# ------------------------------------------------------------------------------ compositing
def _raster_inputs(n, W, H, seed, scale_mean=0.08, sh_degree=3):
    s = make_gaussians(n, seed=seed, scale_mean=scale_mean, sh_degree=sh_degree)
    V, K = make_camera(W, H)
    rc, ra, aux = G.render_reference_sequence(s["means"], s["quats"], s["scales"], s["opacities"], s["shs"], V, K,
                                              W, H, sh_degree, return_aux=True)
    return s, V, K, rc, ra, aux


@pytest.mark.parametrize("n,W,H,scale", [(20000, 640, 400, 0.08), (3000, 333, 211, 0.5), (200, 48, 32, 1.0)])
def test_rasterize_forward(S, dev, n, W, H, scale):
    s, V, K, rc_w, ra_w, aux = _raster_inputs(n, W, H, seed=n, scale_mean=scale)
    a = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in aux.items()}
    rc, ra = S.rasterize_to_pixels(a["means2d"], a["conics"], a["colors"], a["opacities"], W, H, 16,
                                   a["isect_offsets"], a["flatten_ids"], backgrounds=None, packed=False, absgrad=True)
    assert rc.shape == (1, H, W, 4) and ra.shape == (1, H, W, 1)
    assert_image_close(rc[..., :3], rc_w[..., :3], "rgb", margin=aux["margin"])
    assert_image_close(ra, ra_w, "alpha", margin=aux["margin"])
    dscale = float(aux["depths"].max())
    assert_image_close(rc[..., 3], rc_w[..., 3], "depth-weighted", scale=max(1.0, dscale), margin=aux["margin"])
    assert float(ra.max()) > 0.3
    # backgrounds
    bg = torch.tensor([[0.2, 0.4, 0.6, 0.0]])
    rcb, _ = S.rasterize_to_pixels(a["means2d"], a["conics"], a["colors"], a["opacities"], W, H, 16,
                                   a["isect_offsets"], a["flatten_ids"], backgrounds=bg.to(dev))
    want = rc_w + bg[:, None, None, :] * (1 - ra_w)
    assert_image_close(rcb[..., :3], want[..., :3], "rgb+bg")


@pytest.mark.parametrize("CH", [1, 3, 5])
def test_rasterize_channel_counts(S, dev, CH):
    s, V, K, rc_w, ra_w, aux = _raster_inputs(2500, 200, 120, seed=9, scale_mean=0.4)
    g = torch.Generator().manual_seed(CH)
    cols = torch.rand(1, 2500, CH, generator=g)
    want, wa = G.rasterize_to_pixels(aux["means2d"], aux["conics"], cols, aux["opacities"], 200, 120, 16,
                                     aux["isect_offsets"], aux["flatten_ids"])
    got, ga = S.rasterize_to_pixels(aux["means2d"].to(dev), aux["conics"].to(dev), cols.to(dev),
                                    aux["opacities"].to(dev), 200, 120, 16, aux["isect_offsets"].to(dev),
                                    aux["flatten_ids"].to(dev))
    assert_image_close(got, want, f"CH={CH}")
    assert_image_close(ga, wa, "alpha")


@pytest.mark.parametrize("n,W,H,scale,bg", [(6000, 320, 200, 0.25, False), (800, 100, 75, 0.8, True)])
def test_rasterize_backward(S, dev, n, W, H, scale, bg):
    s, V, K, _, _, aux = _raster_inputs(n, W, H, seed=n + 1, scale_mean=scale)
    g = torch.Generator().manual_seed(0)
    v_rc = torch.randn(1, H, W, 4, generator=g)
    v_ra = torch.randn(1, H, W, 1, generator=g)
    bgt = torch.tensor([[0.3, 0.1, 0.7, 0.5]]) if bg else None
    names = ("means2d", "conics", "colors", "opacities")
    leaves = [aux[k].detach().clone().requires_grad_(True) for k in names]
    rc, ra = G.rasterize_to_pixels(*leaves, W, H, 16, aux["isect_offsets"], aux["flatten_ids"], backgrounds=bgt)
    ((rc * v_rc).sum() + (ra * v_ra).sum()).backward()
    cl = [aux[k].detach().to(dev).requires_grad_(True) for k in names]
    rc2, ra2 = S.rasterize_to_pixels(*cl, W, H, 16, aux["isect_offsets"].to(dev), aux["flatten_ids"].to(dev),
                                     backgrounds=None if bgt is None else bgt.to(dev), absgrad=True)
    ((rc2 * v_rc.to(dev)).sum() + (ra2 * v_ra.to(dev)).sum()).backward()
    for nm, a, b in zip(names, cl, leaves):
        assert_grad_close(a.grad, b.grad, f"v_{nm}")
    ag = cl[0].absgrad
    assert ag.shape == (1, n, 2) and (ag >= 0).all()
    assert (ag.cpu() + 1e-6 >= cl[0].grad.abs().cpu() * (1 - 1e-4)).all(), "absgrad >= |grad|"


def test_absgrad_matches_per_pixel_sum(S, dev):
    """absgrad := sum over pixels of |d L_pixel / d means2d|, checked by per-pixel autograd."""
    n, W, H = 40, 16, 16
    s, V, K, _, _, aux = _raster_inputs(n, W, H, seed=77, scale_mean=2.0, sh_degree=1)
    g = torch.Generator().manual_seed(1)
    v_rc = torch.randn(1, H, W, 4, generator=g)
    v_ra = torch.randn(1, H, W, 1, generator=g)
    m2 = aux["means2d"].detach().clone().requires_grad_(True)
    rc, ra = G.rasterize_to_pixels(m2, aux["conics"], aux["colors"], aux["opacities"], W, H, 16,
                                   aux["isect_offsets"], aux["flatten_ids"])
    per_pix = (rc * v_rc).sum(-1) + (ra * v_ra).sum(-1)  # [1,H,W]
    want = torch.zeros(1, n, 2)
    for i in range(H):
        for j in range(W):
            (gr,) = torch.autograd.grad(per_pix[0, i, j], m2, retain_graph=True)
            want += gr.abs()
    cl = aux["means2d"].detach().to(dev).requires_grad_(True)
    rc2, ra2 = S.rasterize_to_pixels(cl, aux["conics"].to(dev), aux["colors"].to(dev), aux["opacities"].to(dev), W,
                                     H, 16, aux["isect_offsets"].to(dev), aux["flatten_ids"].to(dev), absgrad=True)
    ((rc2 * v_rc.to(dev)).sum() + (ra2 * v_ra.to(dev)).sum()).backward()
    assert want.abs().max() > 0
    assert_grad_close(cl.absgrad, want, "absgrad")


# ------------------------------------------------------------------------------ whole pipeline
def _pipeline_gpu(S, dev, s, V, K, W, H, deg, train=True, fused=True):
    from street_crafter_b200.renderer import render_kernel_gsplat
    c = {k: v.to(dev) for k, v in s.items()}
    for k in c:
        c[k].requires_grad_(train)
    cam_center = (-(V[:3, :3].T @ V[:3, 3])).to(dev)
    out = render_kernel_gsplat(c["means"], c["shs"], c["scales"], c["quats"], c["opacities"][:, None],
                               V[None].to(dev), K[None].to(dev), cam_center, W, H, deg, train=train, fused=fused)
    return c, out


@pytest.mark.parametrize("fused", [True, False])
def test_full_sequence_cfg1_forward_and_backward(S, dev, fused):
    """BASELINE.json configs[0]: 100k Gaussians, SH deg 3, one 960x640 camera, fwd+bwd; through
    the single-node fused path and through the reference's five-op sequence."""
    n, W, H, deg = 100000, 960, 640, 3
    s = make_gaussians(n, sh_degree=deg, seed=0)
    V, K = make_camera(W, H)
    leaves = {k: v.clone().requires_grad_(True) for k, v in s.items()}
    rc_w, ra_w, aux_w = G.render_reference_sequence(leaves["means"], leaves["quats"], leaves["scales"],
                                                    leaves["opacities"], leaves["shs"], V, K, W, H, deg,
                                                    return_aux=True)
    g = torch.Generator().manual_seed(0)
    v_rc = torch.randn(1, H, W, 4, generator=g)
    v_ra = torch.randn(1, H, W, 1, generator=g)
    v_rc[..., 3] *= 0.02  # depth channel is O(50): keep its share of the loss comparable
    ((rc_w * v_rc).sum() + (ra_w * v_ra).sum()).backward()

    c, out = _pipeline_gpu(S, dev, s, V, K, W, H, deg, fused=fused)
    aux = out["aux"]
    for k in ("radii", "tiles_per_gauss", "isect_ids", "flatten_ids", "isect_offsets"):
        assert torch.equal(aux[k].cpu(), aux_w[k]), f"{k} not bit-exact"
    assert_image_close(out["render_colors"][..., :3], rc_w[..., :3].detach(), "rgb", margin=aux_w["margin"])
    assert_image_close(out["render_alphas"], ra_w.detach(), "alpha", margin=aux_w["margin"])
    assert_image_close(out["render_colors"][..., 3], rc_w[..., 3].detach(), "depth", scale=100.0, margin=aux_w["margin"])
    ((out["render_colors"] * v_rc.to(dev)).sum() + (out["render_alphas"] * v_ra.to(dev)).sum()).backward()
    for k in ("means", "quats", "scales", "opacities", "shs"):
        assert_grad_close(c[k].grad, torch.nan_to_num(leaves[k].grad), f"v_{k}")
    vp = out["viewspace_points"]
    assert vp.grad is not None and vp.absgrad.shape == vp.shape      # retain_grad + absgrad side effect
    assert out["rgb"].shape == (3, H, W) and out["acc"].shape == (1, H, W) and out["depth"].shape == (1, H, W)
    assert out["visibility_filter"].dtype == torch.bool and out["radii"].dtype == torch.float32


@pytest.mark.parametrize("yaw", [45.0, -45.0])
def test_cfg4_side_camera_1920x886_vs_oracle(S, dev, yaw):
    """BASELINE.json configs[3]: the Waymo SIDE cameras are 1920x886 — a height that is NOT a multiple of the
    16-pixel tile (56 tile rows, the last one 6 pixels tall).  Whole fused path, forward + backward, against
    the oracle at 100 k Gaussians."""
    n, W, H, deg = 100000, 1920, 886, 3
    s = make_gaussians(n, sh_degree=deg, seed=44)
    V, K = make_camera(W, H, yaw_deg=yaw, z_shift=3.0)
    leaves = {k: v.clone().requires_grad_(True) for k, v in s.items()}
    rc_w, ra_w, aux_w = G.render_reference_sequence(leaves["means"], leaves["quats"], leaves["scales"],
                                                    leaves["opacities"], leaves["shs"], V, K, W, H, deg, return_aux=True)
    g = torch.Generator().manual_seed(4)
    v_rc = torch.randn(1, H, W, 4, generator=g)
    v_ra = torch.randn(1, H, W, 1, generator=g)
    v_rc[..., 3] *= 0.02
    ((rc_w * v_rc).sum() + (ra_w * v_ra).sum()).backward()
    c, out = _pipeline_gpu(S, dev, s, V, K, W, H, deg, fused=True)
    aux = out["aux"]
    assert aux["isect_offsets"].shape == (1, 56, 120)
    for k in ("radii", "tiles_per_gauss", "isect_ids", "flatten_ids", "isect_offsets"):
        assert torch.equal(aux[k].cpu(), aux_w[k]), f"{k} not bit-exact"
    assert_image_close(out["render_colors"][..., :3], rc_w[..., :3].detach(), "cfg4 rgb", margin=aux_w["margin"])
    assert_image_close(out["render_alphas"], ra_w.detach(), "cfg4 alpha", margin=aux_w["margin"])
    assert_image_close(out["render_colors"][..., 3], rc_w[..., 3].detach(), "cfg4 depth", scale=100.0, margin=aux_w["margin"])
    ((out["render_colors"] * v_rc.to(dev)).sum() + (out["render_alphas"] * v_ra.to(dev)).sum()).backward()
    for k in ("means", "quats", "scales", "opacities", "shs"):
        assert_grad_close(c[k].grad, torch.nan_to_num(leaves[k].grad), f"cfg4 v_{k}")


@pytest.mark.parametrize("deg,antialiasing", [(3, True), (1, True), (3, False)])
def test_fused_view_equals_five_op_sequence(S, dev, deg, antialiasing):
    """the single-node path computes the same tensors as the op-by-op sequence: forward
    bit-identical (same kernels, the glue folded in), gradients equal to rounding"""
    from street_crafter_b200.renderer import render_kernel_gsplat
    n, W, H = 40000, 640, 400
    s = make_gaussians(n, sh_degree=deg, seed=11)
    V, K = make_camera(W, H, yaw_deg=10.0)
    g = torch.Generator().manual_seed(2)
    v_rc = torch.randn(1, H, W, 4, generator=g).to(dev)
    v_ra = torch.randn(1, H, W, 1, generator=g).to(dev)
    v_rc[..., 3] *= 0.02
    res = []
    for fused in (False, True):
        c = {k: v.to(dev).requires_grad_(True) for k, v in s.items()}
        cc = (-(V[:3, :3].T @ V[:3, 3])).to(dev)
        out = render_kernel_gsplat(c["means"], c["shs"], c["scales"], c["quats"], c["opacities"][:, None],
                                   V[None].to(dev), K[None].to(dev), cc, W, H, deg, antialiasing=antialiasing,
                                   fused=fused)
        torch.autograd.backward([out["render_colors"], out["render_alphas"]], [v_rc, v_ra])
        res.append((c, out))
    (c0, o0), (c1, o1) = res
    for k in ("radii", "means2d", "depths", "conics", "isect_ids", "flatten_ids", "isect_offsets", "colors",
              "opacities"):
        assert torch.equal(o0["aux"][k], o1["aux"][k]), k
    assert torch.equal(o0["render_colors"], o1["render_colors"]) and torch.equal(o0["render_alphas"],
                                                                                 o1["render_alphas"])
    for k in ("means", "quats", "scales", "opacities", "shs"):
        assert_grad_close(c1[k].grad, c0[k].grad.cpu(), f"fused v_{k}")
    # the renderer's result packing (depth / clamp(alpha), radii > 0, radii / max(H, W)) is written by
    # the kernels on the single-node path and by torch ops on the op-by-op path: identical bits
    for k in ("rgb", "acc", "depth", "visibility_filter", "radii"):
        assert o0[k].shape == o1[k].shape and o0[k].dtype == o1[k].dtype, k
        assert torch.equal(o0[k], o1[k]), k
    vp0, vp1 = o0["viewspace_points"], o1["viewspace_points"]
    assert_grad_close(vp1.grad, vp0.grad.cpu(), "viewspace grad")
    assert_grad_close(vp1.absgrad, vp0.absgrad.cpu(), "viewspace absgrad")


def test_fused_view_depth_loss_gradient(S, dev):
    """a loss on the renderer's normalised depth map flows through the in-kernel division exactly
    as through the torch ops of the op-by-op path"""
    from street_crafter_b200.renderer import render_kernel_gsplat
    n, W, H, deg = 20000, 320, 200, 3
    s = make_gaussians(n, sh_degree=deg, seed=13)
    V, K = make_camera(W, H, yaw_deg=-5.0)
    g = torch.Generator().manual_seed(3)
    w_depth = torch.rand(1, H, W, generator=g).to(dev)
    w_rgb = torch.randn(3, H, W, generator=g).to(dev)
    res = []
    for fused in (False, True):
        c = {k: v.to(dev).requires_grad_(True) for k, v in s.items()}
        cc = (-(V[:3, :3].T @ V[:3, 3])).to(dev)
        out = render_kernel_gsplat(c["means"], c["shs"], c["scales"], c["quats"], c["opacities"][:, None],
                                   V[None].to(dev), K[None].to(dev), cc, W, H, deg, fused=fused)
        loss = (out["depth"] * w_depth).sum() * 1e-3 + (out["rgb"] * w_rgb).sum() + out["acc"].sum() * 0.1
        loss.backward()
        res.append((c, out, loss.item()))
    (c0, o0, l0), (c1, o1, l1) = res
    assert abs(l0 - l1) <= 1e-5 * abs(l0)
    for k in ("means", "quats", "scales", "opacities", "shs"):
        assert_grad_close(c1[k].grad, c0[k].grad.cpu(), f"depth-loss v_{k}")


@pytest.mark.parametrize("fused", [True, False])
def test_grad_sinks_match_autograd_accumulation(S, dev, fused):
    """gradient-arena mode (projection / SH backward += into the flat arena) == plain autograd
    accumulation over several views"""
    from street_crafter_b200.parallel import GaussianArena
    from street_crafter_b200.renderer import render_kernel_gsplat
    n, W, H, deg = 30000, 640, 400, 3
    scene = make_gaussians(n, sh_degree=deg, seed=5)
    cams = [make_camera(W, H, yaw_deg=y) for y in (0.0, 45.0, -45.0)]
    g = torch.Generator().manual_seed(3)
    v_rc = [torch.randn(1, H, W, 4, generator=g).to(dev) for _ in cams]
    v_ra = [torch.randn(1, H, W, 1, generator=g).to(dev) for _ in cams]
    grads = []
    for use_sinks in (False, True):
        arena = GaussianArena(scene, dev)
        arena.zero_grad()
        sinks = arena.sinks() if use_sinks else None
        for (V, K), a, b in zip(cams, v_rc, v_ra):
            cc = (-(V[:3, :3].T @ V[:3, 3])).to(dev)
            out = render_kernel_gsplat(arena.means, arena.shs, arena.scales, arena.quats, arena.opacities,
                                       V[None].to(dev), K[None].to(dev), cc, W, H, deg, grad_sinks=sinks,
                                       fused=fused)
            torch.autograd.backward([out["render_colors"], out["render_alphas"]], [a, b])
        for name, off, numel, shape in arena.layout:
            assert getattr(arena, name).grad.data_ptr() == arena.grads.data_ptr() + 4 * off, name
        grads.append(arena.grads.clone())
    ref, got = grads
    assert ref.abs().max() > 0
    # raster-backward atomics reorder sums between runs: compare at gradient tolerance
    assert_grad_close(got, ref.cpu(), "arena grads via sinks")


def test_rasterization_entry_matches_sequence(S, dev):
    n, W, H = 20000, 640, 384
    s = make_gaussians(n, seed=4)
    V, K = make_camera(W, H)
    c, out = _pipeline_gpu(S, dev, s, V, K, W, H, 3, train=False)
    rc, ra, meta = S.rasterization(c["means"], c["quats"], c["scales"], c["opacities"], c["shs"], V[None].to(dev),
                                   K[None].to(dev), W, H, near_plane=0.001, far_plane=1000.0, sh_degree=3,
                                   render_mode="RGB+ED", rasterize_mode="antialiased", absgrad=True, packed=False)
    assert torch.equal(meta["isect_ids"], out["aux"]["isect_ids"])
    assert torch.allclose(rc[..., :3], out["render_colors"][..., :3], atol=1e-5)
    assert torch.allclose(ra, out["render_alphas"], atol=1e-6)
    assert torch.allclose(rc[0, ..., 3], out["depth"][0], atol=1e-4)


# ------------------------------------------------------------------------------ size-independent properties
def _assert_binning_equals_oracle(S, means2d, radii, depths, tile, W, H, tpg, ids, flat, off):
    """FULL-SIZE bit-exact binning: the fused pipeline's sorted list against (1) the CPU oracle's isect_tiles /
    isect_offset_encode on the same projected inputs and (2) a stable torch.sort on the device of this library's
    own UNSORTED emission (a different set of kernels: count, 3-kernel scan, entry-order emit)"""
    tw, th = math.ceil(W / float(tile)), math.ceil(H / float(tile))
    w_tpg, w_ids, w_flat = G.isect_tiles(means2d.cpu(), radii.cpu(), depths.cpu(), tile, tw, th, n_cameras=radii.shape[0])
    assert torch.equal(tpg.cpu(), w_tpg), "tiles_per_gauss"
    assert torch.equal(ids.cpu(), w_ids), "sorted isect_ids"
    assert torch.equal(flat.cpu(), w_flat), "flatten_ids"
    assert torch.equal(off.cpu(), G.isect_offset_encode(w_ids, radii.shape[0], tw, th)), "isect_offsets"
    _, ids_u, flat_u = S.isect_tiles(means2d, radii, depths, tile, tw, th, sort=False)
    ids_s, order = torch.sort(ids_u, stable=True)
    assert torch.equal(ids, ids_s) and torch.equal(flat, flat_u[order]), "stable sort of the unsorted emission"


def test_cfg2_scale_properties(S, dev):
    """BASELINE.json configs[1] shape (2M Gaussians, 1920x1280): structural invariants +
    linearity in colour + determinism of the forward pass."""
    n, W, H = 2_000_000, 1920, 1280
    s = make_gaussians(n, seed=0)
    V, K = make_camera(W, H)
    c, out = _pipeline_gpu(S, dev, s, V, K, W, H, 3, train=False)
    aux = out["aux"]
    ids, flat, off, tpg = aux["isect_ids"], aux["flatten_ids"], aux["isect_offsets"], aux["tiles_per_gauss"]
    I = ids.numel()
    assert I == int(tpg.sum()) and I > 1_000_000
    assert bool((ids[1:] >= ids[:-1]).all()), "keys sorted"
    tile = (ids >> 32)
    offs = torch.cat([off.reshape(-1).long(), torch.tensor([I], device=dev)])
    assert bool((offs[1:] >= offs[:-1]).all())
    counts = torch.bincount(tile, minlength=off.numel())
    assert torch.equal(counts, offs[1:] - offs[:-1]), "offsets == per-tile histogram"
    # stable tie-break: equal keys keep ascending flatten id
    same = ids[1:] == ids[:-1]
    assert bool((flat[1:][same] > flat[:-1][same]).all())
    # depth bits of every key equal the depth of its Gaussian
    assert torch.equal((ids & 0xFFFFFFFF).to(torch.int32).view(torch.float32), aux["depths"][0][flat.long()])
    # checksum of the multiset of values survives the sort
    assert int(flat.long().sum()) == int((tpg[0].long() * torch.arange(n, device=dev)).sum())
    _assert_binning_equals_oracle(S, aux["means2d"], aux["radii"], aux["depths"], 16, W, H, tpg, ids, flat, off)
    ra = out["render_alphas"]
    assert float(ra.min()) >= 0 and float(ra.max()) <= 1.0
    a = aux
    rc1, ra1 = S.rasterize_to_pixels(a["means2d"], a["conics"], a["colors"], a["opacities"], W, H, 16,
                                     a["isect_offsets"], a["flatten_ids"])
    rc2, ra2 = S.rasterize_to_pixels(a["means2d"], a["conics"], a["colors"] * 2.0, a["opacities"], W, H, 16,
                                     a["isect_offsets"], a["flatten_ids"])
    assert torch.equal(rc1, out["render_colors"]) and torch.equal(ra1, ra), "forward is deterministic"
    assert torch.equal(rc2, rc1 * 2.0) and torch.equal(ra2, ra1), "compositing is linear in colour"


# ------------------------------------------------------------------------------ KNN
@pytest.mark.parametrize("P", [1, 2, 3, 4, 5, 31, 32, 33, 1000, 20000])
def test_knn_vs_bruteforce(S, dev, P):
    g = torch.Generator().manual_seed(P)
    pts = torch.randn(P, 3, generator=g) * torch.tensor([20.0, 3.0, 40.0])
    if P > 100:
        pts[50] = pts[3]
        pts[51] = pts[3]
    got = S.distCUDA2(pts.to(dev)).cpu()
    want = KO.dist_knn3_bruteforce(pts)
    assert got.shape == (P,) and got.dtype == torch.float32
    if P < 4:
        assert (got > 1e37).all() and (want > 1e37).all()
    else:
        assert torch.allclose(got, want, rtol=2e-6, atol=1e-12), (got - want).abs().max()


def test_knn_large_vs_kdtree(S, dev):
    xyz, _ = make_lidar_points(300000, seed=2)
    xyz[:1000] = xyz[1000:2000] * 1.0  # block of exact duplicates
    got = S.distCUDA2(xyz.to(dev)).cpu()
    want = KO.dist_knn3_kdtree(xyz)
    assert torch.allclose(got, want, rtol=1e-4, atol=1e-9), (got - want).abs().max()


def test_knn_vs_compiled_reference(S, dev):
    """oracle/_ref/libsimpleknn_ref.so = the reference's own simple_knn.cu compiled for sm_100a
    (oracle/build_ref.sh).  Skipped when it was not built."""
    import ctypes
    path = os.path.join(os.path.dirname(os.path.dirname(__file__)), "oracle", "_ref", "libsimpleknn_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref not built")
    ref = ctypes.CDLL(path)
    ref.ref_knn.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    xyz, _ = make_lidar_points(200000, seed=5)
    p = xyz.to(dev).contiguous()
    want = torch.zeros(p.shape[0], device=dev)
    torch.cuda.synchronize()
    ref.ref_knn(p.shape[0], p.data_ptr(), want.data_ptr())
    torch.cuda.synchronize()
    got = S.distCUDA2(p)
    assert torch.allclose(got, want, rtol=1e-6, atol=0), (got - want).abs().max()
    exact = (got == want).float().mean().item()
    assert exact > 0.99, f"only {exact:.4f} bit-identical to the reference kernel"


def test_cfg5_knn_5m_vs_compiled_reference_full_size(S, dev):
    """BASELINE configs[4]: distCUDA2 on 5 M points against the reference's own kernel on ALL of them (the subset test
    above checks 512 queries against brute force)"""
    import ctypes
    path = os.path.join(os.path.dirname(os.path.dirname(__file__)), "oracle", "_ref", "libsimpleknn_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref not built")
    ref = ctypes.CDLL(path)
    ref.ref_knn.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    xyz, _ = make_lidar_points(5_000_000, seed=5)
    p = xyz.to(dev).contiguous()
    want = torch.zeros(p.shape[0], device=dev)
    torch.cuda.synchronize()
    ref.ref_knn(p.shape[0], p.data_ptr(), want.data_ptr())
    torch.cuda.synchronize()
    got = S.distCUDA2(p)
    assert torch.allclose(got, want, rtol=1e-6, atol=0), (got - want).abs().max()
    assert (got == want).float().mean().item() > 0.99


# ------------------------------------------------------------------------------ LiDAR point splat
def test_point_rasterizer_max_hit(S, dev):
    N, W, H = 20000, 256, 144
    xyz, rgb = make_lidar_points(N, seed=3)
    V, K = make_camera(W, H)
    fx = float(K[0, 0])
    z = (xyz @ V[:3, :3].T + V[:3, 3])[:, 2:3]
    radius = 0.01 * z / fx * 0.5 * min(H, W) * 4.0  # a few px so that splats overlap
    st = S.PointRasterizationSettings(
        image_height=H, image_width=W, tanfovx=W / (2 * fx), tanfovy=H / (2 * fx), bg=torch.zeros(3),
        scale_modifier=1.0, viewmatrix=V.T.contiguous(), projmatrix=torch.eye(4), sh_degree=0, max_hit=10,
        campos=torch.zeros(3), prefiltered=False, debug=False)
    img, depth, alpha, radii = S.PointRasterizer(st)(xyz.to(dev), torch.zeros_like(xyz).to(dev), rgb.to(dev),
                                                     torch.ones(N, 1, device=dev), radius.to(dev))
    assert img.shape == (3, H, W) and depth.shape == (1, H, W) and alpha.shape == (1, H, W) and radii.shape == (N,)
    # oracle with the same restated semantics
    quats = torch.zeros(N, 4); quats[:, 0] = 1
    r, m2, dp, cn, _ = G.fully_fused_projection(xyz, None, quats, radius.expand(N, 3).contiguous(), V[None], K[None],
                                                W, H, near_plane=1.0, far_plane=100.0)
    assert torch.equal(radii.cpu(), r[0])
    tw, th = math.ceil(W / 16), math.ceil(H / 16)
    _, ids, flat = G.isect_tiles(m2, r, dp, 16, tw, th)
    off = G.isect_offset_encode(ids, 1, tw, th)
    cols = torch.cat([rgb[None], dp[..., None]], -1)
    want, wa = G.rasterize_to_pixels(m2, cn, cols, torch.ones(1, N), W, H, 16, off, flat, max_hit=10)
    assert_image_close(img.permute(1, 2, 0), want[0, ..., :3], "lidar rgb")
    assert_image_close(alpha[0], wa[0, ..., 0], "lidar alpha")
    nolimit, _ = G.rasterize_to_pixels(m2, cn, cols, torch.ones(1, N), W, H, 16, off, flat)
    assert (nolimit - want).abs().max() > 1e-3, "max_hit must matter in this scene"


# ------------------------------------------------------------------------------ photometric loss (N3)
LOSS_TOL = 2e-5   # fp32 kernel vs float64 reference values (loss terms are O(0.1..1))


def test_photometric_loss_vs_reference_golden(S, dev):
    """fused L1 + SSIM kernels against the reference's own l1_loss / ssim outputs
    (tests/golden/loss_l1_ssim.npz, made by importing /root/reference/.../loss_utils.py)"""
    import numpy as np
    from street_crafter_b200 import losses as L
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "loss_l1_ssim.npz"))
    for name in ("a", "b", "c"):
        img = torch.from_numpy(z[f"{name}_img"]).float().to(dev).requires_grad_(True)
        gt = torch.from_numpy(z[f"{name}_gt"]).float().to(dev)
        mask = torch.from_numpy(z[f"{name}_mask"]).to(dev) if z[f"{name}_mask"].size else None
        loss, l1, s = L.photometric_loss(img, gt, mask, lambda_dssim=0.2, lambda_l1=1.0)
        loss.backward()
        assert abs(float(l1) - float(z[f"{name}_l1"])) < LOSS_TOL, name
        assert abs(float(s) - float(z[f"{name}_ssim"])) < LOSS_TOL, name
        assert abs(float(loss.detach()) - float(z[f"{name}_loss"])) < LOSS_TOL, name
        want = torch.from_numpy(z[f"{name}_grad"])
        assert_grad_close(img.grad, want, f"loss grad {name}")
        # the reference's names
        assert abs(float(L.l1_loss(img.detach(), gt, mask)) - float(z[f"{name}_l1"])) < LOSS_TOL
        assert abs(float(L.ssim(img.detach(), gt, mask=mask)) - float(z[f"{name}_ssim"])) < LOSS_TOL


@pytest.mark.parametrize("H,W,use_mask", [(97, 131, True), (128, 160, False), (5, 7, True)])
def test_photometric_loss_vs_oracle_and_layouts(S, dev, H, W, use_mask):
    """ragged sizes (partial tiles, images smaller than the window), the strided `rgb` view of
    the compositor's output and the direct [1,H,W,4] entry all give the oracle's values"""
    from oracle import loss_torch as LO
    from street_crafter_b200 import losses as L
    g = torch.Generator().manual_seed(H * 1000 + W)
    rc = torch.rand(1, H, W, 4, generator=g)
    gt = (0.5 * torch.rand(3, H, W, generator=g) + 0.5 * rc[0, ..., :3].permute(2, 0, 1)).contiguous()
    mask = (torch.rand(1, H, W, generator=g) > 0.25) if use_mask else None
    img_w = rc[0, ..., :3].permute(2, 0, 1).double().requires_grad_(True)
    loss_w, l1_w, s_w = LO.photometric_loss(img_w, gt.double(), mask, 0.2, 1.0)
    (3.0 * loss_w).backward()
    md = mask.to(dev) if mask is not None else None
    # (1) strided view, as the renderer returns it
    rc1 = rc.to(dev).requires_grad_(True)
    img = rc1[..., :-1][0].permute(2, 0, 1)
    loss, l1, s = L.photometric_loss(img, gt.to(dev), md, 0.2, 1.0)
    (3.0 * loss).backward()
    assert abs(float(loss.detach()) - float(loss_w)) < LOSS_TOL and abs(float(s) - float(s_w)) < LOSS_TOL
    assert abs(float(l1) - float(l1_w)) < LOSS_TOL
    assert_grad_close(rc1.grad[0, ..., :3].permute(2, 0, 1), img_w.grad, "grad via rgb view")
    assert (rc1.grad[..., 3] == 0).all()
    # (2) direct entry on render_colors
    rc2 = rc.to(dev).requires_grad_(True)
    loss2, _, _ = L.photometric_loss_render(rc2, gt.to(dev), md, 0.2, 1.0)
    (3.0 * loss2).backward()
    assert abs(float(loss2.detach()) - float(loss_w)) < LOSS_TOL
    assert_grad_close(rc2.grad[0, ..., :3].permute(2, 0, 1), img_w.grad, "grad via render_colors")
    assert (rc2.grad[..., 3] == 0).all()
    # (3) contiguous [3,H,W]
    img3 = rc[0, ..., :3].permute(2, 0, 1).contiguous().to(dev).requires_grad_(True)
    loss3, _, _ = L.photometric_loss(img3, gt.to(dev), md, 0.2, 1.0)
    (3.0 * loss3).backward()
    assert_grad_close(img3.grad, img_w.grad, "grad contiguous")


# ------------------------------------------------------------------------ densification statistics (N3)
@pytest.mark.gpu
def test_densification_stats_vs_reference_golden(S, dev):
    import numpy as np
    from street_crafter_b200.densify import densification_stats
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "densify_stats.npz"))
    f = lambda k: torch.from_numpy(z[k]).float().to(dev)
    vp = torch.zeros(1, z["grad"].shape[1], 2, device=dev)
    vp.grad, vp.absgrad = f("grad"), f("absgrad")
    acc, den, mr = f("accum_in").contiguous(), f("denom_in").contiguous(), f("max_radii2D_in").contiguous()
    densification_stats(vp, torch.from_numpy(z["vis"]).to(dev), f("radii"), int(z["W"]), int(z["H"]), acc, den, mr)
    assert torch.allclose(acc.cpu().double(), torch.from_numpy(z["accum_out"]), rtol=2e-6, atol=1e-7)
    assert torch.equal(den.cpu().double(), torch.from_numpy(z["denom_out"]))
    assert torch.equal(mr.cpu(), torch.from_numpy(z["max_radii2D_out"]).float())


@pytest.mark.gpu
def test_densification_stats_after_render_vs_oracle(S, dev):
    """the call sequence of train.py:283-284 on a real render: statistics from the renderer's own
    viewspace_points / visibility_filter / radii equal the oracle's, untouched where invisible"""
    from oracle.densify_torch import densification_stats as oracle_stats
    from street_crafter_b200.densify import densification_stats
    from street_crafter_b200.renderer import render_kernel_gsplat
    n, W, H, deg = 30000, 480, 320, 3
    s = make_gaussians(n, sh_degree=deg, seed=17)
    V, K = make_camera(W, H, yaw_deg=8.0)
    c = {k: v.to(dev).requires_grad_(True) for k, v in s.items()}
    cc = (-(V[:3, :3].T @ V[:3, 3])).to(dev)
    out = render_kernel_gsplat(c["means"], c["shs"], c["scales"], c["quats"], c["opacities"][:, None],
                               V[None].to(dev), K[None].to(dev), cc, W, H, deg)
    (out["rgb"].square().sum() + out["acc"].sum()).backward()
    g = torch.Generator().manual_seed(4)
    acc0, den0, mr0 = torch.rand(n, 2, generator=g), torch.randint(0, 3, (n, 1), generator=g).float(), torch.rand(n, generator=g) * 0.01
    acc, den, mr = acc0.to(dev), den0.to(dev), mr0.to(dev)
    vp = out["viewspace_points"]
    half = n // 2  # two "models" of the scene graph: rows [0, half) and [half, n)
    densification_stats(vp, out["visibility_filter"], out["radii"], W, H, acc[:half], den[:half], mr[:half], 0, half)
    densification_stats(vp, out["visibility_filter"], out["radii"], W, H, acc[half:], den[half:], mr[half:], half, n)
    w_acc, w_den, w_mr = oracle_stats(vp.grad.cpu(), vp.absgrad.cpu(), out["visibility_filter"].cpu(), out["radii"].cpu(),
                                      W, H, acc0.clone(), den0.clone(), mr0.clone())
    assert out["visibility_filter"].sum().item() > 1000
    assert torch.allclose(acc.cpu(), w_acc, rtol=2e-6, atol=1e-9)
    assert torch.equal(den.cpu(), w_den) and torch.equal(mr.cpu(), w_mr)


# ------------------------------------------------------- BASELINE configs[2] / configs[4] at FULL size
def test_cfg3_lidar_frame_scale_properties(S, dev):
    """configs[2]: 5 M LiDAR points splatted to a 1024x576 frame with max_hit = 10 — too large for the
    oracle, so size-independent properties: binning structure, alpha range, the hit cap only ever removes
    contributions, and a cap above the longest per-pixel list changes nothing."""
    from street_crafter_b200 import rendering as R
    P, W, H = 5_000_000, 1024, 576
    xyz, rgb = make_lidar_points(P, seed=3)
    xyz, rgb = xyz.to(dev), rgb.to(dev)
    V, K = make_camera(W, H)
    V, K = V.to(dev), K.to(dev)
    quats = torch.zeros(P, 4, device=dev); quats[:, 0] = 1
    scales = torch.full((P, 3), 0.02, device=dev)
    radii, m2, dep, con, _ = R.fully_fused_projection(xyz, None, quats, scales, V[None], K[None], W, H,
                                                      near_plane=1.0, far_plane=100.0)
    tw, th = math.ceil(W / 16), math.ceil(H / 16)
    tpg, ids, flat = R.isect_tiles(m2, radii, dep, 16, tw, th, n_cameras=1)
    off = R.isect_offset_encode(ids, 1, tw, th)
    I = ids.numel()
    assert I == int(tpg.sum()) and I > 500_000
    assert bool((ids[1:] >= ids[:-1]).all())
    offs = torch.cat([off.reshape(-1).long(), torch.tensor([I], device=dev)])
    assert torch.equal(torch.bincount(ids >> 32, minlength=off.numel()), offs[1:] - offs[:-1])
    assert bool((radii[0][flat.long()] > 0).all()), "only visible points are binned"
    _assert_binning_equals_oracle(S, m2, radii, dep, 16, W, H, tpg, ids, flat, off)
    cols = torch.cat([rgb, dep[0][:, None]], -1)[None]
    op = torch.ones(1, P, device=dev)
    outs = {}
    for mh in (0, 10, 1 << 20):
        outs[mh] = R.rasterize_to_pixels(m2, con, cols, op, W, H, 16, off, flat, max_hit=mh)
    (rc0, ra0), (rc10, ra10), (rcb, rab) = outs[0], outs[10], outs[1 << 20]
    for t in (rc0, ra0, rc10, ra10):
        assert bool(torch.isfinite(t).all())
    assert float(ra10.min()) >= 0.0 and float(ra10.max()) <= 1.0
    assert bool((ra10 <= ra0 + 1e-6).all()), "the hit cap only removes contributions"
    assert bool((ra10 < ra0 - 1e-3).any()), "and it does bite somewhere at this density"
    assert torch.equal(rcb, rc0) and torch.equal(rab, ra0), "a cap above the longest list is the uncapped render"
    # front-to-back depth: the blended depth lies inside the near/far range wherever the pixel is covered
    cov = ra10[0, ..., 0] > 0.5
    d = (rc10[0, ..., 3] / ra10[0, ..., 0].clamp(min=1e-10))[cov]
    assert cov.float().mean().item() > 0.2 and float(d.min()) >= 1.0 - 1e-3 and float(d.max()) <= 100.0 + 1e-3


def test_cfg5_knn_5m_subset_exact(S, dev):
    """configs[4]: distCUDA2 on 5 M points; 512 random queries are checked against an exhaustive search over
    all 5 M points (the property every output must satisfy: mean squared distance to the 3 nearest OTHER points)"""
    P = 5_000_000
    xyz, _ = make_lidar_points(P, seed=5)
    p = xyz.to(dev)
    got = S.distCUDA2(p)
    assert got.shape == (P,) and bool(torch.isfinite(got).all()) and float(got.min()) >= 0.0
    g = torch.Generator().manual_seed(9)
    q = torch.randint(0, P, (512,), generator=g).to(dev)
    best = torch.full((512, 4), float("inf"), device=dev)
    for a in range(0, P, 500_000):
        blk = p[a:a + 500_000]
        d2 = ((p[q][:, None, :] - blk[None, :, :]) ** 2).sum(-1)          # [512, 500k]
        best = torch.cat([best, d2.topk(4, dim=1, largest=False).values], 1).topk(4, dim=1, largest=False).values
    want = best[:, 1:].mean(1)                                             # drop the self match (0)
    assert float(best[:, 0].max()) == 0.0
    assert torch.allclose(got[q], want, rtol=1e-5, atol=1e-12), (got[q] - want).abs().max()


# ------------------------------------------------------------------------------ degenerate inputs
@pytest.mark.parametrize("case", ["nothing_visible", "single_gaussian", "odd_image"])
@pytest.mark.parametrize("fused", [True, False])
def test_renderer_degenerate_inputs(S, dev, case, fused):
    """empty binning (every Gaussian behind the camera), N = 1, and an image that is not a multiple of the
    tile size, through the whole renderer + backward on both call paths: finite results, same as the oracle"""
    from street_crafter_b200.renderer import render_kernel_gsplat
    deg = 3
    if case == "nothing_visible":
        n, W, H = 500, 160, 96
        s = make_gaussians(n, sh_degree=deg, seed=5)
        s["means"][:, 2] = -s["means"][:, 2].abs() - 1.0       # all behind the camera
    elif case == "single_gaussian":
        n, W, H = 1, 64, 48
        s = make_gaussians(n, sh_degree=deg, seed=6)
        s["means"][0] = torch.tensor([0.05, -0.02, 4.0])
        s["scales"][0] = torch.tensor([0.05, 0.08, 0.03])
    else:
        n, W, H = 3000, 333, 211
        s = make_gaussians(n, sh_degree=deg, seed=7, scale_mean=0.3)
    V, K = make_camera(W, H)
    c = {k: v.to(dev).requires_grad_(True) for k, v in s.items()}
    cc = (-(V[:3, :3].T @ V[:3, 3])).to(dev)
    out = render_kernel_gsplat(c["means"], c["shs"], c["scales"], c["quats"], c["opacities"][:, None],
                               V[None].to(dev), K[None].to(dev), cc, W, H, deg, fused=fused)
    assert out["rgb"].shape == (3, H, W) and out["acc"].shape == (1, H, W) and out["depth"].shape == (1, H, W)
    loss = out["rgb"].square().sum() + out["acc"].sum() + out["depth"].sum() * 1e-2
    loss.backward()
    for k in ("means", "quats", "scales", "opacities", "shs"):
        g = c[k].grad
        assert g is not None and bool(torch.isfinite(g).all()), k
    vis = out["visibility_filter"]
    if case == "nothing_visible":
        assert int(vis.sum()) == 0 and out["aux"]["isect_ids"].numel() == 0
        assert float(out["acc"].detach().abs().max()) == 0.0 and float(out["rgb"].detach().abs().max()) == 0.0
        assert all(float(c[k].grad.abs().max()) == 0.0 for k in ("means", "quats", "scales", "opacities", "shs"))
    else:
        assert int(vis.sum()) >= 1
    # oracle on the same inputs (forward image)
    o = {k: v.detach().cpu() for k, v in c.items()}
    want_rc, want_ra = G.render_reference_sequence(o["means"], o["quats"], o["scales"], o["opacities"], o["shs"], V, K,
                                                   W, H, deg)[:2]
    assert_image_close(out["render_colors"].cpu(), want_rc.detach(), "degenerate render")
    assert_image_close(out["render_alphas"].cpu(), want_ra.detach(), "degenerate alpha")


# ------------------------------------------------------------------------ prune / select compaction (N4)
def test_prune_state_vs_reference_golden(S, dev):
    """parameters, Adam moments and statistics after the reference's own prune_points == one compaction launch"""
    from street_crafter_b200.compaction import prune_state
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "prune_points.npz"))
    names = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation", "semantic")
    stats = ("xyz_gradient_accum", "denom", "max_radii2D")
    d = lambda k: torch.from_numpy(z[k]).to(dev)
    keep = ~d("mask")
    p, m, v, st = prune_state(keep, {k: d(f"in_{k}") for k in names}, {k: d(f"in_m_{k}") for k in names},
                              {k: d(f"in_v_{k}") for k in names}, {k: d(f"in_{k}") for k in stats})
    for k in names:
        for got, key in ((p[k], f"out_{k}"), (m[k], f"out_m_{k}"), (v[k], f"out_v_{k}")):
            assert got.shape == z[key].shape and torch.equal(got.cpu(), torch.from_numpy(z[key])), key
    for k in stats:
        assert torch.equal(st[k].cpu(), torch.from_numpy(z[f"out_{k}"])), k


@pytest.mark.parametrize("N,frac", [(0, 0.5), (1, 1.0), (1, 0.0), (4097, 0.0), (4097, 1.0), (300001, 0.37), (2_000_000, 0.9)])
def test_compact_rows_matches_boolean_indexing(S, dev, N, frac):
    """bit-exact against t[keep] for ragged row widths, empty / full masks and the cfg2 Gaussian count"""
    from street_crafter_b200.compaction import compact_rows
    g = torch.Generator().manual_seed(N + 1)
    keep = (torch.rand(N, generator=g) < frac).to(dev)
    ts = [torch.randn((N,) + s, generator=g).to(dev) for s in ((3,), (4,), (), (16, 3), (1,), (2,))]
    if N >= 1_000_000:
        ts = ts + [t.clone() for t in ts] + [t.clone() for t in ts]       # params + two Adam moments: 18 arrays
    out = compact_rows(keep, ts)
    assert len(out) == len(ts)
    for o, t in zip(out, ts):
        assert o.shape == t[keep].shape and torch.equal(o, t[keep])


# ------------------------------------------------------------------------ densification cycle (N4, clone / split)
DENSIFY_NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation", "semantic")


def _assert_cycle_equal(got, want, tag, exact_all=False):
    """got / want: (params, exp_avg, exp_avg_sq).  Copied rows are bit-exact; the split children's xyz (free
    summation order of the 3x3 product) and log-scales (exp / log of two math libraries) within 2e-6."""
    for gi, wi, what in zip(got, want, ("param", "exp_avg", "exp_avg_sq")):
        for k in wi:
            g, w = gi[k].cpu(), wi[k].cpu()
            assert g.shape == w.shape, (tag, what, k, g.shape, w.shape)
            if what == "param" and k in ("xyz", "scaling") and not exact_all:
                assert (g - w).abs().max() <= 2e-6 * max(1.0, float(w.abs().max())), (tag, k, float((g - w).abs().max()))
            else:
                assert torch.equal(g, w), (tag, what, k)


@pytest.mark.parametrize("tag", ["a", "b"])
def test_densify_cycle_vs_reference_golden(S, dev, tag):
    """parameters, Adam moments, statistics and scalar_dict after the reference's own densify_and_prune (clone +
    split + prune on a real Adam optimizer, tests/golden/densify_cycle.npz) == classify + scan + one move launch"""
    from street_crafter_b200.densify import densify_and_prune
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "densify_cycle.npz"))
    d = lambda k: torch.from_numpy(z[f"{tag}_{k}"]).to(dev)
    max_grad, min_opacity, extent, mss, pd, pbw = [float(x) for x in z[f"{tag}_cfg"]]
    p, m, v, st, sc = densify_and_prune({k: d(f"in_{k}") for k in DENSIFY_NAMES}, {k: d(f"in_m_{k}") for k in DENSIFY_NAMES},
                                        {k: d(f"in_v_{k}") for k in DENSIFY_NAMES}, d("in_xyz_gradient_accum"),
                                        d("in_denom"), max_grad, min_opacity, extent, mss or None, pd, pbw,
                                        noise=d("noise"))
    assert [sc[k] for k in ("points_total", "points_clone", "points_split", "points_pruned")] == z[f"{tag}_scalars"].tolist()
    want = tuple({k: torch.from_numpy(z[f"{tag}_out_{pre}{k}"]) for k in DENSIFY_NAMES} for pre in ("", "m_", "v_"))
    _assert_cycle_equal((p, m, v), want, tag)
    for k in ("xyz_gradient_accum", "denom", "max_radii2D"):
        assert st[k].shape == z[f"{tag}_out_{k}"].shape and not st[k].any(), k


@pytest.mark.parametrize("N,mss,col", [(1, 20, 0), (31, None, 0), (4097, 20, 1), (300_001, 20, 0), (2_000_000, 20, 0)])
def test_densify_cycle_vs_oracle(S, dev, N, mss, col):
    """ragged sizes up to the cfg2 Gaussian count against the pinned oracle run on the same device tensors (torch's
    own exp / log / sigmoid kernels), same noise; row counts and every copied value bit-exact"""
    from oracle.densify_cycle_torch import densify_and_prune as oracle_cycle
    from street_crafter_b200.densify import densify_and_prune
    g = torch.Generator().manual_seed(N)
    extent, max_grad, min_op = 20.0, 0.0002, 0.005
    shapes = {"xyz": (3,), "f_dc": (1, 3), "f_rest": (15, 3), "opacity": (1,), "scaling": (3,), "rotation": (4,)}
    params = {k: torch.randn((N,) + s, generator=g).to(dev) for k, s in shapes.items()}
    params["scaling"] = (math.log(0.01 * extent) + 0.8 * torch.randn(N, 3, generator=g)).to(dev)
    params["opacity"] = (2.5 * torch.randn(N, 1, generator=g) - 1.0).to(dev)
    m = {k: torch.randn((N,) + s, generator=g).to(dev) for k, s in shapes.items()}
    v = {k: torch.rand((N,) + s, generator=g).to(dev) for k, s in shapes.items() if k != "f_rest"}   # one group without exp_avg_sq
    denom = torch.randint(0, 4, (N, 1), generator=g).float().to(dev)
    accum = (torch.rand(N, 2, generator=g) * 2.5 * max_grad).to(dev) * denom.clamp(min=1)
    p1, m1, v1, st, sc = densify_and_prune(params, m, v, accum, denom, max_grad, min_op, extent, mss,
                                           generator=torch.Generator(device=dev).manual_seed(7), grad_column=col)
    noise = torch.randn((2 * sc["points_split"], 3), device=dev, generator=torch.Generator(device=dev).manual_seed(7))
    p0, m0, v0, sc0 = oracle_cycle(params, m, {**v, "f_rest": torch.zeros_like(m["f_rest"])}, accum, denom, noise,
                                   max_grad, min_op, extent, mss, grad_column=col)
    v0.pop("f_rest")
    assert sc == sc0, (sc, sc0)
    _assert_cycle_equal((p1, m1, v1), (p0, m0, v0), N)
    n = p1["xyz"].shape[0]
    assert st["xyz_gradient_accum"].shape == (n, 2) and st["denom"].shape == (n, 1) and st["max_radii2D"].shape == (n,)
    if N >= 4097:
        assert sc["points_clone"] > 0 and sc["points_split"] > 0 and sc["points_pruned"] > 0


def test_densify_cycle_empty_and_cpu_raises(S, dev):
    from street_crafter_b200.densify import densify_and_prune
    shapes = {"xyz": (3,), "opacity": (1,), "scaling": (3,), "rotation": (4,)}
    p = {k: torch.zeros((0,) + s, device=dev) for k, s in shapes.items()}
    out = densify_and_prune(p, {}, {}, torch.zeros(0, 2, device=dev), torch.zeros(0, 1, device=dev), 2e-4, 5e-3, 10.0, 20)
    assert out[0]["xyz"].shape == (0, 3) and out[4]["points_total"] == 0
    with pytest.raises(RuntimeError):
        densify_and_prune({k: t.cpu() for k, t in p.items()}, {}, {}, torch.zeros(0, 2), torch.zeros(0, 1), 2e-4, 5e-3, 10.0, 20)


# ------------------------------------------------------------------------ scene-graph assembly (N2, geometric half)
def _assemble_loss_and_grads(fn, segs, w_xyz, w_rot, keys=("xyz", "rotation", "pose_rot", "pose_trans")):
    xyz, rot = fn(segs)
    ((xyz * w_xyz).sum() + (rot * w_rot).sum()).backward()
    grads = []
    for s in segs:
        for k in keys:
            t = s.get(k)
            if t is not None:
                grads.append((s["name"], k, (t.grad if t.grad is not None else torch.zeros_like(t)).detach().clone()))
                t.grad = None
    return xyz.detach(), rot.detach(), grads


def test_assemble_vs_reference_golden(S, dev):
    """get_xyz / get_rotation values and autograd gradients (Gaussian parameters AND tracked poses) of the reference's
    own property bodies (tests/golden/assemble.npz: background + 3 actors incl. an empty one + sky, flips) == one
    forward + one backward kernel.  Values within 1e-5, gradients within 1e-4 relative to their scale."""
    from street_crafter_b200.scene_graph import assemble_xyz_rotation
    from test_oracle_cpu import _assemble_golden_segments
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "assemble.npz"))
    segs = _assemble_golden_segments(z, to=lambda t: t.to(dev))
    fq, ax = [float(v) for v in z["flip_quat"]], int(z["flip_axis"])
    xyz, rot, grads = _assemble_loss_and_grads(lambda s: assemble_xyz_rotation(s, fq, ax), segs,
                                               torch.from_numpy(z["w_xyz"]).to(dev), torch.from_numpy(z["w_rot"]).to(dev))
    assert (xyz.cpu() - torch.from_numpy(z["out_xyz"])).abs().max() < 1e-5
    assert (rot.cpu() - torch.from_numpy(z["out_rot"])).abs().max() < 1e-5
    key = {"xyz": "v_xyz", "rotation": "v_rot", "pose_rot": "v_pose_rot", "pose_trans": "v_pose_trans"}
    for name, k, g in grads:
        want = torch.from_numpy(z[f"{name}_{key[k]}"])
        assert g.shape == want.shape, (name, k)
        if want.numel():
            assert (g.cpu() - want).abs().max() <= 1e-4 * max(1.0, float(want.abs().max())), (name, k)


@pytest.mark.parametrize("n_bkgd,n_actors,n_per_actor,n_sky", [(1000, 2, 300, 0), (0, 40, 257, 50), (1_900_000, 20, 5000, 100_000)])
def test_assemble_vs_oracle(S, dev, n_bkgd, n_actors, n_per_actor, n_sky):
    """cfg2-like scene graph (1.9 M background + 20 actors x 5000 + sky) and a 41-segment graph (two kernel calls)
    against the pinned oracle run with autograd on the same device tensors"""
    from oracle.assemble_torch import assemble as oracle_assemble
    from street_crafter_b200.scene_graph import FLIP_QUAT_Y, assemble_xyz_rotation
    g = torch.Generator().manual_seed(n_actors)
    leaf = lambda *s, k=1.0: (torch.randn(*s, generator=g) * k).to(dev).requires_grad_(True)
    segs = []
    if n_bkgd:
        segs.append({"name": "background", "xyz": leaf(n_bkgd, 3, k=30.0), "rotation": leaf(n_bkgd, 4)})
    for a in range(n_actors):
        n = n_per_actor + a
        segs.append({"name": f"obj_{a:03d}", "xyz": leaf(n, 3, k=2.0), "rotation": leaf(n, 4), "pose_rot": leaf(4),
                     "pose_trans": leaf(3, k=20.0),
                     "flip": None if a % 3 == 2 else (torch.rand(n, generator=g) < 0.5).to(dev)})
    if n_sky:
        segs.append({"name": "sky", "xyz": leaf(n_sky, 3, k=100.0), "rotation": leaf(n_sky, 4),
                     "shell": ([3.0, -2.0, 1.0], 60.0)})     # GaussianModelSky.get_xyz: ~half the points inside 2 R
    N = sum(s["xyz"].shape[0] for s in segs)
    w_xyz, w_rot = torch.randn(N, 3, generator=g).to(dev), torch.randn(N, 4, generator=g).to(dev)
    fq = torch.tensor(FLIP_QUAT_Y, device=dev)
    x0, r0, g0 = _assemble_loss_and_grads(lambda s: oracle_assemble(s, fq, 1), segs, w_xyz, w_rot)
    x1, r1, g1 = _assemble_loss_and_grads(lambda s: assemble_xyz_rotation(s), segs, w_xyz, w_rot)
    assert x1.shape == (N, 3) and r1.shape == (N, 4)
    assert (x1 - x0).abs().max() <= 1e-5 * max(1.0, float(x0.abs().max())) and (r1 - r0).abs().max() < 1e-5
    for (name, k, a), (_, _, b) in zip(g1, g0):
        assert (a - b).abs().max() <= 2e-4 * max(1.0, float(b.abs().max())), (name, k, float((a - b).abs().max()))


def test_assemble_cpu_raises(S, dev):
    from street_crafter_b200.scene_graph import assemble_xyz_rotation
    with pytest.raises(RuntimeError):
        assemble_xyz_rotation([{"xyz": torch.zeros(4, 3), "rotation": torch.ones(4, 4)}])


def test_assemble_activations_vs_reference_golden(S, dev):
    """get_scaling / get_opacity (sky model clamped) of the reference's own property bodies, values and gradients"""
    from street_crafter_b200.scene_graph import assemble_scaling_opacity
    from test_oracle_cpu import _assemble_golden_segments
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "assemble.npz"))
    segs = _assemble_golden_segments(z, to=lambda t: t.to(dev))
    sca, opa, grads = _assemble_loss_and_grads(assemble_scaling_opacity, segs, torch.from_numpy(z["w_scaling"]).to(dev),
                                               torch.from_numpy(z["w_opacity"]).to(dev), keys=("scaling", "opacity"))
    want_s, want_o = torch.from_numpy(z["out_scaling"]), torch.from_numpy(z["out_opacity"])
    assert sca.shape == want_s.shape and opa.shape == want_o.shape
    assert ((sca.cpu() - want_s).abs() <= 2e-6 * want_s.abs()).all() and (opa.cpu() - want_o).abs().max() < 1e-6
    for name, k, g in grads:
        want = torch.from_numpy(z[f"{name}_v_{k}"])
        assert g.shape == want.shape, (name, k)
        if want.numel():
            assert (g.cpu() - want).abs().max() <= 1e-5 * max(1.0, float(want.abs().max())), (name, k)


def test_assemble_activations_vs_oracle_cfg2(S, dev):
    """1.9 M background + 40 actors (two kernel calls) + clamped sky against torch's own exp / sigmoid / clamp"""
    from oracle.assemble_torch import assemble_activations
    from street_crafter_b200.scene_graph import assemble_scaling_opacity
    g = torch.Generator().manual_seed(5)
    leaf = lambda *s, k=1.0, b=0.0: (torch.randn(*s, generator=g) * k + b).to(dev).requires_grad_(True)
    sizes = [1_900_000] + [2500 + a for a in range(40)] + [100_000]
    segs = [{"name": str(i), "scaling": leaf(n, 3, b=-2.0), "opacity": leaf(n, 1, k=2.5)} for i, n in enumerate(sizes)]
    segs[-1]["scaling_clamp_max"] = 0.2
    N = sum(sizes)
    w_s, w_o = torch.randn(N, 3, generator=g).to(dev), torch.randn(N, 1, generator=g).to(dev)
    s0, o0, g0 = _assemble_loss_and_grads(assemble_activations, segs, w_s, w_o, keys=("scaling", "opacity"))
    s1, o1, g1 = _assemble_loss_and_grads(assemble_scaling_opacity, segs, w_s, w_o, keys=("scaling", "opacity"))
    assert ((s1 - s0).abs() <= 2e-6 * s0.abs()).all() and (o1 - o0).abs().max() < 1e-6
    for (name, k, a), (_, _, b) in zip(g1, g0):
        assert (a - b).abs().max() <= 1e-5 * max(1.0, float(b.abs().max())), (name, k)


def test_assemble_features_vs_reference_golden(S, dev):
    """get_features (Fourier DC of the actors, cat with the rest coefficients) of the reference's own bodies"""
    from street_crafter_b200.scene_graph import assemble_features
    from test_oracle_cpu import _assemble_golden_segments
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "assemble.npz"))
    segs = _assemble_golden_segments(z, to=lambda t: t.to(dev))
    fea = assemble_features(segs)
    want = torch.from_numpy(z["out_features"])
    assert fea.shape == want.shape and (fea.detach().cpu() - want).abs().max() < 1e-6
    (fea * torch.from_numpy(z["w_features"]).to(dev)).sum().backward()
    for s in segs:
        for key, t in (("v_f_dc", s["f_dc"]), ("v_f_rest", s["f_rest"])):
            w = torch.from_numpy(z[f"{s['name']}_{key}"])
            got = t.grad if t.grad is not None else torch.zeros_like(t)
            assert got.shape == w.shape, (s["name"], key)
            if w.numel():
                assert (got.cpu() - w).abs().max() < 1e-6, (s["name"], key)


@pytest.mark.parametrize("K", [16, 4, 1])
def test_assemble_features_vs_oracle_cfg2(S, dev, K):
    """1.9 M background + 40 actors (fourier_dim 1..5, two kernel calls) + sky, SH degree 3 / 1 / 0"""
    from oracle.assemble_torch import assemble_features as oracle_features, idft
    from street_crafter_b200.scene_graph import assemble_features
    g = torch.Generator().manual_seed(K)
    leaf = lambda *s: torch.randn(*s, generator=g).to(dev).requires_grad_(True)
    sizes = [(1_900_000 if K == 16 else 50_000, 1)] + [(2500 + a, 1 + a % 5) for a in range(40)] + [(10_000, 1)]
    segs = []
    for i, (n, C) in enumerate(sizes):
        s = {"name": str(i), "f_dc": leaf(n, C, 3), "f_rest": leaf(n, K - 1, 3)}
        if 0 < i <= 40:
            s["idft_base"] = idft(0.1 + 0.02 * i, C)[0].tolist()
        segs.append(s)
    N = sum(n for n, _ in sizes)
    w = torch.randn(N, K, 3, generator=g).to(dev)
    res = []
    for fn in (oracle_features, assemble_features):
        f = fn(segs)
        (f * w).sum().backward()
        res.append((f.detach(), [(s["f_dc"].grad.clone(), s["f_rest"].grad.clone() if s["f_rest"].grad is not None else None) for s in segs]))
        for s in segs:
            s["f_dc"].grad = s["f_rest"].grad = None
    assert res[1][0].shape == (N, K, 3) and (res[1][0] - res[0][0]).abs().max() < 2e-6
    for (a_dc, a_r), (b_dc, b_r) in zip(res[1][1], res[0][1]):
        assert (a_dc - b_dc).abs().max() < 2e-6
        if b_r is not None and b_r.numel():
            assert torch.equal(a_r, b_r)


def test_prepare_view_two_phase_equals_single_pass(S, dev):
    """rendering.prepare_view (projection + binning first) + fused_view(prepared=...) == one fused_view call,
    forward bit-equal, gradients to the rounding of the atomic accumulation; a state prepared for another image
    size is refused"""
    from street_crafter_b200.rendering import prepare_view
    from street_crafter_b200.renderer import render_kernel_gsplat
    n, W, H, deg = 40000, 640, 400, 3
    s = make_gaussians(n, sh_degree=deg, seed=8)
    V, K = make_camera(W, H, yaw_deg=-12.0)
    Vd, Kd = V[None].to(dev), K[None].to(dev)
    cc = (-(V[:3, :3].T @ V[:3, 3])).to(dev)
    g = torch.Generator().manual_seed(1)
    v_rc, v_ra = torch.randn(1, H, W, 4, generator=g).to(dev), torch.randn(1, H, W, 1, generator=g).to(dev)
    res = []
    for two_phase in (False, True):
        c = {k: v.to(dev).requires_grad_(True) for k, v in s.items()}
        st = prepare_view(c["means"], c["quats"], c["scales"], c["opacities"], Vd, Kd, W, H, near_plane=0.001,
                          far_plane=1000.0, need_isect_ids=True) if two_phase else None
        out = render_kernel_gsplat(c["means"], c["shs"], c["scales"], c["quats"], c["opacities"][:, None], Vd, Kd, cc, W, H,
                                   deg, prepared=st)
        torch.autograd.backward([out["render_colors"], out["render_alphas"]], [v_rc, v_ra])
        res.append((c, out))
    (c0, o0), (c1, o1) = res
    for k in ("render_colors", "render_alphas", "depth", "radii", "visibility_filter"):
        assert torch.equal(o0[k], o1[k]), k
    for k in ("isect_ids", "flatten_ids", "isect_offsets", "tiles_per_gauss"):
        assert torch.equal(o0["aux"][k], o1["aux"][k]), k
    for k in c0:
        # same kernels, the atomic accumulation in another order: the noise is ~1e-7 of the LARGEST partial sum, which
        # after cancellation can exceed 1e-6 of the largest entry (seen once on quats) -> bar at 1e-4 of the scale
        assert float((c0[k].grad - c1[k].grad).abs().max()) <= 1e-4 * float(c0[k].grad.abs().max()), k
    assert o1["viewspace_points"].absgrad is not None
    with pytest.raises(AssertionError):
        st = prepare_view(c0["means"], c0["quats"], c0["scales"], c0["opacities"], Vd, Kd, W, H)
        render_kernel_gsplat(c0["means"], c0["shs"], c0["scales"], c0["quats"], c0["opacities"][:, None], Vd, Kd, cc, W // 2, H,
                             deg, prepared=st)
