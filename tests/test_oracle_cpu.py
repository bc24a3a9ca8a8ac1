"""CPU tests (-m "not gpu"): the oracle against the reference-derived golden vectors and its
own invariants, and the C-ABI library's export table.  No GPU compute here."""
import ctypes
import math
import os
import re

import numpy as np
import pytest
import torch

from oracle import gsplat_torch as G
from oracle import knn as KO
from tools.scene import make_camera, make_gaussians

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ---------------------------------------------------------------- SH pinned on the reference
@pytest.mark.parametrize("deg", [0, 1, 2, 3, 4])
def test_sh_matches_reference_golden(deg):
    """tests/golden/sh_eval.npz was produced by /root/reference/street_gaussian/utils/sh_utils.py
    eval_sh (tools/make_golden.py); the oracle must reproduce it."""
    z = np.load(os.path.join(ROOT, "tests", "golden", "sh_eval.npz"))
    dirs = torch.from_numpy(z["dirs"])
    coeffs = torch.from_numpy(z["coeffs"])
    K = (deg + 1) ** 2
    got = G.spherical_harmonics(deg, dirs, coeffs[:, :K])
    want = torch.from_numpy(z[f"colors_deg{deg}"])
    assert torch.allclose(got, want, atol=2e-5, rtol=1e-5), (got - want).abs().max()


def test_sh_mask_zeroes():
    g = torch.Generator().manual_seed(1)
    dirs = torch.randn(10, 3, generator=g)
    coeffs = torch.randn(10, 16, 3, generator=g)
    m = torch.arange(10) % 2 == 0
    out = G.spherical_harmonics(3, dirs, coeffs, m)
    assert (out[~m] == 0).all() and (out[m] != 0).any()


# ---------------------------------------------------------------- projection invariants
def _small_scene(n=4000, W=320, H=192, seed=3):
    s = make_gaussians(n, seed=seed)
    V, K = make_camera(W, H)
    return s, V, K, W, H


def test_projection_matches_matrix_form_fp64():
    """the scalar-op restatement equals the textbook matrix form  J W Σ Wᵀ Jᵀ  in fp64."""
    s, V, K, W, H = _small_scene()
    d = {k: v.double() for k, v in s.items()}
    Vd, Kd = V.double(), K.double()
    radii, m2, dep, con, comp = G.fully_fused_projection(d["means"], None, d["quats"], d["scales"], Vd[None],
                                                         Kd[None], W, H, near_plane=0.001, far_plane=1000.0,
                                                         calc_compensations=True)
    q = torch.nn.functional.normalize(d["quats"], dim=-1)
    w, x, y, z = q.unbind(-1)
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                     2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                     2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], -1).reshape(-1, 3, 3)
    M = R * d["scales"][:, None, :]
    S = M @ M.transpose(1, 2)
    Rw, t = Vd[:3, :3], Vd[:3, 3]
    pc = d["means"] @ Rw.T + t
    Sc = Rw @ S @ Rw.T
    vis = radii[0] > 0
    fx, fy, cx, cy = Kd[0, 0], Kd[1, 1], Kd[0, 2], Kd[1, 2]
    zc = pc[:, 2]
    assert torch.allclose(dep[0][vis], zc[vis])
    assert torch.allclose(m2[0][vis, 0], fx * pc[vis, 0] / zc[vis] + cx)
    # unclamped Gaussians only (well inside the frustum)
    inside = vis & (m2[0, :, 0] > 0) & (m2[0, :, 0] < W) & (m2[0, :, 1] > 0) & (m2[0, :, 1] < H)
    J = torch.zeros(len(zc), 2, 3, dtype=torch.float64)
    J[:, 0, 0] = fx / zc
    J[:, 0, 2] = -fx * pc[:, 0] / zc ** 2
    J[:, 1, 1] = fy / zc
    J[:, 1, 2] = -fy * pc[:, 1] / zc ** 2
    c2 = J @ Sc @ J.transpose(1, 2)
    c2[:, 0, 0] += 0.3
    c2[:, 1, 1] += 0.3
    inv = torch.linalg.inv(c2[inside])
    assert torch.allclose(con[0][inside, 0], inv[:, 0, 0], rtol=1e-9)
    assert torch.allclose(con[0][inside, 1], inv[:, 0, 1], rtol=1e-9, atol=1e-12)
    assert torch.allclose(con[0][inside, 2], inv[:, 1, 1], rtol=1e-9)
    assert inside.sum() > 100
    assert (comp[0][vis] <= 1.0 + 1e-12).all() and (comp[0][vis] > 0).all()


def test_projection_culls_and_zeroes():
    s, V, K, W, H = _small_scene()
    radii, m2, dep, con, comp = G.fully_fused_projection(s["means"], None, s["quats"], s["scales"], V[None],
                                                         K[None], W, H, near_plane=0.001, far_plane=1000.0,
                                                         calc_compensations=True)
    culled = radii[0] == 0
    assert culled.any() and (~culled).any()
    for t in (m2[0], con[0]):
        assert (t[culled] == 0).all()
    assert (dep[0][culled] == 0).all() and (comp[0][culled] == 0).all()
    assert radii.dtype == torch.int32
    behind = (s["means"] @ V[:3, :3].T + V[:3, 3])[:, 2] < 0.001
    assert (radii[0][behind] == 0).all()


# ---------------------------------------------------------------- binning invariants
def test_isect_keys_sorted_and_offsets_consistent():
    s, V, K, W, H = _small_scene()
    rc, ra, aux = G.render_reference_sequence(s["means"], s["quats"], s["scales"], s["opacities"], s["shs"], V,
                                              K, W, H, 3, return_aux=True)
    ids, flat, off, tpg = aux["isect_ids"], aux["flatten_ids"], aux["isect_offsets"], aux["tiles_per_gauss"]
    assert ids.dtype == torch.int64 and flat.dtype == torch.int32 and off.dtype == torch.int32
    assert int(tpg.sum()) == ids.numel() > 0
    assert (ids[1:] >= ids[:-1]).all()
    tw, th = math.ceil(W / 16), math.ceil(H / 16)
    tile = (ids >> 32).int()
    offs = off.reshape(-1).tolist() + [ids.numel()]
    for t in range(tw * th):
        seg = tile[offs[t]:offs[t + 1]]
        assert (seg == t).all()
    # depth bits round-trip
    dbits = (ids & 0xFFFFFFFF).to(torch.int32).view(torch.float32)
    assert torch.equal(dbits, aux["depths"][0][flat.long()])
    # brute-force recount of tile coverage for a few Gaussians
    vis = torch.nonzero(aux["radii"][0] > 0).flatten()[:50]
    for g in vis.tolist():
        x, y = aux["means2d"][0, g].tolist()
        r = int(aux["radii"][0, g])
        x0 = min(max(0, math.floor((x - r) / 16)), tw); x1 = min(max(0, math.ceil((x + r) / 16)), tw)
        y0 = min(max(0, math.floor((y - r) / 16)), th); y1 = min(max(0, math.ceil((y + r) / 16)), th)
        assert int(tpg[0, g]) == (x1 - x0) * (y1 - y0)


def test_isect_empty_and_multicamera():
    tpg, ids, flat = G.isect_tiles(torch.zeros(2, 5, 2), torch.zeros(2, 5, dtype=torch.int32),
                                   torch.zeros(2, 5), 16, 4, 4, n_cameras=2)
    assert ids.numel() == 0 and flat.numel() == 0 and int(tpg.sum()) == 0
    off = G.isect_offset_encode(ids, 2, 4, 4)
    assert off.shape == (2, 4, 4) and (off == 0).all()
    m2 = torch.tensor([[[8.0, 8.0]], [[40.0, 40.0]]])
    rad = torch.tensor([[3], [3]], dtype=torch.int32)
    dep = torch.tensor([[2.0], [1.0]])
    tpg, ids, flat = G.isect_tiles(m2, rad, dep, 16, 4, 4, n_cameras=2)
    assert tpg.tolist() == [[1], [1]]
    bits = 5  # floor(log2(16)) + 1
    assert (ids >> (32 + bits)).tolist() == [0, 1]
    assert ((ids >> 32) & 31).tolist() == [0, 2 * 4 + 2]
    assert flat.tolist() == [0, 1]


# ---------------------------------------------------------------- compositing
def test_rasterize_matches_naive_loop():
    """vectorised oracle == literal per-pixel front-to-back loop on a tiny scene."""
    s = make_gaussians(300, seed=5, scale_mean=0.6)
    W, H = 48, 32
    V, K = make_camera(W, H)
    rc, ra, aux = G.render_reference_sequence(s["means"], s["quats"], s["scales"], s["opacities"], s["shs"], V,
                                              K, W, H, 3, return_aux=True)
    off = aux["isect_offsets"].reshape(-1).tolist() + [aux["flatten_ids"].numel()]
    tw = math.ceil(W / 16)
    m2, con, col, op = aux["means2d"][0], aux["conics"][0], aux["colors"][0], aux["opacities"][0]
    worst = 0.0
    for (i, j) in [(0, 0), (5, 7), (17, 33), (31, 47), (16, 16), (20, 40)]:
        t = (i // 16) * tw + j // 16
        T, pix = 1.0, torch.zeros(4)
        for g in aux["flatten_ids"][off[t]:off[t + 1]].tolist():
            dx, dy = float(m2[g, 0]) - (j + 0.5), float(m2[g, 1]) - (i + 0.5)
            sig = 0.5 * (float(con[g, 0]) * dx * dx + float(con[g, 2]) * dy * dy) + float(con[g, 1]) * dx * dy
            a = min(0.999, float(op[g]) * math.exp(-sig))
            if sig < 0 or a < 1 / 255:
                continue
            nT = T * (1 - a)
            if nT <= 1e-4:
                break
            pix += col[g] * a * T
            T = nT
        worst = max(worst, float((pix - rc[0, i, j]).abs().max()), abs((1 - T) - float(ra[0, i, j, 0])))
    assert worst < 1e-4, worst
    assert ra.max() > 0.2


def test_oracle_gradcheck_fp64():
    """finite-difference check of the oracle's autograd backward on a tiny fp64 scene."""
    s = make_gaussians(24, seed=7, scale_mean=1.5, sh_degree=1)
    W, H = 32, 32
    V, K = make_camera(W, H)
    d = {k: v.double() for k, v in s.items()}
    d["means"][:, 2] = d["means"][:, 2].abs() + 20.0
    d["means"][:, 0] *= 0.05
    d["means"][:, 1] *= 0.2
    Vd, Kd = V.double(), K.double()
    g = torch.Generator().manual_seed(0)
    wc = torch.randn(1, H, W, 4, generator=g, dtype=torch.float64)
    wa = torch.randn(1, H, W, 1, generator=g, dtype=torch.float64)

    def loss(means, quats, scales, opac, shs):
        rc, ra = G.render_reference_sequence(means, quats, scales, opac, shs, Vd, Kd, W, H, 1)
        return (rc * wc).sum() + (ra * wa).sum()

    leaves = [d[k].clone().requires_grad_(True) for k in ("means", "quats", "scales", "opacities", "shs")]
    L = loss(*leaves)
    grads = torch.autograd.grad(L, leaves)
    eps = 1e-6
    rng = np.random.RandomState(0)
    for li, (leaf, gr) in enumerate(zip(leaves, grads)):
        flat = leaf.detach().reshape(-1)
        for idx in rng.choice(flat.numel(), size=4, replace=False):
            args_p = [l.detach().clone() for l in leaves]
            args_m = [l.detach().clone() for l in leaves]
            args_p[li].reshape(-1)[idx] += eps
            args_m[li].reshape(-1)[idx] -= eps
            fd = (loss(*args_p) - loss(*args_m)) / (2 * eps)
            an = gr.reshape(-1)[idx]
            assert abs(fd - an) <= 1e-4 * max(1.0, abs(fd)), (li, idx, float(fd), float(an))


# ---------------------------------------------------------------- KNN oracle
def test_knn_bruteforce_vs_kdtree():
    g = torch.Generator().manual_seed(0)
    pts = torch.randn(3000, 3, generator=g) * torch.tensor([10.0, 2.0, 30.0])
    pts[100] = pts[7]  # exact duplicate -> distance 0 counts
    a = KO.dist_knn3_bruteforce(pts)
    b = KO.dist_knn3_kdtree(pts)
    assert torch.allclose(a, b, rtol=1e-4, atol=1e-7)
    d = ((pts[7] - pts) ** 2).sum(-1)
    d[7] = float("inf")
    assert abs(float(a[7]) - float(d.topk(3, largest=False).values.mean())) < 1e-5
    assert float(d.min()) == 0.0


def test_knn_tiny():
    pts = torch.tensor([[0.0, 0, 0], [1, 0, 0], [0, 2, 0], [0, 0, 3]])
    out = KO.dist_knn3_bruteforce(pts)
    assert torch.allclose(out[0], torch.tensor((1.0 + 4.0 + 9.0) / 3))
    # P < 4: FLT_MAX placeholders survive into the mean (simple_knn.cu:149,183)
    assert (KO.dist_knn3_bruteforce(pts[:3]) > 1e37).all()


# ---------------------------------------------------------------- C ABI export table
def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "scb200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)  # comments mention names loosely
    names = sorted(set(re.findall(r"\b(sc_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 18, names
    from street_crafter_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"libscb200.so does not export {n}"
    assert set(_lib.SIGNATURES) == set(names), set(_lib.SIGNATURES) ^ set(names)
    assert _lib.lib.sc_version() >= 100


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "street_crafter_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f


def test_cpu_tensors_raise():
    import street_crafter_b200 as S
    with pytest.raises(RuntimeError):
        S.distCUDA2(torch.zeros(8, 3))
    with pytest.raises(RuntimeError):
        S.fully_fused_projection(torch.zeros(4, 3), None, torch.ones(4, 4), torch.ones(4, 3),
                                 torch.eye(4)[None], torch.eye(3)[None], 32, 32)


# ------------------------------------------------------------------------------ photometric loss (N3)
def test_loss_oracle_matches_reference_golden():
    """oracle/loss_torch.py against outputs of the reference's own l1_loss / ssim
    (tests/golden/loss_l1_ssim.npz, generated by tools/make_golden_loss.py importing
    /root/reference/street_gaussian/utils/loss_utils.py)"""
    import numpy as np
    from oracle import loss_torch as LO
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "loss_l1_ssim.npz"))
    for name in ("a", "b", "c"):
        img = torch.from_numpy(z[f"{name}_img"]).requires_grad_(True)
        gt = torch.from_numpy(z[f"{name}_gt"])
        mask = torch.from_numpy(z[f"{name}_mask"]) if z[f"{name}_mask"].size else None
        loss, l1, s = LO.photometric_loss(img, gt, mask, lambda_dssim=0.2, lambda_l1=1.0)
        loss.backward()
        assert abs(float(l1) - float(z[f"{name}_l1"])) < 1e-12
        assert abs(float(s) - float(z[f"{name}_ssim"])) < 1e-10
        assert abs(float(loss) - float(z[f"{name}_loss"])) < 1e-10
        assert np.abs(img.grad.numpy() - z[f"{name}_grad"]).max() < 1e-10


def test_densify_oracle_matches_reference_golden():
    """oracle/densify_torch.py against vectors produced by EXECUTING the reference's own
    set_max_radii2D / add_densification_stats (tools/make_golden_densify.py)"""
    import numpy as np
    from oracle.densify_torch import densification_stats
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "densify_stats.npz"))
    t = lambda k: torch.from_numpy(z[k])
    acc, den, mr = densification_stats(t("grad"), t("absgrad"), t("vis"), t("radii"), int(z["W"]), int(z["H"]),
                                       t("accum_in").clone(), t("denom_in").clone(), t("max_radii2D_in").clone())
    assert torch.allclose(acc, t("accum_out"), rtol=1e-12, atol=0)
    assert torch.equal(den, t("denom_out")) and torch.equal(mr, t("max_radii2D_out"))
    # Gaussians outside the visibility filter are untouched
    inv = ~t("vis")
    assert torch.equal(acc[inv], t("accum_in")[inv]) and torch.equal(den[inv], t("denom_in")[inv])


PRUNE_NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation", "semantic")
PRUNE_STATS = ("xyz_gradient_accum", "denom", "max_radii2D")


def test_prune_oracle_matches_reference_golden():
    """oracle/prune_torch.py against vectors produced by EXECUTING the reference's prune_optimizer /
    prune_points on a real Adam optimizer (tools/make_golden_prune.py)"""
    import numpy as np
    from oracle.prune_torch import prune_points
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "prune_points.npz"))
    t = lambda k: torch.from_numpy(z[k])
    p, m, v, st = prune_points(t("mask"), {k: t(f"in_{k}") for k in PRUNE_NAMES}, {k: t(f"in_m_{k}") for k in PRUNE_NAMES},
                               {k: t(f"in_v_{k}") for k in PRUNE_NAMES}, {k: t(f"in_{k}") for k in PRUNE_STATS})
    for k in PRUNE_NAMES:
        assert torch.equal(p[k], t(f"out_{k}")) and torch.equal(m[k], t(f"out_m_{k}")) and torch.equal(v[k], t(f"out_v_{k}")), k
    for k in PRUNE_STATS:
        assert torch.equal(st[k], t(f"out_{k}")), k


@pytest.mark.parametrize("tag", ["a", "b"])
def test_densify_cycle_oracle_matches_reference_golden(tag):
    """oracle/densify_cycle_torch.py against vectors produced by EXECUTING the reference's densify_and_prune
    (clone + split + prune, with cat_optimizer / prune_optimizer on a real Adam optimizer;
    tools/make_golden_densify_cycle.py).  Case b has max_screen_size=None (no big-point pruning)."""
    import numpy as np
    from oracle.densify_cycle_torch import densify_and_prune, n_split
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "densify_cycle.npz"))
    t = lambda k: torch.from_numpy(z[f"{tag}_{k}"])
    max_grad, min_opacity, extent, mss, pd, pbw = [float(x) for x in z[f"{tag}_cfg"]]
    params = {k: t(f"in_{k}") for k in PRUNE_NAMES}
    assert 2 * n_split(params, t("in_xyz_gradient_accum"), t("in_denom"), max_grad, extent, pd) == t("noise").shape[0]
    p, m, v, sc = densify_and_prune(params, {k: t(f"in_m_{k}") for k in PRUNE_NAMES},
                                    {k: t(f"in_v_{k}") for k in PRUNE_NAMES}, t("in_xyz_gradient_accum"),
                                    t("in_denom"), t("noise"), max_grad, min_opacity, extent, mss or None, pd, pbw)
    assert [sc[k] for k in ("points_total", "points_clone", "points_split", "points_pruned")] == z[f"{tag}_scalars"].tolist()
    for k in PRUNE_NAMES:
        assert p[k].shape == t(f"out_{k}").shape, k
        if k == "xyz":      # the only arithmetic with a free summation order (bmm)
            assert (p[k] - t(f"out_{k}")).abs().max() < 1e-6
        else:
            assert torch.equal(p[k], t(f"out_{k}")), k
        assert torch.equal(m[k], t(f"out_m_{k}")) and torch.equal(v[k], t(f"out_v_{k}")), k
    n = p["xyz"].shape[0]
    for k in PRUNE_STATS:
        assert z[f"{tag}_out_{k}"].shape[0] == n and not z[f"{tag}_out_{k}"].any(), k


def _assemble_golden_segments(z, to=lambda t: t, grad=True):
    segs = []
    for name in z["names"]:
        name = str(name)
        mk = lambda k: to(torch.from_numpy(z[f"{name}_{k}"])).requires_grad_(grad)
        s = {"name": name, "xyz": mk("xyz"), "rotation": mk("rot"), "scaling": mk("scaling"), "opacity": mk("opacity"),
             "f_dc": mk("f_dc"), "f_rest": mk("f_rest")}
        if name.startswith("obj_"):
            s["idft_base"] = [float(v) for v in z["actor_idft_base"]]
        if name == "sky":
            s["scaling_clamp_max"] = float(z["sky_clamp_max"])
            s["shell"] = ([float(v) for v in z["sky_center"]], float(z["sky_radius"]))
        if name.startswith("obj_"):
            s.update(pose_rot=mk("pose_rot"), pose_trans=mk("pose_trans"), flip=to(torch.from_numpy(z[f"{name}_flip"])))
        segs.append(s)
    return segs


def test_assemble_oracle_matches_reference_golden():
    """oracle/assemble_torch.py against values and autograd gradients produced by EXECUTING the reference's
    StreetGaussianModel.get_xyz / get_rotation (tools/make_golden_assemble.py)"""
    import numpy as np
    from oracle.assemble_torch import assemble
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "assemble.npz"))
    segs = _assemble_golden_segments(z)
    xyz, rot = assemble(segs, torch.from_numpy(z["flip_quat"]), int(z["flip_axis"]))
    assert torch.allclose(xyz, torch.from_numpy(z["out_xyz"]), atol=1e-6, rtol=1e-6)
    assert torch.allclose(rot, torch.from_numpy(z["out_rot"]), atol=1e-6, rtol=1e-6)
    ((xyz * torch.from_numpy(z["w_xyz"])).sum() + (rot * torch.from_numpy(z["w_rot"])).sum()).backward()
    for s in segs:
        n = s["name"]
        for key, t in (("v_xyz", s["xyz"]), ("v_rot", s["rotation"]), ("v_pose_rot", s.get("pose_rot")),
                       ("v_pose_trans", s.get("pose_trans"))):
            if t is None:
                continue
            got = t.grad if t.grad is not None else torch.zeros_like(t)
            want = torch.from_numpy(z[f"{n}_{key}"])
            assert torch.allclose(got, want, atol=2e-5, rtol=1e-5), (n, key, float((got - want).abs().max()))


def test_assemble_activations_oracle_matches_reference_golden():
    """get_scaling / get_opacity of the reference's StreetGaussianModel (sky model clamped) incl. autograd gradients"""
    import numpy as np
    from oracle.assemble_torch import assemble_activations
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "assemble.npz"))
    segs = _assemble_golden_segments(z)
    sca, opa = assemble_activations(segs)
    assert torch.allclose(sca, torch.from_numpy(z["out_scaling"]), atol=0, rtol=1e-6)
    assert torch.allclose(opa, torch.from_numpy(z["out_opacity"]), atol=1e-7, rtol=1e-6)
    assert (torch.from_numpy(z["out_scaling"][-int(z["sizes"][-1]):]) == float(z["sky_clamp_max"])).any()   # the clamp is exercised
    ((sca * torch.from_numpy(z["w_scaling"])).sum() + (opa * torch.from_numpy(z["w_opacity"])).sum()).backward()
    for s in segs:
        for key, t in (("v_scaling", s["scaling"]), ("v_opacity", s["opacity"])):
            got = t.grad if t.grad is not None else torch.zeros_like(t)
            assert torch.allclose(got, torch.from_numpy(z[f"{s['name']}_{key}"]), atol=1e-7, rtol=1e-5), (s["name"], key)


def test_assemble_features_oracle_matches_reference_golden():
    """get_features of the reference's StreetGaussianModel (actors through get_features_fourier + IDFT) incl. gradients"""
    import numpy as np
    from oracle.assemble_torch import assemble_features, idft
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "assemble.npz"))
    assert torch.allclose(idft((31 - 10) / (58 - 10), 5)[0], torch.from_numpy(z["actor_idft_base"]), atol=1e-7)
    segs = _assemble_golden_segments(z)
    fea = assemble_features(segs)
    assert torch.allclose(fea, torch.from_numpy(z["out_features"]), atol=1e-6, rtol=1e-6)
    (fea * torch.from_numpy(z["w_features"])).sum().backward()
    for s in segs:
        for key, t in (("v_f_dc", s["f_dc"]), ("v_f_rest", s["f_rest"])):
            got = t.grad if t.grad is not None else torch.zeros_like(t)
            assert torch.allclose(got, torch.from_numpy(z[f"{s['name']}_{key}"]), atol=1e-6, rtol=1e-6), (s["name"], key)


def test_scene_graph_and_densify_wrappers_have_no_cpu_path():
    """the product wrappers of the N2 / N4 rows refuse CPU tensors before touching the library (no fallback)"""
    from street_crafter_b200.compaction import compact_rows
    from street_crafter_b200.densify import densify_and_prune
    from street_crafter_b200.scene_graph import assemble_features, assemble_scaling_opacity, assemble_xyz_rotation
    z3, z4, z1 = torch.zeros(2, 3), torch.ones(2, 4), torch.zeros(2, 1)
    with pytest.raises(RuntimeError):
        assemble_xyz_rotation([{"xyz": z3, "rotation": z4}])
    with pytest.raises(RuntimeError):
        assemble_scaling_opacity([{"scaling": z3, "opacity": z1}])
    with pytest.raises(RuntimeError):
        assemble_features([{"f_dc": torch.zeros(2, 1, 3), "f_rest": torch.zeros(2, 15, 3)}])
    with pytest.raises(RuntimeError):
        densify_and_prune({"xyz": z3, "scaling": z3, "rotation": z4, "opacity": z1}, {}, {}, torch.zeros(2, 2), z1,
                          2e-4, 5e-3, 10.0, 20)
    with pytest.raises(RuntimeError):
        compact_rows(torch.ones(2, dtype=torch.bool), [z3])
