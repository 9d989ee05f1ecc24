"""CPU: pins the oracle (oracle/gs_oracle.*) on the reference's own golden vectors
(DGR/tests/data/args.pth -> DGR/tests/render_data/*.csv, repacked in tests/golden/dgr_fixture.npz) and checks its
self-consistency where the reference holds no golden vector (FD-vs-JVP, adjoint identity -- the reference's own
test pattern, DGR/tests/preprocess_jvp_test.cu, tests/test_matvec.py:15-20,81-83)."""
import numpy as np
import pytest

import gs_oracle as o
import scene_synth as ss
from helpers import relerr, rel_l2


def _fixture_inputs(d):
    return dict(means3D=d["means3D"], opacities=d["opacities"], viewmatrix=d["viewmatrix"],
                projmatrix=d["projmatrix"], campos=d["campos_raw"][:3], W=int(d["W"]), H=int(d["H"]),
                tan_fovx=float(d["tan_fovx"]), tan_fovy=float(d["tan_fovy"]), scales=d["scales"],
                rotations=d["rotations"], shs=d["shs"], sh_degree=int(d["sh_degree"]),
                scale_modifier=float(d["scale_modifier"]))


def _camera_tangents():
    # DGR/tests/forward_tests.cu:398-409
    i = np.arange(16)
    return {"viewmatrix": ((i % 4) * 0.03 + (i + 2) * 0.01).astype(np.float32),
            "projmatrix": ((i % 4) * 0.02 + (i + 2) * 0.02).astype(np.float32),
            "campos": (np.arange(3) * 0.014 + 0.03).astype(np.float32)}


def _close_csv(a, ref, tol=2e-5):
    """the CSVs carry 6 significant digits of fp32 numbers"""
    a = np.asarray(a, np.float64)
    ref = np.asarray(ref, np.float64)
    return np.abs(a - ref) <= tol * np.maximum(1.0, np.abs(ref)) + 1e-6 * np.abs(ref)


def test_preprocess_and_binning_match_reference_fixture(golden):
    d = golden
    kw = _fixture_inputs(d)
    args = [kw.pop(k) for k in ("means3D", "opacities", "viewmatrix", "projmatrix", "campos", "W", "H", "tan_fovx",
                                "tan_fovy")]
    pre = o.preprocess(*args, **kw)
    tiles_ref = np.diff(np.concatenate([[0], d["point_offsets"].astype(np.int64)]))
    assert np.array_equal(pre["tiles_touched"].astype(np.int64), tiles_ref)          # exact
    vis = tiles_ref > 0
    assert np.array_equal(pre["radii"] > 0, vis)
    # floats to CSV precision
    assert _close_csv(pre["xy"][vis], d["means2D"][vis], 1e-5).mean() > 0.995
    assert np.abs(pre["xy"][vis] - d["means2D"][vis]).max() < 2e-3
    assert _close_csv(pre["depths"][vis], d["depth"][vis]).all()
    assert _close_csv(pre["conic_opacity"][vis], d["conic_opacity"][vis]).mean() > 0.995
    assert np.abs(pre["conic_opacity"][vis] - d["conic_opacity"][vis]).max() < 1e-4
    assert _close_csv(pre["rgb"][vis], d["rgb"][vis]).all()
    b = o.binning(pre["xy"], pre["depths"], pre["radii"], pre["tiles_touched"], args[5], args[6])
    assert b["R"] == 91378
    assert np.array_equal(b["point_offsets"], d["point_offsets"])
    assert np.array_equal(b["point_list"], d["point_list"])                          # bit-exact sort order
    assert np.array_equal(b["ranges"], d["ranges"])                                  # bit-exact tile ranges


def test_preprocess_jvp_matches_reference_fixture(golden):
    d = golden
    kw = _fixture_inputs(d)
    args = [kw.pop(k) for k in ("means3D", "opacities", "viewmatrix", "projmatrix", "campos", "W", "H", "tan_fovx",
                                "tan_fovy")]
    pre = o.preprocess(*args, tangents=_camera_tangents(), **kw)
    vis = pre["radii"] > 0
    for mine, ref, scale_tol in ((pre["t_depths"], d["t_depth"], 2e-5), (pre["t_xy"], d["t_means2D"], 2e-5),
                                 (pre["t_conic_opacity"], d["t_conic_opacity"], 1e-4), (pre["t_rgb"], d["t_rgb"], 2e-5)):
        a, r = np.asarray(mine)[vis], np.asarray(ref)[vis]
        assert rel_l2(a, r) < 1e-5
        assert np.abs(a - r).max() <= scale_tol * max(1.0, np.abs(r).max())


def test_f32_and_f64_instantiations_agree():
    g, cams = ss.make_config("A", views=1)
    c = ss.camera_dict(cams[0])
    a = o.render_forward(g, c, dtype=np.float32)
    b = o.render_forward(g, c, dtype=np.float64)
    # a handful of pixels can flip a hit/skip decision between precisions; the bulk must agree tightly
    err = np.abs(a["color"] - b["color"])
    assert np.quantile(err, 0.999) < 1e-5
    assert (a["radii"] != b["radii"]).mean() < 1e-3


def _small_scene(seed=3, P=300, W=48, H=32, deg=3, opacity_cap=0.9, extent=1.0):
    g = ss.make_gaussians(P, sh_degree=deg, seed=seed, scale_mult=0.6, extent=extent)
    g["opacities"] = np.minimum(g["opacities"], opacity_cap).astype(np.float32)   # no 0.99 clamp: J and J^T consistent
    cam = ss.hemisphere_rig(3, W, H)[1]
    return g, ss.camera_dict(cam, bg=(0.2, 0.1, 0.3))


@pytest.mark.parametrize("deg,aa", [(0, False), (3, False), (2, True)])
def test_oracle_jvp_matches_finite_differences(deg, aa):
    g, c = _small_scene(deg=deg)
    c["antialiasing"] = aa
    t = ss.random_tangent(g, seed=11)
    out = o.render_forward(g, c, tangents=t, dtype=np.float64)
    eps = 1e-6

    def shifted(s):
        g2 = dict(g)
        for k in t:
            g2[k] = g[k].astype(np.float64) + s * eps * t[k]
        return o.render_forward(g2, c, dtype=np.float64)

    p, m = shifted(+1), shifted(-1)
    fd = (p["color"] - m["color"]) / (2 * eps)
    fdd = (p["invdepth"] - m["invdepth"]) / (2 * eps)
    # pixels whose hit set changes under the perturbation are discontinuities, as in the reference tests
    # (preprocess_jvp_test.cu:61-63 skips Gaussians whose radius flips): compare where n_contrib agrees
    ok = (p["n_contrib"] == m["n_contrib"]) & (p["n_contrib"] == out["n_contrib"])
    assert ok.mean() > 0.9
    err = np.abs(fd - out["t_color"])[:, ok]
    assert np.quantile(err, 0.99) <= 1e-3 * np.abs(fd).max()
    assert np.quantile(np.abs(fdd - out["t_invdepth"])[:, ok], 0.99) <= 1e-3 * np.abs(fdd).max()


@pytest.mark.parametrize("deg", [0, 3])
def test_oracle_adjoint_identity(deg):
    """<u, J v> == <J^T u, v>  (tests/test_matvec.py:81-83, rel 1e-4) in the float64 instantiation.

    The identity holds for the reference's VJP only away from three documented reference quirks, which the
    oracle reproduces on purpose (DESIGN.md "quirks"): (1) alpha clamped at 0.99 is not masked in backward.cu:569
    (opacities are capped at 0.9 here), (2) the 1.3*tanfov clamp of t.x/t.y treats t.x as independent of t.z in
    backward.cu:300-312 (extent = 0.5 keeps every Gaussian inside the limit), (3) the antialiasing branch
    evaluates its closed-form derivative at the DILATED covariance (backward.cu:236-246), so J^T is not the
    adjoint of J when antialiasing = True (not exercised here; see test_aa_backward_is_reference_quirk)."""
    g, c = _small_scene(deg=deg, seed=5, extent=0.5)
    t = ss.random_tangent(g, seed=12)
    st = o.render_forward(g, c, tangents=t, dtype=np.float64)
    rng = np.random.default_rng(1)
    u = rng.normal(size=st["color"].shape)
    ud = rng.normal(size=st["invdepth"].shape)
    gr = o.render_backward(g, c, st, u, ud, dtype=np.float64)
    lhs = float((u * st["t_color"]).sum() + (ud * st["t_invdepth"]).sum())
    rhs = float((gr["dL_dmeans3D"] * t["means3D"]).sum() + (gr["dL_dscales"] * t["scales"]).sum()
                + (gr["dL_drotations"] * t["rotations"]).sum() + (gr["dL_dopacity"].reshape(-1, 1) * t["opacities"]).sum()
                + (gr["dL_dsh"] * t["shs"]).sum())
    assert abs(lhs - rhs) <= 1e-6 * max(abs(lhs), abs(rhs), 1.0)


def test_aa_backward_is_reference_quirk():
    """With antialiasing the reference VJP is NOT the adjoint of its JVP (quirk 3 above); with the opacity tangent
    only (which does not pass through the dilation derivative) the identity still holds."""
    g, c = _small_scene(deg=1, seed=5, extent=0.5)
    c["antialiasing"] = True
    t = {k: np.zeros_like(v) for k, v in ss.random_tangent(g, seed=12).items()}
    t["opacities"] = ss.random_tangent(g, seed=12)["opacities"]
    st = o.render_forward(g, c, tangents=t, dtype=np.float64)
    u = np.random.default_rng(1).normal(size=st["color"].shape)
    gr = o.render_backward(g, c, st, u, None, dtype=np.float64)
    lhs = float((u * st["t_color"]).sum())
    rhs = float((gr["dL_dopacity"].reshape(-1, 1) * t["opacities"]).sum())
    assert abs(lhs - rhs) <= 1e-6 * max(abs(lhs), abs(rhs), 1.0)


def test_binning_edge_cases():
    # empty scene: zero instances, all ranges (0,0)
    b = o.binning(np.zeros((0, 2)), np.zeros(0), np.zeros(0, np.int32), np.zeros(0, np.uint32), 64, 48)
    assert b["R"] == 0 and not b["ranges"].any()
    # equal depths in the same tile keep ascending Gaussian id (stable sort)
    xy = np.array([[8.0, 8.0]] * 5, np.float32)
    dep = np.array([2.0, 1.0, 2.0, 1.0, 2.0], np.float32)
    rad = np.full(5, 3, np.int32)
    tt = np.ones(5, np.uint32)
    b = o.binning(xy, dep, rad, tt, 16, 16)
    assert b["point_list"].tolist() == [1, 3, 0, 2, 4]
    assert b["ranges"].tolist() == [[0, 5]]
    assert o.higher_msb(1) == 1 and o.higher_msb(1200) == 11 and o.higher_msb(8160) == 13
