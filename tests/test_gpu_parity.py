"""GPU parity tests (run with -m gpu on the B200 box): the sm_100a kernels, called through the C ABI
(diff_gaussian_rasterization._C -> libgslm_b200.so), against
  * the reference's own golden vectors (tests/golden/dgr_fixture.npz)   -- bit-exact integers,
  * the CPU oracle on the same seeded inputs                            -- BASELINE.json tolerances:
        images rel 1e-4; gradients, JVPs and matvecs rel 1e-3.
Tolerances are written next to each assertion."""
import numpy as np
import pytest
import torch

import gs_oracle as o
import scene_synth as ss
import helpers as h

pytestmark = pytest.mark.gpu

IMG_TOL = 1e-4     # rendered images (north_star)
DER_TOL = 1e-3     # gradients, JVPs, matvecs (north_star)


@pytest.fixture(scope="module")
def C():
    from diff_gaussian_rasterization import _C
    _C.lib()
    return _C


def _scene(name="A", view=1, deg=3, P=None, bg=(0.1, 0.2, 0.3), aa=False, extent=None, seed=None, W=None, H=None):
    if name == "small":
        g = ss.make_gaussians(P or 1500, sh_degree=deg, seed=seed or 21, scale_mult=0.8, extent=extent or 1.0)
        cam = ss.hemisphere_rig(3, W or 150, H or 100)[view]
    else:
        g, cams = ss.make_config(name, sh_degree=deg, views=max(view + 1, 2), P=P)
        cam = cams[view]
    return g, cam, ss.camera_dict(cam, bg=bg, antialiasing=aa)


def test_golden_fixture_binning_bit_exact(C, golden):
    """The reference's own dump (P=17073, 640x480, 40x30 tiles, R=91378): tiles_touched, point_offsets,
    point_list and tile ranges must be IDENTICAL."""
    d = golden
    dev = "cuda"
    t = lambda k, shape=None: torch.from_numpy(np.ascontiguousarray(d[k])).to(dev)
    gd = {"means3D": t("means3D"), "opacities": t("opacities"), "scales": t("scales"), "rotations": t("rotations"),
          "shs": t("shs"), "sh_degree": int(d["sh_degree"])}
    W, H = int(d["W"]), int(d["H"])
    ct = dict(bg=t("bg"), viewmatrix=t("viewmatrix"), projmatrix=t("projmatrix"),
              campos=torch.from_numpy(d["campos_raw"][:3].copy()).to(dev), tan_fovx=float(d["tan_fovx"]),
              tan_fovy=float(d["tan_fovy"]), H=H, W=W)
    R, color, radii, geom, binning, img, invd = h.c_forward(C, gd, ct, scale_modifier=float(d["scale_modifier"]),
                                                            keep_sort_scratch=True)
    assert R == 91378
    v = C.buffer_view(gd["means3D"].shape[0], W, H, R, geom, binning, img, C.LAST_SORT_SCRATCH)
    assert np.array_equal(v["point_offsets"].cpu().numpy().astype(np.uint32), d["point_offsets"])
    assert np.array_equal(v["point_list"].cpu().numpy().astype(np.uint32), d["point_list"])
    assert np.array_equal(v["ranges"].cpu().numpy().astype(np.uint32), d["ranges"])
    vis = (radii > 0).cpu().numpy()
    assert np.array_equal(vis, np.diff(np.concatenate([[0], d["point_offsets"].astype(np.int64)])) > 0)
    # primal floats to the CSV's 6 significant digits
    assert np.abs(v["means2D"].cpu().numpy()[vis] - d["means2D"][vis]).max() < 2e-3
    assert np.abs(v["depths"].cpu().numpy()[vis] - d["depth"][vis]).max() < 1e-4
    assert np.abs(v["conic_opacity"].cpu().numpy()[vis] - d["conic_opacity"][vis]).max() < 1e-4
    assert np.abs(v["rgb"].cpu().numpy()[vis] - d["rgb"][vis]).max() < 1e-5


@pytest.mark.parametrize("name,deg,aa", [("A", 3, False), ("A", 0, False), ("A", 2, True), ("small", 1, False)])
def test_forward_vs_oracle(C, name, deg, aa):
    g, cam, c = _scene(name, deg=deg, aa=aa)
    gd, ct = h.to_dev(g), h.cam_tensors(cam, bg=c["bg"])
    R, color, radii, geom, binning, img, invd = h.c_forward(C, gd, ct, antialiasing=aa, keep_sort_scratch=True)
    ref = o.render_forward(g, c)
    P, W, H = g["means3D"].shape[0], c["W"], c["H"]
    v = C.buffer_view(P, W, H, R, geom, binning, img, C.LAST_SORT_SCRATCH)
    # integers from the per-Gaussian stage: a one-ulp difference between nvcc's FMA contraction and the
    # uncontracted CPU oracle may flip a radius; count, do not tolerate more than 0.1 %
    radii_np = radii.cpu().numpy()
    assert (radii_np != ref["radii"]).mean() <= 1e-3
    # binning is integer work: bit-exact given the GPU's own (xy, depth, radii)
    b = o.binning(v["means2D"].cpu().numpy(), v["depths"].cpu().numpy(), radii_np,
                  v["tiles_touched"].cpu().numpy().astype(np.uint32), W, H)
    assert b["R"] == R
    assert np.array_equal(v["point_offsets"].cpu().numpy().astype(np.uint32), b["point_offsets"])
    assert np.array_equal(v["keys_sorted"].cpu().numpy().astype(np.uint64), b["keys"])
    assert np.array_equal(v["point_list"].cpu().numpy().astype(np.uint32), b["point_list"])
    assert np.array_equal(v["ranges"].cpu().numpy().astype(np.uint32), b["ranges"])
    # image: rel 1e-4 (scale = largest reference value)
    assert h.relerr(color.cpu().numpy(), ref["color"]) < IMG_TOL
    assert h.relerr(invd.cpu().numpy(), ref["invdepth"]) < IMG_TOL
    assert h.relerr(v["final_T"].cpu().numpy(), ref["final_T"]) < IMG_TOL
    assert (v["n_contrib"].cpu().numpy() != ref["n_contrib"]).mean() < 1e-3


@pytest.mark.parametrize("name,deg,aa,invdepth", [("A", 3, False, False), ("A", 1, True, True), ("small", 0, False, True)])
def test_backward_vs_oracle(C, name, deg, aa, invdepth):
    g, cam, c = _scene(name, deg=deg, aa=aa)
    gd, ct = h.to_dev(g), h.cam_tensors(cam, bg=c["bg"])
    fwd = h.c_forward(C, gd, ct, antialiasing=aa)
    ref = o.render_forward(g, c)
    rng = np.random.default_rng(5)
    u = rng.normal(size=ref["color"].shape).astype(np.float32)
    ud = rng.normal(size=ref["invdepth"].shape).astype(np.float32) if invdepth else None
    grads = h.c_backward(C, gd, ct, fwd, torch.from_numpy(u).cuda(),
                         torch.from_numpy(ud).cuda() if invdepth else None, antialiasing=aa)
    gref = o.render_backward(g, c, ref, u, ud)
    names = ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
             "dL_drotations")
    for k, got in zip(names, grads):
        if k == "dL_dcolors":
            continue   # only meaningful with colors_precomp (see test_precomputed_inputs)
        a, r = got.cpu().numpy().reshape(-1), np.asarray(gref[k]).reshape(-1)
        assert h.rel_l2(a, r) < DER_TOL, k
        assert h.relerr(a, r) < 5 * DER_TOL, k


@pytest.mark.parametrize("name,deg,aa", [("A", 3, False), ("A", 2, True), ("small", 0, False)])
def test_jvp_vs_oracle(C, name, deg, aa):
    g, cam, c = _scene(name, deg=deg, aa=aa)
    t = ss.random_tangent(g, seed=9)
    gd, td, ct = h.to_dev(g), h.to_dev(t), h.cam_tensors(cam, bg=c["bg"])
    e = torch.empty(0, device="cuda")
    z = lambda x: torch.zeros_like(x)
    out = C.rasterize_gaussians_jvp(ct["bg"], z(ct["bg"]), gd["means3D"], td["means3D"], e, e, gd["opacities"],
                                    td["opacities"], gd["scales"], td["scales"], gd["rotations"], td["rotations"],
                                    1.0, 0.0, e, e, ct["viewmatrix"], z(ct["viewmatrix"]), ct["projmatrix"],
                                    z(ct["projmatrix"]), ct["tan_fovx"], 0.0, ct["tan_fovy"], 0.0, ct["H"], ct["W"],
                                    gd["shs"], td["shs"], deg, ct["campos"], z(ct["campos"]), False, aa, False)
    R, color, radii, geom, binning, img, invd, color_t, invd_t = out
    ref = o.render_forward(g, c, tangents=t)
    assert h.relerr(color.cpu().numpy(), ref["color"]) < IMG_TOL            # primal(JVP) == primal
    assert h.rel_l2(color_t.cpu().numpy(), ref["t_color"]) < DER_TOL
    assert h.rel_l2(invd_t.cpu().numpy(), ref["t_invdepth"]) < DER_TOL
    # per-pixel: 99.9 % of pixels within 1e-3 of the image-tangent scale (a flipped hit decision moves a pixel)
    err = np.abs(color_t.cpu().numpy() - ref["t_color"])
    assert np.quantile(err, 0.999) < DER_TOL * np.abs(ref["t_color"]).max()
    # the cached entry point reproduces it bit for bit
    jv, jd = C.jvp_cached(h.gauss_args(gd), h.tangent_args(td), h.cam_args(ct, aa), radii, geom, R, binning, img,
                          want_invdepth=True)
    assert torch.equal(jv, color_t) and torch.equal(jd, invd_t)


def test_camera_tangents_vs_oracle(C):
    g, cam, c = _scene("small", deg=2)
    gd, ct = h.to_dev(g), h.cam_tensors(cam, bg=c["bg"])
    i = np.arange(16)
    tc = {"viewmatrix": (1e-2 * ((i % 4) * 0.03 + (i + 2) * 0.01)).astype(np.float32),
          "projmatrix": (1e-2 * ((i % 4) * 0.02 + (i + 2) * 0.02)).astype(np.float32),
          "campos": (np.arange(3) * 0.014 + 0.03).astype(np.float32), "bg": np.array([0.3, -0.2, 0.1], np.float32),
          "tan_fovx": 0.01, "tan_fovy": -0.02, "scale_modifier": 0.05}
    e = torch.empty(0, device="cuda")
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    out = C.rasterize_gaussians_jvp(ct["bg"], T(tc["bg"]), gd["means3D"], e, e, e, gd["opacities"], e, gd["scales"], e,
                                    gd["rotations"], e, 1.0, tc["scale_modifier"], e, e, ct["viewmatrix"],
                                    T(tc["viewmatrix"]).view(4, 4), ct["projmatrix"], T(tc["projmatrix"]).view(4, 4),
                                    ct["tan_fovx"], tc["tan_fovx"], ct["tan_fovy"], tc["tan_fovy"], ct["H"], ct["W"],
                                    gd["shs"], e, 2, ct["campos"], T(tc["campos"]), False, False, False)
    ref = o.render_forward(g, c, tangents=tc)
    assert h.rel_l2(out[7].cpu().numpy(), ref["t_color"]) < DER_TOL
    assert h.rel_l2(out[8].cpu().numpy(), ref["t_invdepth"]) < DER_TOL


@pytest.mark.parametrize("name,deg,weighted", [("A", 3, False), ("small", 1, True)])
def test_fused_jtjv_equals_vjp_of_jvp(C, name, deg, weighted):
    """J^T W J v from the fused tile kernel == backward(W * jvp(v)) from the two stand-alone kernels, and the
    by-product v^T J^T W J v == sum W (J v)^2.  rel 1e-3 (north_star: matvecs)."""
    g, cam, c = _scene(name, deg=deg)
    t = ss.random_tangent(g, seed=10)
    gd, td, ct = h.to_dev(g), h.to_dev(t), h.cam_tensors(cam, bg=c["bg"])
    fwd = h.c_forward(C, gd, ct)
    R, color, radii, geom, binning, img, invd = fwd
    jv, _ = C.jvp_cached(h.gauss_args(gd), h.tangent_args(td), h.cam_args(ct), radii, geom, R, binning, img)
    w = (torch.rand_like(jv) > 0.3).float() * 2.0 if weighted else None
    two = h.c_backward(C, gd, ct, fwd, jv * w if weighted else jv)
    P, M = g["means3D"].shape[0], g["shs"].shape[1]
    zz = lambda *s: torch.zeros(s, device="cuda")
    outs = (zz(P, 3), zz(P, 3), zz(P, 1), zz(P, 3), zz(P, 6), zz(P, M, 3), zz(P, 3), zz(P, 4))
    quad = torch.zeros(1, device="cuda")
    C.jtjv(h.gauss_args(gd), h.tangent_args(td), h.cam_args(ct), radii, geom, R, binning, img, w, outs, out_vJJv=quad)
    for i in (0, 2, 3, 4, 5, 6, 7):
        assert h.rel_l2(outs[i].cpu().numpy(), two[i].cpu().numpy()) < DER_TOL, i
    q_ref = float(((jv * jv * w) if weighted else (jv * jv)).double().sum())
    assert abs(float(quad) - q_ref) <= DER_TOL * abs(q_ref)
    # against the oracle: J^T (W J v) with J v from the oracle too
    ref = o.render_forward(g, c, tangents=t)
    u = ref["t_color"] * (w.cpu().numpy() if weighted else 1.0)
    gref = o.render_backward(g, c, ref, u.astype(np.float32))
    for i, k in ((2, "dL_dopacity"), (3, "dL_dmeans3D"), (5, "dL_dsh"), (6, "dL_dscales"), (7, "dL_drotations")):
        assert h.rel_l2(outs[i].cpu().numpy().reshape(-1), gref[k].reshape(-1)) < 2 * DER_TOL, k
    # accumulate semantics: a second call adds the same again
    C.jtjv(h.gauss_args(gd), h.tangent_args(td), h.cam_args(ct), radii, geom, R, binning, img, w, outs)
    assert h.rel_l2(outs[5].cpu().numpy(), 2 * two[5].cpu().numpy()) < DER_TOL


def test_jtjv_batch_equals_sum_of_views(C):
    """gslm_jtjv_batch over 19 views (two chunks of the 16-view launch limit, two image sizes, some weights None)
    == the per-view fused calls accumulated in the same order (to the last bits), in accumulate and in overwrite
    mode, and bit-reproducible from run to run."""
    g = ss.make_gaussians(3000, sh_degree=2, seed=21, scale_mult=0.8)
    t = ss.random_tangent(g, seed=22)
    gd, td = h.to_dev(g), h.to_dev(t)
    cams = ss.hemisphere_rig(10, 112, 80) + ss.hemisphere_rig(9, 96, 64)
    views, singles = [], []
    gen = torch.Generator().manual_seed(5)
    for i, cam in enumerate(cams):
        ct = h.cam_tensors(cam, bg=(0.1, 0.0, 0.2))
        R, color, radii, geom, binning, img, invd = h.c_forward(C, gd, ct)
        w = None if i % 3 == 0 else (torch.rand(color.shape, generator=gen) > 0.3).float().cuda() * 2.0
        views.append((h.cam_args(ct), radii, geom, R, binning, img, w))
    P, M = g["means3D"].shape[0], g["shs"].shape[1]
    zz = lambda *s: torch.zeros(s, device="cuda")
    mk = lambda: (zz(P, 3), None, zz(P, 1), zz(P, 3), None, zz(P, M, 3), zz(P, 3), zz(P, 4))
    ref, quad_ref = mk(), torch.zeros(1, device="cuda")
    for v in views:
        C.jtjv(h.gauss_args(gd), h.tangent_args(td), *v, ref, out_vJJv=quad_ref)
    plan = C.JtjvBatchPlan(views, torch.device("cuda"))
    out, quad = mk(), torch.zeros(1, device="cuda")
    C.jtjv_batch(h.gauss_args(gd), h.tangent_args(td), plan, out, out_vJJv=quad)
    # same arithmetic and the same view order; the multi-view adjoint kernel is a different instantiation of the
    # per-Gaussian adjoint, so multiply-adds may contract differently: last-bit differences only
    for a, b in zip(out, ref):
        if a is not None:
            assert h.relerr(a.cpu().numpy(), b.cpu().numpy()) < 1e-6
    # run to run the batched path is bit-reproducible (no atomics anywhere on the gradient path)
    again = mk()
    C.jtjv_batch(h.gauss_args(gd), h.tangent_args(td), plan, again)
    for a, b in zip(out, again):
        if a is not None:
            assert torch.equal(a, b)
    assert abs(float(quad) - float(quad_ref)) <= 1e-5 * abs(float(quad_ref))
    assert float(ref[5].abs().sum()) > 0
    # overwrite mode: garbage in the sinks is replaced, not added to
    out2 = tuple(None if o is None else torch.full_like(o, 7.0) for o in out)
    C.jtjv_batch(h.gauss_args(gd), h.tangent_args(td), plan, out2, accumulate=False)
    # (the overwrite and accumulate kernels contract their last multiply-add differently: last-bit differences)
    for a, b in zip(out2, ref):
        if a is not None:
            assert h.relerr(a.cpu().numpy(), b.cpu().numpy()) < 1e-6


def test_adjoint_identity_gpu(C):
    """<u, J v> == <J^T u, v>, rel 1e-4 (tests/test_matvec.py:81-83) on a scene away from the reference's
    non-adjoint quirks (see tests/test_oracle_golden.py::test_oracle_adjoint_identity)."""
    g, cam, c = _scene("small", deg=3, extent=0.5, P=3000)
    g["opacities"] = np.minimum(g["opacities"], 0.9).astype(np.float32)
    t = ss.random_tangent(g, seed=13)
    gd, td, ct = h.to_dev(g), h.to_dev(t), h.cam_tensors(cam, bg=c["bg"])
    fwd = h.c_forward(C, gd, ct)
    R, color, radii, geom, binning, img, invd = fwd
    jv, jd = C.jvp_cached(h.gauss_args(gd), h.tangent_args(td), h.cam_args(ct), radii, geom, R, binning, img,
                          want_invdepth=True)
    u, ud = torch.randn_like(jv), torch.randn_like(jd)
    grads = h.c_backward(C, gd, ct, fwd, u, ud)
    lhs = float((u.double() * jv.double()).sum() + (ud.double() * jd.double()).sum())
    rhs = sum(float((grads[i].double().reshape(-1) * td[k].double().reshape(-1)).sum())
              for i, k in ((2, "opacities"), (3, "means3D"), (5, "shs"), (6, "scales"), (7, "rotations")))
    assert abs(lhs - rhs) <= 1e-4 * max(abs(lhs), abs(rhs))


def test_precomputed_inputs(C):
    """colors_precomp + cov3D_precomp paths (rasterize_points.cu:90-102; forward.cu:204-213,252-260)."""
    g, cam, c = _scene("small", deg=0)
    pre = o.preprocess(g["means3D"], g["opacities"], c["viewmatrix"], c["projmatrix"], c["campos"], c["W"], c["H"],
                       c["tan_fovx"], c["tan_fovy"], scales=g["scales"], rotations=g["rotations"], shs=g["shs"])
    rng = np.random.default_rng(3)
    g2 = {"means3D": g["means3D"], "opacities": g["opacities"], "sh_degree": 0,
          "colors_precomp": rng.uniform(0, 1, size=g["means3D"].shape).astype(np.float32),
          "cov3D_precomp": None}
    # Sigma = R diag(s^2) R^T (forward.cu:117-151), for every Gaussian
    r, x, y, z = (g["rotations"][:, i].astype(np.float64) for i in range(4))
    Rm = np.stack([np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y)], -1),
                   np.stack([2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x)], -1),
                   np.stack([2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], -1)], -2)
    Sig = Rm @ (g["scales"].astype(np.float64)[:, :, None] ** 2 * np.swapaxes(Rm, 1, 2))
    cov = np.stack([Sig[:, 0, 0], Sig[:, 0, 1], Sig[:, 0, 2], Sig[:, 1, 1], Sig[:, 1, 2], Sig[:, 2, 2]], -1)
    cov = cov.astype(np.float32)
    vis = pre["radii"] > 0
    assert np.abs(cov[vis] - pre["cov3D"][vis]).max() < 1e-5 * np.abs(cov).max()
    g2["cov3D_precomp"] = cov
    gd, ct = h.to_dev(g2), h.cam_tensors(cam, bg=c["bg"])
    fwd = h.c_forward(C, gd, ct)
    ref = o.render_forward(g2, c)
    assert fwd[0] == ref["num_rendered"]
    assert h.relerr(fwd[1].cpu().numpy(), ref["color"]) < IMG_TOL
    u = rng.normal(size=ref["color"].shape).astype(np.float32)
    grads = h.c_backward(C, gd, ct, fwd, torch.from_numpy(u).cuda())
    gref = o.render_backward(g2, c, ref, u)
    for i, k in ((1, "dL_dcolors"), (2, "dL_dopacity"), (3, "dL_dmeans3D"), (4, "dL_dcov3D")):
        assert h.rel_l2(grads[i].cpu().numpy().reshape(-1), np.asarray(gref[k]).reshape(-1)) < DER_TOL, k


def test_edge_cases(C):
    dev = "cuda"
    cam = ss.hemisphere_rig(2, 50, 37)[0]            # ragged: neither side a multiple of 16
    ct = h.cam_tensors(cam, bg=(0.3, 0.6, 0.9))
    # P == 0 (rasterize_points.cu:105): zero outputs, empty buffers
    g0 = {"means3D": torch.zeros(0, 3, device=dev), "opacities": torch.zeros(0, 1, device=dev),
          "scales": torch.zeros(0, 3, device=dev), "rotations": torch.zeros(0, 4, device=dev),
          "shs": torch.zeros(0, 1, 3, device=dev), "sh_degree": 0}
    R, color, radii, geom, binning, img, invd = h.c_forward(C, g0, ct)
    assert R == 0 and radii.numel() == 0 and float(color.abs().sum()) == 0.0
    # everything behind the camera: R == 0, image == background
    g = ss.make_gaussians(64, sh_degree=0, seed=1)
    g["means3D"] = (g["means3D"] * 0.1 + 3.0 * cam.camera_center).astype(np.float32)
    gd = h.to_dev(g)
    fwd = h.c_forward(C, gd, ct)
    assert fwd[0] == 0 and int((fwd[2] != 0).sum()) == 0
    assert torch.allclose(fwd[1], ct["bg"].view(3, 1, 1).expand(3, 37, 50))
    grads = h.c_backward(C, gd, ct, fwd, torch.ones_like(fwd[1]))
    assert all(float(x.abs().sum()) == 0.0 for x in grads)
    assert not bool(C.mark_visible(gd["means3D"], ct["viewmatrix"], ct["projmatrix"]).any())
    # ragged image with content: parity + mark_visible
    g, _, c = _scene("small", deg=1, W=50, H=37)
    c = ss.camera_dict(cam, bg=(0.3, 0.6, 0.9))
    gd = h.to_dev(g)
    fwd = h.c_forward(C, gd, ct)
    ref = o.render_forward(g, c)
    assert fwd[0] == ref["num_rendered"]
    assert h.relerr(fwd[1].cpu().numpy(), ref["color"]) < IMG_TOL
    vis = C.mark_visible(gd["means3D"], ct["viewmatrix"], ct["projmatrix"]).cpu().numpy()
    assert np.array_equal(vis, o.mark_visible(g["means3D"], c["viewmatrix"]))
    # one very opaque wall of Gaussians in front: early termination (T < 1e-4) path, long tile lists (> 1 batch)
    g = ss.make_gaussians(6000, sh_degree=0, seed=2, extent=0.3, scale_mult=6.0)
    g["opacities"][:] = 0.95
    gd = h.to_dev(g)
    fwd = h.c_forward(C, gd, ct)
    ref = o.render_forward(g, c)
    assert fwd[0] == ref["num_rendered"]
    assert h.relerr(fwd[1].cpu().numpy(), ref["color"]) < IMG_TOL
    v = C.buffer_view(6000, 50, 37, fwd[0], fwd[3], fwd[4], fwd[5])
    assert (v["n_contrib"].cpu().numpy() != ref["n_contrib"]).mean() < 1e-2
    u = torch.randn_like(fwd[1])
    grads = h.c_backward(C, gd, ct, fwd, u)
    gref = o.render_backward(g, c, ref, u.cpu().numpy())
    assert h.rel_l2(grads[3].cpu().numpy(), gref["dL_dmeans3D"]) < DER_TOL


def test_autograd_surface(C):
    """GaussianRasterizer: autograd backward and forward-AD through autograd.Function.jvp, incl. under no_grad
    (solver/solver_functions.py:86)."""
    import math
    import torch.autograd.forward_ad as fwAD
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    g, cam, c = _scene("small", deg=2)
    t = ss.random_tangent(g, seed=4)
    gd, td, ct = h.to_dev(g), h.to_dev(t), h.cam_tensors(cam, bg=c["bg"])
    rs = GaussianRasterizationSettings(ct["H"], ct["W"], ct["tan_fovx"], ct["tan_fovy"], ct["bg"], 1.0,
                                       ct["viewmatrix"].view(4, 4), ct["projmatrix"].view(4, 4), 2, ct["campos"],
                                       False, False, False)
    rast = GaussianRasterizer(rs)
    leaves = {k: gd[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "shs", "scales", "rotations")}
    m2d = torch.zeros_like(leaves["means3D"], requires_grad=True)
    color, radii, invd = rast(leaves["means3D"], m2d, leaves["opacities"], shs=leaves["shs"],
                              scales=leaves["scales"], rotations=leaves["rotations"])
    u = torch.randn_like(color)
    (color * u).sum().backward()
    ref = o.render_forward(g, c, tangents=t)
    gref = o.render_backward(g, c, ref, u.cpu().numpy())
    assert h.relerr(color.detach().cpu().numpy(), ref["color"]) < IMG_TOL
    assert h.rel_l2(leaves["means3D"].grad.cpu().numpy(), gref["dL_dmeans3D"]) < DER_TOL
    assert h.rel_l2(leaves["shs"].grad.cpu().numpy(), gref["dL_dsh"]) < DER_TOL
    assert h.rel_l2(m2d.grad.cpu().numpy(), gref["dL_dmeans2D"]) < DER_TOL
    with torch.no_grad(), fwAD.dual_level():
        d = {k: fwAD.make_dual(gd[k], td[k]) for k in ("means3D", "opacities", "shs", "scales", "rotations")}
        out, _, outd = rast(d["means3D"], torch.zeros_like(gd["means3D"]), d["opacities"], shs=d["shs"],
                            scales=d["scales"], rotations=d["rotations"])
        tan = fwAD.unpack_dual(out).tangent
        assert tan is not None
        assert h.rel_l2(tan.cpu().numpy(), ref["t_color"]) < DER_TOL
        assert h.rel_l2(fwAD.unpack_dual(outd).tangent.cpu().numpy(), ref["t_invdepth"]) < DER_TOL


def test_compact_cache_structure(C):
    """The compact linearisation cache: tile offsets are a scan, every compact record carries its id, the CSR
    (g_off, g_list) is a permutation of the compact positions grouped by Gaussian and ascending inside a group,
    n_contrib_c never exceeds the tile's compact count, and a pixel's compact count equals the number of
    instances it blended or skipped before its last hit that SOME pixel of the tile blended."""
    g, cam, c = _scene("A", deg=1)
    gd, ct = h.to_dev(g), h.cam_tensors(cam, bg=c["bg"])
    R, color, radii, geom, binning, img, invd = h.c_forward(C, gd, ct, keep_sort_scratch=True)
    P, W, H = g["means3D"].shape[0], c["W"], c["H"]
    v = C.buffer_view(P, W, H, R, geom, binning, img, C.LAST_SORT_SCRATCH)
    Rc = v["num_cached"]
    assert 0 < Rc <= int(R) and R.cached == Rc
    cstart = v["cstart"].cpu().numpy().astype(np.int64)
    assert cstart[0] == 0 and cstart[-1] == Rc and (np.diff(cstart) >= 0).all()
    ranges = v["ranges"].cpu().numpy().astype(np.int64)
    assert (np.diff(cstart) <= ranges[:, 1] - ranges[:, 0]).all()
    rec = v["records"]
    cid = v["cid"].cpu().numpy()
    assert np.array_equal(rec[:, 10].contiguous().view(torch.int32).cpu().numpy(), cid)
    # each tile's compact list is a subsequence of its full sorted list
    pl = v["point_list"].cpu().numpy()
    for t in (0, len(ranges) // 2, len(ranges) - 1):
        full = pl[ranges[t, 0]:ranges[t, 1]].tolist()
        it = iter(full)
        assert all(any(x == y for y in it) for x in cid[cstart[t]:cstart[t + 1]].tolist())
    g_off = v["g_off"].cpu().numpy().astype(np.int64)
    g_list = v["g_list"].cpu().numpy().astype(np.int64)
    assert g_off[0] == 0 and g_off[-1] == Rc
    assert np.array_equal(np.sort(g_list), np.arange(Rc))
    owner = np.repeat(np.arange(P), np.diff(g_off))
    assert np.array_equal(cid[g_list], owner)
    same = owner[1:] == owner[:-1]
    assert (np.diff(g_list)[same] > 0).all()
    # per pixel
    ncc = v["n_contrib_c"].cpu().numpy().astype(np.int64)
    gx = (W + 15) // 16
    tile_of = (np.arange(H)[:, None] // 16) * gx + (np.arange(W)[None, :] // 16)
    assert (ncc <= np.diff(cstart)[tile_of]).all()
    assert ((ncc == 0) == (v["n_contrib"].cpu().numpy() == 0)).all()
    # hit masks: bit l of word k of a compact instance <=> pixel (row 2k + l // 16, column l % 16) of its tile
    # blended it.  No mask is empty, and a pixel's last set bit sits at its n_contrib_c - 1.
    masks = v["masks"].cpu().numpy().view(np.uint32)
    assert masks.shape == (Rc, 8) and (masks != 0).any(axis=1).all()
    for t in (0, len(ranges) // 3, len(ranges) // 2, len(ranges) - 1):
        m = masks[cstart[t]:cstart[t + 1]]
        ty, tx = divmod(t, gx)
        for k in range(8):
            for l in range(32):
                y, x = ty * 16 + 2 * k + l // 16, tx * 16 + l % 16
                hit = np.nonzero((m[:, k] >> np.uint32(l)) & np.uint32(1))[0]
                last = int(hit[-1]) + 1 if len(hit) else 0
                if y < H and x < W:
                    assert last == ncc[y, x], (t, k, l)
                else:
                    assert last == 0
    # derivative-form records: q1.y = log2(opacity kept in q2.w)
    assert torch.allclose(rec[:, 5], torch.log2(rec[:, 11]), atol=1e-5)


@pytest.mark.parametrize("deg", [3, 1])
def test_dense_multi_view_adjoint_with_repeated_views(C, deg):
    """The dense multi-view adjoint kernel (k_preprocess_bwd_views_dense: SH rows of 16 / 4 coefficients) against the
    per-view fused calls accumulated in the same order.  Six of the views are the SAME camera, so nearly every Gaussian
    is held by several lanes of one pass and the phased store sites (rank among equal ids, up to six deep) carry the
    whole sum; the remaining views mix sizes across the 16-view chunk boundary."""
    g = ss.make_gaussians(4000, sh_degree=deg, seed=31, scale_mult=0.9)
    t = ss.random_tangent(g, seed=32)
    gd, td = h.to_dev(g), h.to_dev(t)
    rig = ss.hemisphere_rig(9, 128, 96)
    cams = [rig[2]] * 4 + rig[:5] + [rig[2]] * 2 + rig[5:] + ss.hemisphere_rig(6, 96, 64)
    assert len(cams) > 16
    views = []
    gen = torch.Generator().manual_seed(6)
    for i, cam in enumerate(cams):
        ct = h.cam_tensors(cam, bg=(0.0, 0.1, 0.2))
        R, color, radii, geom, binning, img, invd = h.c_forward(C, gd, ct)
        w = None if i % 4 == 0 else (torch.rand(color.shape, generator=gen) > 0.2).float().cuda() * 2.0
        views.append((h.cam_args(ct), radii, geom, R, binning, img, w))
    P, M = g["means3D"].shape[0], g["shs"].shape[1]
    assert M % 4 == 0          # the layout the dense kernel's float4 row updates cover
    zz = lambda *s: torch.zeros(s, device="cuda")
    mk = lambda: (zz(P, 3), None, zz(P, 1), zz(P, 3), None, zz(P, M, 3), zz(P, 3), zz(P, 4))
    ref = mk()
    for v in views:
        C.jtjv(h.gauss_args(gd), h.tangent_args(td), *v, ref)
    plan = C.JtjvBatchPlan(views, torch.device("cuda"))
    out = mk()
    C.jtjv_batch(h.gauss_args(gd), h.tangent_args(td), plan, out)
    for name, a, b in zip(("means2D", "", "opacity", "means3D", "", "sh", "scales", "rotations"), out, ref):
        if a is not None:
            assert float(b.abs().sum()) > 0, name
            assert h.relerr(a.cpu().numpy(), b.cpu().numpy()) < 1e-6, name
    again = mk()
    C.jtjv_batch(h.gauss_args(gd), h.tangent_args(td), plan, again)
    for a, b in zip(out, again):
        if a is not None:
            assert torch.equal(a, b)
    # J^T u through the batched backward of the same plan (matvec_T's path) against per-view backward calls
    cots = [torch.randn((3, v[0][5], v[0][6]), generator=gen).cuda() for v in views]
    ref_b = mk()
    for v, cot in zip(views, cots):
        bgc, vm, pm, tfx, tfy, H, W, campos, aa, dbg = v[0]
        e = torch.empty(0, device="cuda")
        C.rasterize_gaussians_backward(bgc, gd["means3D"], v[1], e, gd["opacities"], gd["scales"], gd["rotations"], 1.0, e,
                                       vm, pm, tfx, tfy, cot, None, gd["shs"], deg, campos, v[2], v[3], v[4], v[5], aa, dbg,
                                       accumulate_into=tuple(x if x is not None else zz(P, 3 if k == 1 else 6)
                                                             for k, x in enumerate(ref_b)))
    out_b = mk()
    C.backward_batch(h.gauss_args(gd), plan, cots, out_b)
    for name, a, b in zip(("means2D", "", "opacity", "means3D", "", "sh", "scales", "rotations"), out_b, ref_b):
        if a is not None:
            assert h.relerr(a.cpu().numpy(), b.cpu().numpy()) < 1e-6, name
