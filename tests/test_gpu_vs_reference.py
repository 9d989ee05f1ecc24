"""GPU: our sm_100a kernels against the REFERENCE's OWN CUDA kernels (unmodified forward.cu / backward.cu /
rasterizer_impl.cu compiled for sm_100a into oracle/_ref/libgsref.so, see oracle/Makefile and DESIGN.md 2.1),
on the same inputs:
  * radii, tiles_touched, point_offsets, point_list and tile ranges  BIT-EXACT (north_star),
  * rendered image / inverse depth rel 1e-4, gradients rel 1e-3 (north_star; atomic ordering differs).
Skipped when the prebuilt library is absent (it can only be built where /root/reference exists)."""
import numpy as np
import pytest
import torch

import scene_synth as ss
import helpers as h

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def R():
    import gs_ref
    if not gs_ref.available():
        pytest.skip("oracle/_ref/libgsref.so not built (needs the reference tree at build time)")
    gs_ref.lib()
    return gs_ref


@pytest.fixture(scope="module")
def C():
    from diff_gaussian_rasterization import _C
    _C.lib()
    return _C


CASES = [  # (config, view, sh degree, antialiasing, P override)
    ("A", 0, 3, False, None), ("A", 2, 1, True, None), ("B", 5, 3, False, None), ("B", 11, 0, True, None),
    ("C", 17, 3, False, None),      # full headline size: 1M Gaussians, 1920x1080, R ~ 20M
]


@pytest.mark.parametrize("name,view,deg,aa,P", CASES)
def test_forward_bit_exact_binning_and_image(C, R, name, view, deg, aa, P):
    g, cams = ss.make_config(name, sh_degree=deg, views=view + 1, P=P)
    cam = cams[view]
    gd, ct = h.to_dev(g), h.cam_tensors(cam, bg=(0.1, 0.2, 0.3))
    n_ref, color_ref, invd_ref, radii_ref = R.forward(gd, ct, antialiasing=aa)
    ours = h.c_forward(C, gd, ct, antialiasing=aa, keep_sort_scratch=True)
    n, color, radii, geom, binning, img, invd = ours
    Pn, W, H = g["means3D"].shape[0], ct["W"], ct["H"]
    assert n == n_ref
    assert torch.equal(radii, radii_ref)
    st = R.state(Pn, W, H, n_ref)
    v = C.buffer_view(Pn, W, H, n, geom, binning, img, C.LAST_SORT_SCRATCH)
    assert torch.equal(v["tiles_touched"], st["tiles_touched"])
    assert torch.equal(v["point_offsets"], st["point_offsets"])
    assert torch.equal(v["point_list"], st["point_list"])          # tile|depth sort order, bit-exact
    assert torch.equal(v["ranges"], st["ranges"])                  # tile ranges, bit-exact
    vis = radii_ref > 0
    assert torch.equal(v["depths"][vis].view(torch.int32), st["depths"][vis].view(torch.int32))   # sort keys' depth bits
    assert torch.equal(v["means2D"][vis].view(torch.int32), st["means2D"][vis].view(torch.int32))
    # images: rel 1e-4 of the largest value
    assert h.relerr(color.cpu().numpy(), color_ref.cpu().numpy()) < 1e-4
    assert h.relerr(invd.cpu().numpy(), invd_ref.cpu().numpy()) < 1e-4
    assert h.relerr(v["final_T"].cpu().numpy(), st["final_T"].cpu().numpy()) < 1e-4
    # n_contrib may move by one where alpha sits within an ulp of a threshold (exp vs ex2 rounding)
    assert float((v["n_contrib"] != st["n_contrib"]).float().mean()) < 1e-3


@pytest.mark.parametrize("name,view,deg,aa,P", CASES[:4])
def test_backward_matches_reference_kernels(C, R, name, view, deg, aa, P):
    g, cams = ss.make_config(name, sh_degree=deg, views=view + 1, P=P)
    cam = cams[view]
    gd, ct = h.to_dev(g), h.cam_tensors(cam, bg=(0.1, 0.2, 0.3))
    n_ref, color_ref, invd_ref, radii_ref = R.forward(gd, ct, antialiasing=aa)
    gen = torch.Generator(device="cuda").manual_seed(5)
    u = torch.randn(color_ref.shape, device="cuda", generator=gen)
    ud = torch.randn(invd_ref.shape, device="cuda", generator=gen)
    ref = R.backward(gd, ct, radii_ref, u, ud, antialiasing=aa)
    ours_f = h.c_forward(C, gd, ct, antialiasing=aa)
    grads = h.c_backward(C, gd, ct, ours_f, u, ud, antialiasing=aa)
    names = ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
             "dL_drotations")
    for k, got in zip(names, grads):
        if k == "dL_dcolors":
            continue   # with SHs the reference returns the colour gradient before the clamp mask; unused downstream
        a, r = got.cpu().numpy().reshape(-1), ref[k].cpu().numpy().reshape(-1)
        assert h.rel_l2(a, r) < 1e-3, k
        assert h.relerr(a, r) < 5e-3, k


def test_precomputed_colors_match_reference(C, R):
    g, cams = ss.make_config("A", sh_degree=0, views=2)
    cam = cams[1]
    g2 = {"means3D": g["means3D"], "opacities": g["opacities"], "scales": g["scales"], "rotations": g["rotations"],
          "colors_precomp": np.random.default_rng(0).uniform(0, 1, g["means3D"].shape).astype(np.float32),
          "sh_degree": 0}
    gd, ct = h.to_dev(g2), h.cam_tensors(cam, bg=(0.0, 0.0, 0.0))
    n_ref, color_ref, invd_ref, radii_ref = R.forward(gd, ct)
    ours = h.c_forward(C, gd, ct)
    assert ours[0] == n_ref and h.relerr(ours[1].cpu().numpy(), color_ref.cpu().numpy()) < 1e-4
    u = torch.randn_like(color_ref)
    ref = R.backward(gd, ct, radii_ref, u)
    grads = h.c_backward(C, gd, ct, ours, u)
    assert h.rel_l2(grads[1].cpu().numpy(), ref["dL_dcolors"].cpu().numpy()) < 1e-3
    # the reference's own backward leaves NaN in isolated rows on this (SH-less) path; compare the finite rows
    rm = ref["dL_dmeans3D"].cpu().numpy()
    ok = np.isfinite(rm).all(axis=1)
    assert ok.mean() > 0.3      # (thousands of NaN rows in the reference's output: uninitialised reads on its side)
    assert h.rel_l2(grads[3].cpu().numpy()[ok], rm[ok]) < 1e-3
    assert np.isfinite(grads[3].cpu().numpy()).all()
