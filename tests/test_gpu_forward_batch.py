"""GPU: the batched linearising forward (gslm_forward_batch, what LinearSolverFunctions.linearize issues) against the
reference-shaped per-view entry point (gslm_forward, pinned on the reference's kernels in test_gpu_vs_reference.py):
images, radii, instance counts and the whole linearisation cache bit for bit, with mixed view sizes and lanes on and
off; the fused residual epilogue against the torch composition it replaces (gaussian_renderer/batch_render.py:118,
solver/batch_training_loss.py:10-17); gslm_count_instances against the counts of the real passes."""
import math

import numpy as np
import pytest
import torch

import scene_synth as ss
import helpers as h

pytestmark = pytest.mark.gpu


def _views(sizes, V):
    views, cams = [], []
    for i, (W, H) in enumerate(sizes):
        cam = ss.hemisphere_rig(V, W, H)[i % V]
        ct = h.cam_tensors(cam, bg=(0.1, 0.2, 0.3))
        cams.append(ct)
        views.append((ct["viewmatrix"], ct["projmatrix"], ct["tan_fovx"], ct["tan_fovy"], ct["H"], ct["W"], ct["campos"]))
    return views, cams


@pytest.mark.parametrize("culling", [False, True])
def test_forward_batch_equals_per_view_forward_bit_for_bit(culling):
    from diff_gaussian_rasterization import _C
    g = ss.make_gaussians(20000, sh_degree=3, seed=5, scale_mult=0.9)
    gd = h.to_dev(g)
    sizes = [(320, 200), (160, 96), (320, 200), (208, 144), (320, 200), (160, 96), (176, 176)]
    views, cams = _views(sizes, 5)
    e = torch.empty(0, device="cuda")
    batch = _C.forward_batch(cams[0]["bg"], gd["means3D"], e, gd["opacities"], gd["scales"], gd["rotations"], 1.0, e,
                             views, gd["shs"], 3, False, False, tile_culling=culling, want_invdepth=True)
    assert len(batch) == len(views)
    counts = _C.count_instances(cams[0]["bg"], gd["means3D"], e, gd["opacities"], gd["scales"], gd["rotations"], 1.0, e,
                                views, gd["shs"], 3, False, tile_culling=culling)
    for i, (ct, got) in enumerate(zip(cams, batch)):
        ref = _C.rasterize_gaussians(ct["bg"], gd["means3D"], e, gd["opacities"], gd["scales"], gd["rotations"], 1.0, e,
                                     ct["viewmatrix"], ct["projmatrix"], ct["tan_fovx"], ct["tan_fovy"], ct["H"], ct["W"],
                                     gd["shs"], 3, ct["campos"], False, False, False, tile_culling=culling)
        R, color, radii, geom, binning, img, invd = got
        R0, color0, radii0, geom0, binning0, img0, invd0 = ref
        assert int(R) == int(R0) == counts[i] and R.cached == R0.cached > 0
        assert torch.equal(color, color0) and torch.equal(invd, invd0) and torch.equal(radii, radii0)
        P, W, H = g["means3D"].shape[0], ct["W"], ct["H"]
        a = _C.buffer_view(P, W, H, R, geom, binning, img)
        b = _C.buffer_view(P, W, H, R0, geom0, binning0, img0)
        for k in ("records", "masks", "cid", "g_list", "ranges", "n_contrib", "n_contrib_c", "final_T", "cstart",
                  "tiles_touched", "point_offsets", "depths", "means2D", "conic_opacity", "rgb", "cov3D", "clamped", "g_off"):
            x, y = a[k], b[k]
            if x.is_floating_point():
                x, y = x.view(torch.int32), y.view(torch.int32)
            if k in ("depths", "means2D", "conic_opacity", "rgb", "cov3D", "clamped"):      # written for visible Gaussians only
                vis = radii0 > 0
                x, y = x.view(P, -1)[vis], y.view(P, -1)[vis]
            assert torch.equal(x, y), (i, k)


def test_residual_epilogue_matches_the_torch_composition():
    from diff_gaussian_rasterization import _C
    g = ss.make_gaussians(8000, sh_degree=2, seed=9, scale_mult=1.2)
    g["shs"][:, 0] = np.abs(g["shs"][:, 0]) * 2.5     # colours exceed 1 (per-Gaussian rgb is clamped at 0 by forward.cu:66-72): the clamp mask matters
    gd = h.to_dev(g)
    sizes = [(200, 120), (136, 88)]
    views, cams = _views(sizes, 3)
    e = torch.empty(0, device="cuda")
    Hm, Wm = 120, 200
    gen = torch.Generator(device="cuda").manual_seed(1)
    gt = torch.rand((2, 3, Hm, Wm), device="cuda", generator=gen)
    am = (torch.rand((2, 1, Hm, Wm), device="cuda", generator=gen) > 0.3).float()
    resid = torch.zeros((2, 3, Hm, Wm), device="cuda")
    specs, masks, weights = [], [], []
    for i, (W, H) in enumerate(sizes):
        m, w = torch.empty((3, H, W), device="cuda"), torch.empty((3, H, W), device="cuda")
        masks.append(m); weights.append(w)
        specs.append(_C.ResidualSpec(image=resid[i, :, :H, :W], mask=m, weight=w, gt=gt[i, :, :H, :W],
                                     alpha_mask=am[i, 0, :H, :W], weight_scale=2.0))
    out = _C.forward_batch(cams[0]["bg"], gd["means3D"], e, gd["opacities"], gd["scales"], gd["rotations"], 1.0, e, views,
                           gd["shs"], 2, False, False, residual=specs, tile_culling=True)
    for i, (W, H) in enumerate(sizes):
        color = out[i][1]
        assert float((color > 1).float().mean()) > 0.001
        m = ((color >= 0) & (color <= 1)).float() * am[i, :, :H, :W]
        want = color.clamp(0, 1) * am[i, :, :H, :W] - gt[i, :, :H, :W]
        assert torch.equal(resid[i, :, :H, :W], want)
        assert torch.equal(masks[i], m) and torch.equal(weights[i], 2.0 * m * m)
    assert float(resid[1, :, 88:, :].abs().sum()) == 0.0 and float(resid[1, :, :, 136:].abs().sum()) == 0.0   # padding untouched
    # ssim-mode outputs: clamped image without mask / target, clamp derivative into a strided stack
    img = torch.zeros((2, 3, Hm, Wm), device="cuda")
    cm = torch.zeros((2, 3, Hm, Wm), device="cuda")
    specs = [_C.ResidualSpec(image=img[i, :, :H, :W], mask=cm[i, :, :H, :W], apply_alpha_mask=False, subtract_gt=False)
             for i, (W, H) in enumerate(sizes)]
    out2 = _C.forward_batch(cams[0]["bg"], gd["means3D"], e, gd["opacities"], gd["scales"], gd["rotations"], 1.0, e, views,
                            gd["shs"], 2, False, False, residual=specs, tile_culling=True)
    for i, (W, H) in enumerate(sizes):
        color = out2[i][1]
        assert torch.equal(color, out[i][1])
        assert torch.equal(img[i, :, :H, :W], color.clamp(0, 1))
        assert torch.equal(cm[i, :, :H, :W], ((color >= 0) & (color <= 1)).float())


def test_residual_epilogue_of_a_scene_without_gaussians():
    """P = 0: the images are all zero, as in the reference (DGR/rasterize_points.cu:69,96: out_color is zero-filled and
    the rasterizer is not called), and the residual / mask / weight outputs are still written (the solver hands them
    over uninitialised)."""
    from diff_gaussian_rasterization import _C
    sizes = [(72, 40), (48, 56)]
    views, cams = _views(sizes, 2)
    e = torch.empty(0, device="cuda")
    z = lambda *s: torch.zeros(s, device="cuda")
    bg = torch.tensor([0.25, 1.5, -0.5], device="cuda")
    gen = torch.Generator(device="cuda").manual_seed(2)
    specs, keep = [], []
    for W, H in sizes:
        gt = torch.rand((3, H, W), device="cuda", generator=gen)
        am = (torch.rand((1, H, W), device="cuda", generator=gen) > 0.3).float()
        r, m, w = (torch.full((3, H, W), float("nan"), device="cuda") for _ in range(3))
        keep.append((gt, am, r, m, w))
        specs.append(_C.ResidualSpec(image=r, mask=m, weight=w, gt=gt, alpha_mask=am[0], weight_scale=2.0))
    out = _C.forward_batch(bg, z(0, 3), e, z(0, 1), z(0, 3), z(0, 4), 1.0, e, views, z(0, 16, 3), 3, False, False,
                           residual=specs, tile_culling=True)
    for (W, H), o, (gt, am, r, m, w) in zip(sizes, out, keep):
        color = o[1]
        assert tuple(color.shape) == (3, H, W) and float(color.abs().max()) == 0.0
        assert torch.equal(m, am.expand(3, H, W)) and torch.equal(w, 2.0 * m * m)      # 0 lies inside the clamp
        assert torch.equal(r, -gt)


def test_render_batch_lanes_are_bit_identical(monkeypatch):
    """gslm_render_batch on its internal streams == one gslm_render call per view."""
    from diff_gaussian_rasterization import _C
    g = ss.make_gaussians(15000, sh_degree=1, seed=3)
    gd = h.to_dev(g)
    sizes = [(256, 160)] * 9 + [(128, 96)] * 2
    views, cams = _views(sizes, 6)
    e = torch.empty(0, device="cuda")
    counts, cols = _C.render_batch(cams[0]["bg"], gd["means3D"], e, gd["opacities"], gd["scales"], gd["rotations"], 1.0, e,
                                   views, gd["shs"], 1, True, False, tile_culling=True)
    for ct, n, c in zip(cams, counts, cols):
        n0, c0, _, _ = _C.render_only(ct["bg"], gd["means3D"], e, gd["opacities"], gd["scales"], gd["rotations"], 1.0, e,
                                      ct["viewmatrix"], ct["projmatrix"], ct["tan_fovx"], ct["tan_fovy"], ct["H"], ct["W"],
                                      gd["shs"], 1, ct["campos"], False, True, False, tile_culling=True)
        assert n == n0 and torch.equal(c, c0)
