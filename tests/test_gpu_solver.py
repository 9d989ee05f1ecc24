"""GPU tests of the solver-facing path (run with -m gpu): batch_render == per-view render, the cached
(linearisation-reusing, fused) mode of LinearSolverFunctions == the generic autograd / forward-AD mode (which is
the reference's own composition), adjoint identity, and a CG solve on the fused J^T J v."""
import functools
from types import SimpleNamespace

import numpy as np
import pytest
import torch

import scene_synth as ss
import helpers as h

pytestmark = pytest.mark.gpu


def _setup(P=4000, V=4, W=128, H=96, deg=2, seed=3, masks=True):
    from scene.gaussian_model import GaussianModel
    from solver.batch_training_loss import batch_training_loss
    g = ss.make_gaussians(P, sh_degree=deg, seed=seed, scale_mult=0.8)
    g["rotations"] = (g["rotations"] * 1.7).astype(np.float32)       # raw quaternions are NOT unit: exercises normalize
    model = GaussianModel.from_activated(g, device="cuda", num_cameras=V)
    gen = torch.Generator(device="cpu").manual_seed(seed)
    cams = []
    for i, c in enumerate(ss.hemisphere_rig(V, W, H)):
        gt = torch.rand(3, H, W, generator=gen).cuda()
        am = (torch.rand(1, H, W, generator=gen) > 0.2).float().cuda() if (masks and i == 1) else None
        cams.append(ss.torch_camera(c, "cuda", original_image=gt, alpha_mask=am, uid=i))
    pipe = SimpleNamespace(debug=False, antialiasing=False, compute_cov3D_python=False, convert_SHs_python=False)
    opt = SimpleNamespace(lambda_dssim=0.2)
    bg = torch.tensor([0.0, 0.0, 0.0], device="cuda")
    loss_func = functools.partial(batch_training_loss, iteration=1, opt=opt, pipe=pipe, bg=bg, train_test_exp=False,
                                  depth_l1_weight=lambda it: 0.0, disable_ssim=True)
    return model, cams, loss_func, pipe, bg


def _rand_state(model, seed):
    from solver.gaussian_model_state import GaussianModelState
    s = GaussianModelState.from_gaussians(model)
    s.packed().copy_(torch.randn(s.length, generator=torch.Generator().manual_seed(seed)).cuda())
    s.exposure_grad.zero_()
    return s


def test_batch_render_equals_per_view_render():
    """tests/test_batch_render.py:82-89 of the reference: render/depth allclose(atol=1e-6), radii exact."""
    from gaussian_renderer import render
    from gaussian_renderer.batch_render import batch_render
    model, cams, _, pipe, bg = _setup()
    with torch.no_grad():
        pkg = batch_render(cams, model, pipe, bg)
        singles = [render(c, model, pipe, bg) for c in cams]
    assert pkg["render"].shape == (len(cams), 3, 96, 128)
    for i, s in enumerate(singles):
        assert torch.allclose(pkg["render"][i], s["render"], atol=1e-6)
        assert torch.allclose(pkg["depth"][i], s["depth"], atol=1e-6)
    radii = torch.stack([s["radii"] for s in singles])
    assert torch.equal(pkg["max_radii"], radii.max(dim=0).values)
    assert torch.equal(pkg["viewcount"], (radii > 0).sum(dim=0))
    assert torch.equal(pkg["visibility_filter"], (radii.max(dim=0).values > 0).nonzero())


def test_batch_gradients_equal_sum_of_single_view_gradients():
    """tests/test_batch_training_loss.py:104-110 of the reference (atol 1e-5 on grads)."""
    from gaussian_renderer import render
    from gaussian_renderer.batch_render import batch_render
    model, cams, _, pipe, bg = _setup(P=1500)
    u = torch.randn(len(cams), 3, 96, 128, device="cuda")
    model.zero_grad()
    pkg = batch_render(cams, model, pipe, bg)
    (pkg["render"] * u).sum().backward()
    names = ("_xyz", "_features_dc", "_features_rest", "_scaling", "_rotation", "_opacity")
    batch = [getattr(model, n).grad.clone() for n in names]
    model.zero_grad()
    for i, c in enumerate(cams):
        (render(c, model, pipe, bg)["render"] * u[i]).sum().backward()
    for n, b in zip(names, batch):
        s = getattr(model, n).grad
        assert torch.allclose(b, s, atol=1e-5 * max(1.0, float(s.abs().max()))), n


def test_cached_mode_equals_generic_mode():
    from solver.solver_functions import LinearSolverFunctions
    model, cams, loss_func, pipe, bg = _setup()
    Fc = LinearSolverFunctions(loss_func, model, cams, batch_size=3)
    Fg = LinearSolverFunctions(loss_func, model, cams, batch_size=3, cached=False)
    assert Fc.cached and not Fg.cached
    Lc, Lg = Fc.evaluate_loss(), Fg.evaluate_loss()
    assert abs(float(Fc.loss_scalar) - float(Fg.loss_scalar)) <= 1e-5 * float(Fg.loss_scalar)
    for a, b in zip(Lc.batch_losses, Lg.batch_losses):
        assert torch.allclose(a.Ll1_per_pixel, b.Ll1_per_pixel, atol=1e-6)
    v = _rand_state(model, 7)
    Jc, Jg = Fc.matvec(v), Fg.matvec(v)
    for a, b in zip(Jc.batch_losses, Jg.batch_losses):
        assert h.rel_l2(a.Ll1_per_pixel.cpu().numpy(), b.Ll1_per_pixel.cpu().numpy()) < 1e-3
    u = Jg * 0.5 + Lg * 1.0
    Tc, Tg = Fc.matvec_T(u), Fg.matvec_T(u)
    for n, a, b in zip(("xyz", "dc", "rest", "scaling", "rotation", "opacity"), Tc.tensors(), Tg.tensors()):
        assert h.rel_l2(a.cpu().numpy(), b.cpu().numpy()) < 1e-3, n
    # fused J^T J v == J^T (J v) of the generic (reference-composition) mode; quad == <Jv, Jv>
    Nc, quad = Fc.normal_matvec(v, return_quad=True)
    Ng = Fg.matvec_T(Fg.matvec(v))
    for n, a, b in zip(("xyz", "dc", "rest", "scaling", "rotation", "opacity"), Nc.tensors(), Ng.tensors()):
        assert h.rel_l2(a.cpu().numpy(), b.cpu().numpy()) < 1e-3, n
    q_ref = Fg.dot(Jg, Jg)
    assert abs(float(quad) - q_ref) <= 1e-3 * abs(q_ref)
    # damping
    from solver.gaussian_model_state import GaussianModelDampMatrix
    D = GaussianModelDampMatrix(5e2, 5e2, 5e2, 5e-2, 5e-2, 1e1, 0.0)
    Nd = Fc.normal_matvec(v, damp=D)
    assert h.rel_l2(Nd.packed().cpu().numpy(), (Nc + v * D).packed().cpu().numpy()) < 1e-6


def test_adjoint_identity_through_solver_interface():
    """tests/test_matvec.py of the reference (rel 1e-4), on the cached path, away from the documented non-adjoint
    quirks (alpha clamp, cov clamp): small extent, capped opacity."""
    from scene.gaussian_model import GaussianModel
    from solver.solver_functions import LinearSolverFunctions
    model, cams, loss_func, pipe, bg = _setup(masks=False)
    g = ss.make_gaussians(3000, sh_degree=2, seed=11, scale_mult=0.8, extent=0.4)
    g["opacities"] = np.minimum(g["opacities"], 0.9).astype(np.float32)
    model = GaussianModel.from_activated(g, device="cuda", num_cameras=len(cams))
    F = LinearSolverFunctions(loss_func, model, cams)
    F.evaluate_loss()
    v = _rand_state(model, 8)
    Jv = F.matvec(v)
    u = F.loss * 1.0 + Jv * 0.1
    lhs, rhs = F.dot(u, Jv), F.dot(F.matvec_T(u), v)
    assert abs(lhs - rhs) <= 1e-4 * max(abs(lhs), abs(rhs))


def test_cg_on_fused_normal_matvec_and_cgls_agree(capsys):
    """Solve (J^T J + D) s = -J^T r with conjugate_gradient(matvec=normal_matvec) and compare with cgls_damped on
    (matvec, matvec_T): same Krylov space => same iterate after the same number of steps (rel 1e-3)."""
    from solver.conjugate_gradient import cgls_damped, conjugate_gradient
    from solver.gaussian_model_state import GaussianModelDampMatrix, GaussianModelParamGroupMask
    from solver.solver_functions import LinearSolverFunctions
    model, cams, loss_func, pipe, bg = _setup(P=2000)
    pm = GaussianModelParamGroupMask(mask_xyz=True, mask_exposure=True)        # train_jvp.py:221-227
    D = GaussianModelDampMatrix(5e2, 5e2, 5e2, 5e-2, 5e-2, 1e1, 1.0)
    F = LinearSolverFunctions(loss_func, model, cams, param_mask=pm)
    F.evaluate_loss()
    b = -1 * F.loss
    rhs = F.matvec_T(b)
    x0 = F.get_initial_solution()
    s_cg = conjugate_gradient(lambda p: F.normal_matvec(p, damp=D), lambda x, y: x + y, lambda x, y: x - y,
                              lambda x, y: F.dot(x, y), lambda a, x: a * x, rhs, x0, max_iter=3, tol=0.0)
    s_ls = cgls_damped(F.matvec, F.matvec_T, F.dot, F.saxpy, b, x0, damp=D, max_iter=3, restart_iter=3,
                       recompute_residual=False)
    capsys.readouterr()
    assert float(s_cg.xyz_grad.abs().sum()) == 0.0
    assert h.rel_l2(s_cg.packed().cpu().numpy(), s_ls.packed().cpu().numpy()) < 1e-3
    lin = F.saxpy(-1.0, F.matvec(s_cg), b)
    assert F.dot(lin, lin) + F.dot(s_cg, s_cg, D) < F.dot(b, b)


@pytest.mark.parametrize("masked", [True, False])
def test_device_resident_cg_matches_host_cg(masked):
    """solver.lm.cg_normal_equations (all scalars on the device, p^T A p by a fixed-order dot) == the reference-style
    conjugate_gradient driven through Python floats, after the same number of iterations (rel 1e-4), and it never
    synchronises: the residual norms come back as one device tensor."""
    from solver.conjugate_gradient import conjugate_gradient
    from solver.gaussian_model_state import GaussianModelDampMatrix, GaussianModelParamGroupMask
    from solver.lm import cg_normal_equations
    from solver.solver_functions import LinearSolverFunctions
    model, cams, loss_func, pipe, bg = _setup(P=2000)
    pm = GaussianModelParamGroupMask(mask_xyz=True, mask_exposure=True) if masked else None
    D = GaussianModelDampMatrix(5e2, 5e-2, 5e-2, 5e-2, 5e-2, 5e-2, 1e1)        # train_jvp.py:229-235
    F = LinearSolverFunctions(loss_func, model, cams, param_mask=pm)
    F.evaluate_loss()
    rhs = F.matvec_T(-1 * F.loss)
    s_dev, info = cg_normal_equations(F, damp=D, max_iter=4, tol=0.0)
    s_host = conjugate_gradient(lambda p: F.normal_matvec(p, damp=D), lambda x, y: x + y, lambda x, y: x - y,
                                lambda x, y: F.dot(x, y), lambda a, x: a * x, rhs, F.get_initial_solution(),
                                max_iter=4, tol=0.0)
    assert h.rel_l2(s_dev.packed().cpu().numpy(), s_host.packed().cpu().numpy()) < 1e-4
    norms = info["residual_norms"].cpu().numpy()
    assert norms.shape == (5,) and norms[-1] < norms[0]
    if masked:
        assert float(s_dev.xyz_grad.abs().sum()) == 0.0
    # a converged solve keeps a static launch sequence: with a huge tolerance every step is masked to zero
    s0, _ = cg_normal_equations(F, damp=D, max_iter=2, tol=10.0)
    assert float(s0.packed().abs().sum()) == 0.0


def test_cuda_graphed_cg_iteration_is_bit_identical():
    """cg_normal_equations(graph=True): iteration 1 eager, iteration 2 captured, 2..n replayed -- the same bits as the
    eager loop (the CG iteration captured in a CUDA graph), in both residual modes."""
    from solver.gaussian_model_state import GaussianModelDampMatrix
    from solver.lm import cg_normal_equations
    from solver.solver_functions import LinearSolverFunctions
    model, cams, loss_func, pipe, bg = _setup(P=2000)
    D = GaussianModelDampMatrix(5e2, 5e-2, 5e-2, 5e-2, 5e-2, 5e-2, 1e1)
    F = LinearSolverFunctions(loss_func, model, cams, batch_size=2)
    F.evaluate_loss()
    s_e, info_e = cg_normal_equations(F, damp=D, max_iter=5, tol=0.0)
    s_g, info_g = cg_normal_equations(F, damp=D, max_iter=5, tol=0.0, graph=True)
    assert info_g["graphed"] and not info_e["graphed"]
    assert torch.equal(s_e.packed(), s_g.packed())
    assert torch.equal(info_e["residual_norms"], info_g["residual_norms"])
    # and the solver object is still usable eagerly afterwards
    s_2, _ = cg_normal_equations(F, damp=D, max_iter=5, tol=0.0)
    assert torch.equal(s_e.packed(), s_2.packed())


def test_lm_step_reduces_validation_loss():
    """One LM step of solver.lm.lm_step (train_jvp.py:237-288: linearise, 2 CG iterations on the damped normal
    equations, back-tracking 2, 1, ..., 1/16) moves a perturbed model towards the images it was rendered from."""
    from gaussian_renderer import render
    from scene.gaussian_model import GaussianModel
    from solver.gaussian_model_state import GaussianModelParamGroupMask
    from solver.lm import lm_step
    model, cams, loss_func, pipe, bg = _setup(P=3000, V=6, masks=False)
    with torch.no_grad():
        for c in cams:
            c.original_image = render(c, model, pipe, bg)["render"].clamp(0, 1)
        gen = torch.Generator().manual_seed(4)
        model._features_dc.data += 0.2 * torch.randn(model._features_dc.shape, generator=gen).cuda()
        model._opacity.data += 0.3 * torch.randn(model._opacity.shape, generator=gen).cuda()
        model._scaling.data += 0.05 * torch.randn(model._scaling.shape, generator=gen).cuda()
    pm = GaussianModelParamGroupMask(mask_xyz=True)
    out = lm_step(loss_func, model, cams[:4], val_cams=cams, param_mask=pm, cg_iters=3, batch_size=2)
    assert out["val_loss"] < 0.7 * out["trace"][-1][1] or out["val_loss"] < out["start_loss"]
    assert out["best_alpha"] in (2.0, 1.0, 0.5, 0.25, 0.125, 0.0625)
    first = lm_step(loss_func, model, cams[:4], val_cams=cams, param_mask=pm, cg_iters=3, batch_size=2)
    assert first["val_loss"] <= out["val_loss"] * 1.0001     # a second step does not make things worse


@pytest.mark.parametrize("P,M", [(1001, 16), (257, 1), (4096, 4)])
def test_activation_layer_kernels_match_torch_autograd(P, M):
    """csrc/param_space.cu against PyTorch differentiating the reference's activations
    (scene/gaussian_model.py:192-220: exp, normalize, sigmoid, cat): forward AD for the tangent, autograd for the
    pull-back, odd P (unaligned quaternion block inside the packed state) and M = 1 (no features_rest) included."""
    import torch.autograd.forward_ad as fwAD
    from diff_gaussian_rasterization import _C
    from solver.gaussian_model_state import GaussianModelDampMatrix, GaussianModelState
    gen = torch.Generator(device="cuda").manual_seed(P + M)
    rn = lambda *s: torch.randn(*s, device="cuda", generator=gen)
    raw = [rn(P, 3), rn(P, 1, 3), rn(P, M - 1, 3), rn(P, 3) * 0.3, rn(P, 4), rn(P, 1), torch.zeros(2, 3, 4, device="cuda")]

    def activate(xyz, dc, rest, sc, rot, op):
        return xyz, torch.cat((dc, rest), dim=1), torch.exp(sc), torch.nn.functional.normalize(rot), torch.sigmoid(op)

    v = GaussianModelState._from_shapes(raw)
    v.packed().copy_(rn(v.packed().numel()))
    norm = raw[4].norm(dim=1, keepdim=True)
    layer = _C.Activations(torch.exp(raw[3]), raw[4] / norm, norm, torch.sigmoid(raw[5]), M)
    t = [torch.empty(P, M, 3, device="cuda"), torch.empty(P, 3, device="cuda"), torch.empty(P, 4, device="cuda"),
         torch.empty(P, 1, device="cuda")]
    layer.jvp(v, *t)
    with torch.no_grad(), fwAD.dual_level():
        duals = [fwAD.make_dual(p, tv) for p, tv in zip(raw[:6], v.tensors()[:6])]
        ref_t = [fwAD.unpack_dual(o).tangent for o in activate(*duals)]
    for a, b in zip(t, ref_t[1:]):
        assert torch.allclose(a, b, rtol=1e-5, atol=1e-6)
    g = [rn(P, 3), rn(P, M, 3), rn(P, 3), rn(P, 4), rn(P, 1)]
    leaves = [p.clone().requires_grad_(True) for p in raw[:6]]
    torch.autograd.backward(activate(*leaves), g)
    out = GaussianModelState._from_shapes(raw)
    layer.vjp(*g, out)
    for a, leaf in zip(out.tensors()[:6], leaves):
        assert torch.allclose(a, leaf.grad, rtol=1e-5, atol=1e-6)
    D = GaussianModelDampMatrix(5e2, 5e-2, 3e-2, 7e-2, 2e-1, 4.0, 1e1)
    out2 = GaussianModelState._from_shapes(raw)
    layer.vjp(*g, out2, v, D.values()[:6])
    for a, b, tv, d in zip(out2.tensors()[:6], out.tensors()[:6], v.tensors()[:6], D.values()):
        assert torch.allclose(a, b + d * tv, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("ssim", [False, True])
def test_render_only_loss_value_matches_linearised_loss(ssim):
    """gslm_render (no linearisation cache) gives the same image bits as gslm_forward, and
    LinearSolverFunctions.loss_value() -- the line-search evaluation -- the same ||r||^2 as evaluate_loss()."""
    import math
    from diff_gaussian_rasterization import _C
    from solver.solver_functions import LinearSolverFunctions
    model, cams, loss_func, pipe, bg = _setup()
    if ssim:
        loss_func = functools.partial(loss_func.func, **{**loss_func.keywords, "disable_ssim": False})
    F = LinearSolverFunctions(loss_func, model, cams, batch_size=3)
    F.evaluate_loss()
    ref = float(F.loss_scalar)
    val = float(F.loss_value())
    assert abs(val - ref) <= 1e-5 * abs(ref)
    a, cam, e = F._act, cams[0], torch.empty(0, device="cuda")
    args = (bg, a["means3D"], e, a["opacities"], a["scales"], a["rotations"], 1.0, e, cam.world_view_transform,
            cam.full_proj_transform, math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5), cam.image_height,
            cam.image_width, a["shs"], a["sh_degree"], cam.camera_center, False, False, False)
    full = _C.rasterize_gaussians(*args)
    n, color, radii, invd = _C.render_only(*args)
    assert n == int(full[0]) and torch.equal(color, full[1]) and torch.equal(radii, full[2]) and torch.equal(invd, full[6])


def test_render_batch_equals_per_view_render_bit_for_bit():
    """gslm_render_batch (one host synchronisation for all views) == one gslm_render per view: same instance counts,
    same image bits, on views of different sizes; calling it twice must not depend on the reused scratch; an empty
    view list and an empty scene are legal; loss_value() is a pure function of the parameters."""
    import math
    from diff_gaussian_rasterization import _C
    from solver.solver_functions import LinearSolverFunctions
    model, cams, loss_func, pipe, bg = _setup(P=6000, V=5)
    small = ss.hemisphere_rig(5, 80, 56)[2]
    cams[2] = ss.torch_camera(small, "cuda", original_image=torch.rand(3, 56, 80).cuda(), uid=2)
    F = LinearSolverFunctions(loss_func, model, cams, batch_size=1)
    a, e = F._activated(with_layer=False), torch.empty(0, device="cuda")
    views = [(c.world_view_transform, c.full_proj_transform, math.tan(c.FoVx * 0.5), math.tan(c.FoVy * 0.5),
              int(c.image_height), int(c.image_width), c.camera_center) for c in cams]
    gargs = (bg, a["means3D"], e, a["opacities"], a["scales"], a["rotations"], 1.0, e)
    counts, colors = _C.render_batch(*gargs, views, a["shs"], a["sh_degree"], False, False)
    counts2, colors2 = _C.render_batch(*gargs, views, a["shs"], a["sh_degree"], False, False)
    for v, n, col, col2 in zip(views, counts, colors, colors2):
        n1, c1, _, _ = _C.render_only(*gargs, *v[:6], a["shs"], a["sh_degree"], v[6], False, False, False)
        assert n == n1 and n > 0
        assert torch.equal(col, c1) and torch.equal(col2, c1)
    assert counts2 == counts
    assert _C.render_batch(*gargs, [], a["shs"], a["sh_degree"], False, False) == ([], [])
    z = torch.zeros((0, 3), device="cuda")
    n0, c0 = _C.render_batch(bg, z, e, z[:, :1], z, torch.zeros((0, 4), device="cuda"), 1.0, e, views[:2],
                             torch.zeros((0, 9, 3), device="cuda"), 2, False, False)
    assert n0 == [0, 0] and all(float(c.abs().max()) == 0.0 for c in c0)
    F.evaluate_loss()
    assert abs(float(F.loss_value()) - float(F.loss_scalar)) <= 1e-5 * float(F.loss_scalar)
    assert float(F.loss_value()) == float(F.loss_value())


def test_normal_product_over_several_chunks_equals_sum_of_chunks():
    """More than 16 views: gslm_jtjv_batch walks them in chunks of 16 (one tangent, one tile and one adjoint launch
    per chunk).  The product over 37 views must equal the sum of the products over the three chunks taken alone
    (same per-view arithmetic, same view order => same bits), repeated calls must be bit-identical, and the result
    must be visible to the caller's stream when the call returns."""
    from solver.solver_functions import LinearSolverFunctions
    model, cams, loss_func, pipe, bg = _setup(P=5000, V=37, W=96, H=64)
    with torch.no_grad():
        F = LinearSolverFunctions(loss_func, model, cams, batch_size=8)
        F.evaluate_loss()
        v = _rand_state(model, 21)
        ref, qref = F.normal_matvec(v, return_quad=True)
        ref, qref = ref.packed().clone(), float(qref)
        for _ in range(3):
            assert torch.equal(F.normal_matvec(v).packed(), ref)
        parts, quad = [], 0.0
        for lo in (0, 16, 32):
            G = LinearSolverFunctions(loss_func, model, cams[lo:lo + 16], batch_size=8)
            G.evaluate_loss()
            out, q = G.normal_matvec(v, return_quad=True)
            parts.append(out.packed().clone())
            quad += float(q)
        total = parts[0] + parts[1] + parts[2]
        assert float((total - ref).abs().max()) <= 1e-5 * float(ref.abs().max())   # regrouped fp32 sums
        assert abs(quad - qref) <= 1e-4 * abs(qref)
        assert float(ref.abs().max()) > 0


def test_batched_cached_maps_with_mixed_view_sizes():
    """gslm_jvp_batch / gslm_backward_batch through LinearSolverFunctions.matvec / matvec_T on views of two different
    sizes in one padded batch == the per-view entry points (gslm_jvp_cached, gslm_backward) view by view."""
    from diff_gaussian_rasterization import _C
    from scene.gaussian_model import GaussianModel
    from solver.batch_training_loss import batch_training_loss
    from solver.solver_functions import LinearSolverFunctions
    model, _, loss_func, pipe, bg = _setup(P=3000, V=2)
    gen = torch.Generator().manual_seed(5)
    cams = []
    for (W, H) in ((128, 96), (80, 64)):
        for c in ss.hemisphere_rig(2, W, H):
            cams.append(ss.torch_camera(c, "cuda", original_image=torch.rand(3, H, W, generator=gen).cuda(),
                                        uid=len(cams)))
    cams = [cams[0], cams[2], cams[1], cams[3]]          # sizes interleaved inside the batches
    F = LinearSolverFunctions(loss_func, model, cams, batch_size=3)
    F.evaluate_loss()
    v = _rand_state(model, 9)
    Jv = F.matvec(v)
    gargs, targs = F._gauss_args(), F._tangent_inputs(v)
    k = 0
    for b in Jv.batch_losses:
        for i in range(b.Ll1_per_pixel.shape[0]):
            vc = F._views[k]
            k += 1
            ref, _ = _C.jvp_cached(gargs, targs, vc.cam_args, vc.radii, vc.geom, vc.R, vc.binning, vc.img)
            got = b.Ll1_per_pixel[i]
            assert torch.allclose(got[:, :vc.H, :vc.W], ref * vc.mask, rtol=1e-5, atol=1e-7)
            assert float(got[:, vc.H:, :].abs().sum()) == 0.0 and float(got[:, :, vc.W:].abs().sum()) == 0.0
    u = Jv * 0.5 + F.loss * 1.0
    T = F.matvec_T(u)
    flat, views, sinks = F._act_accumulators()
    flat.zero_()                       # the accumulators come uninitialised (the batched calls overwrite)
    a, e = F._act, torch.empty(0, device="cuda")
    k = 0
    for b in u.batch_losses:
        cot = b.Ll1_per_pixel + b.ssim_loss_per_pixel
        for i in range(cot.shape[0]):
            vc = F._views[k]
            k += 1
            bgc, vm, pm, tfx, tfy, H, W, campos, aa, dbg = vc.cam_args
            _C.rasterize_gaussians_backward(bgc, a["means3D"], vc.radii, e, a["opacities"], a["scales"], a["rotations"],
                                            1.0, e, vm, pm, tfx, tfy, (cot[i, :, :H, :W] * vc.mask).contiguous(), None,
                                            a["shs"], a["sh_degree"], campos, vc.geom, vc.R, vc.binning, vc.img, aa, dbg,
                                            accumulate_into=sinks)
    ref = F._to_state(views)
    assert h.rel_l2(T.packed().cpu().numpy(), ref.packed().cpu().numpy()) < 1e-5


def test_cached_linearisation_follows_the_parameters():
    """the reference's matvec re-renders at the CURRENT parameters on every call
    (solver/solver_functions.py:75-113).  After update_step the cached mode must not answer with J at the old
    parameters: it re-linearises, and the result equals a fresh object's."""
    from solver.solver_functions import LinearSolverFunctions
    model, cams, loss_func, pipe, bg = _setup(P=2500)
    F = LinearSolverFunctions(loss_func, model, cams, batch_size=2)
    F.evaluate_loss()
    v = _rand_state(model, 5)
    before = F.normal_matvec(v).packed().clone()
    with torch.no_grad():
        model.update_step(_rand_state(model, 6) * 0.05)
        stale_views = F._views
        after = F.normal_matvec(v).packed().clone()
        assert F._views is not stale_views, "the linearisation was not rebuilt after update_step"
        fresh = LinearSolverFunctions(loss_func, model, cams, batch_size=2).normal_matvec(v).packed()
    assert torch.equal(after, fresh)
    assert h.rel_l2(after.cpu().numpy(), before.cpu().numpy()) > 1e-3      # the parameters really moved
    # same for J v and J^T u, and for a replaced parameter tensor (load_ply-style)
    with torch.no_grad():
        model._opacity = torch.nn.Parameter(model._opacity.detach() * 0.9)
        jv = F.matvec(v)
        jv_fresh = LinearSolverFunctions(loss_func, model, cams, batch_size=2).matvec(v)
    for a, b in zip(jv.batch_losses, jv_fresh.batch_losses):
        assert torch.equal(a.Ll1_per_pixel, b.Ll1_per_pixel)


def test_cached_mode_raises_on_the_depth_branch_like_the_reference():
    """depth_l1_weight > 0 with a depth_reliable camera raises NotImplementedError in the reference
    (solver/batch_training_loss.py:88-93) and in the generic mode; the cached mode must not drop the term silently."""
    from solver.solver_functions import LinearSolverFunctions
    model, cams, loss_func, pipe, bg = _setup(P=500, V=2)
    cams[1].depth_reliable = True
    lf = functools.partial(loss_func.func, **{**loss_func.keywords, "depth_l1_weight": lambda it: 1.0})
    F = LinearSolverFunctions(lf, model, cams)
    assert F.cached
    with pytest.raises(NotImplementedError):
        F.evaluate_loss()
    with pytest.raises(NotImplementedError):
        F.loss_value()


def test_densification_bookkeeping_matches_the_reference_lines():
    """SURVEY 8 f4 / train_jvp.py:311-316 + scene/gaussian_model.py:561-563: max_radii2D, xyz_gradient_accum and denom
    from ONE kernel (gslm_densification_stats) fed by the cached adjoint's dL_dmeans2D == the reference's torch lines
    applied to what autograd leaves in viewspace_points.grad on the generic path."""
    from gaussian_renderer.batch_render import batch_render
    from solver.solver_functions import LinearSolverFunctions
    model, cams, loss_func, pipe, bg = _setup(P=3000)
    F = LinearSolverFunctions(loss_func, model, cams, batch_size=4)
    F.track_viewspace = True
    F.evaluate_loss()
    F.matvec_T(F.loss)                                   # gradient of 0.5 ||r||^2: what a first-order step would use
    radii = F.max_radii()
    model.init_densification_stats()
    model.max_radii2D[::3] = 7.0                          # pre-existing state must be max-ed / added, not overwritten
    model.xyz_gradient_accum[::5] = 0.25
    ref_max, ref_acc, ref_den = model.max_radii2D.clone(), model.xyz_gradient_accum.clone(), model.denom.clone()
    model.update_densification_stats(radii, F.viewspace_grad)
    # the reference's lines on the autograd path (batch_render's viewspace_points.grad)
    model.zero_grad()
    pkg = batch_render(cams, model, pipe, bg)
    gt = torch.stack([c.original_image for c in cams])
    img = pkg["render"]
    masks = torch.zeros_like(img[:, :1])
    for i, c in enumerate(cams):
        masks[i] = c.alpha_mask if c.alpha_mask is not None else 0.0     # batch quirk: mask-less views of a masked batch are zeroed
    r = img * masks - gt
    (2.0 * 0.5 * (r * r).sum()).backward()               # two aliased residual blocks
    vs_grad = pkg["viewspace_points"].grad
    vis = pkg["visibility_filter"].view(-1)
    assert torch.equal((radii > 0).nonzero().view(-1), vis)
    assert h.rel_l2(F.viewspace_grad.cpu().numpy(), vs_grad.cpu().numpy()) < 1e-3
    rr = pkg["max_radii"].to(ref_max.dtype)
    ref_max[vis] = torch.max(ref_max[vis], rr[vis])
    ref_acc[vis] += torch.norm(vs_grad[vis, :2], dim=-1, keepdim=True)
    ref_den[vis] += 1
    assert torch.equal(model.max_radii2D, ref_max.view(-1))
    assert torch.equal(model.denom, ref_den)
    assert h.rel_l2(model.xyz_gradient_accum.cpu().numpy(), ref_acc.cpu().numpy()) < 1e-3


def test_host_pipeline_streams_host_vectors_through_the_product():
    """solver.HostPipeline (bench.py's end-to-end leg): host-resident vectors go through normal_matvec with the copies
    double-buffered on side streams -- every result equals the device-resident product of the same vector, bit for bit,
    also when the host reuses its buffers as soon as submit() returns the slot."""
    from solver.solver_functions import LinearSolverFunctions
    model, cams, loss_func, pipe, bg = _setup(P=3000)
    F = LinearSolverFunctions(loss_func, model, cams, batch_size=2)
    F.evaluate_loss()
    states = [_rand_state(model, 20 + k) for k in range(5)]
    want = [F.normal_matvec(s).packed().clone() for s in states]
    n = states[0].length
    ins = [torch.empty(n).pin_memory() for _ in states]
    outs = [torch.empty(n).pin_memory() for _ in states]
    for h_in, s in zip(ins, states):
        h_in.copy_(s.packed().cpu())
    hp = F.host_pipeline(depth=2)
    for h_in, h_out in zip(ins, outs):
        hp.submit(h_in, h_out)
    hp.flush()
    for k in range(len(states)):
        assert torch.equal(outs[k].cuda(), want[k]), k
    assert hp.bytes_in == 4 * n and hp.bytes_out == 4 * n
