"""Size-independent properties of the matrix-free normal product at BASELINE config C's full per-view size
(1 000 000 Gaussians, 1920x1080, SH degree 3; 3 of the 64 views so the test stays in seconds), where the CPU oracle
cannot follow: the fused J^T J v must equal matvec_T(matvec(v)) on the same cached linearisation, be symmetric and
linear, its quadratic form must equal <Jv, Jv>, and every map must be bit-reproducible (no atomics on the path).
The reference pins the same path by self-consistency only (SURVEY 8c "not pinned by anything"; tests/test_matvec.py
checks <u, Jv> == <J^T u, v> at rel 1e-4)."""
import functools
from types import SimpleNamespace

import pytest
import torch

import scene_synth as ss

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def full():
    from scene.gaussian_model import GaussianModel
    from solver.batch_training_loss import batch_training_loss
    from solver.solver_functions import LinearSolverFunctions
    g, cams_np = ss.make_config("C", sh_degree=3, views=3)
    assert g["means3D"].shape[0] == 1_000_000 and (cams_np[0].image_width, cams_np[0].image_height) == (1920, 1080)
    model = GaussianModel.from_activated(g, device="cuda", num_cameras=len(cams_np), requires_grad=False)
    cams = [ss.torch_camera(c, "cuda", original_image=torch.zeros(3, 1080, 1920, device="cuda"), uid=i)
            for i, c in enumerate(cams_np)]
    pipe = SimpleNamespace(debug=False, antialiasing=False, compute_cov3D_python=False, convert_SHs_python=False)
    loss_func = functools.partial(batch_training_loss, iteration=1, opt=SimpleNamespace(lambda_dssim=0.2), pipe=pipe,
                                  bg=torch.zeros(3, device="cuda"), train_test_exp=False,
                                  depth_l1_weight=lambda it: 0.0, disable_ssim=True)
    with torch.no_grad():
        F = LinearSolverFunctions(loss_func, model, cams, batch_size=1)
        F.evaluate_loss()
        yield model, F


def _state(model, seed):
    from solver.gaussian_model_state import GaussianModelState
    s = GaussianModelState.from_gaussians(model)
    s.packed().copy_(torch.randn(s.length, generator=torch.Generator().manual_seed(seed)).cuda())
    s.exposure_grad.zero_()
    return s


def _rel(a, b):
    a, b = a.packed().double(), b.packed().double()
    return float((a - b).norm() / b.norm())


def test_fused_normal_product_equals_composition_at_full_size(full):
    model, F = full
    with torch.no_grad():
        v = _state(model, 1)
        N, quad = F.normal_matvec(v, return_quad=True)
        Jv = F.matvec(v)
        ref = F.matvec_T(Jv)
        assert _rel(N, ref) < 1e-4                      # same linearisation, different summation order only
        jj = F.dot(Jv, Jv)
        assert abs(float(quad) - jj) <= 1e-4 * jj
        assert abs(F.dot(N, v) - jj) <= 1e-3 * jj       # v^T (J^T J v) == |J v|^2


def test_normal_product_symmetric_and_linear_at_full_size(full):
    model, F = full
    with torch.no_grad():
        v, w = _state(model, 2), _state(model, 3)
        Nv, Nw = F.normal_matvec(v), F.normal_matvec(w)
        a, b = F.dot(w, Nv), F.dot(v, Nw)
        scale = (F.dot(v, Nv) * F.dot(w, Nw)) ** 0.5    # Cauchy-Schwarz bound of the bilinear form
        assert abs(a - b) <= 1e-4 * scale
        comb = F.normal_matvec(v * 2.0 + w * -0.5)
        assert _rel(comb, Nv * 2.0 + Nw * -0.5) < 1e-4


def test_maps_are_bit_reproducible_at_full_size(full):
    model, F = full
    with torch.no_grad():
        v = _state(model, 4)
        n1 = F.normal_matvec(v).packed().clone()
        n2 = F.normal_matvec(v).packed()
        assert torch.equal(n1, n2)
        j1 = [b.Ll1_per_pixel.clone() for b in F.matvec(v).batch_losses]
        j2 = [b.Ll1_per_pixel for b in F.matvec(v).batch_losses]
        assert all(torch.equal(x, y) for x, y in zip(j1, j2))
        u = F.loss * 1.0
        t1 = F.matvec_T(u).packed().clone()
        assert torch.equal(t1, F.matvec_T(u).packed())


def test_adjoint_identity_at_full_size(full):
    """<u, J v> == <J^T u, v> up to the two documented non-adjoint quirks of the reference's kernels (a clamped alpha
    has zero tangent but a live gradient; same for the 1.3 tan(fov) clamp), which touch a small
    fraction of the hits of this scene: 1e-2 here, 1e-4 in the quirk-free test of test_gpu_solver.py."""
    model, F = full
    with torch.no_grad():
        v = _state(model, 5)
        Jv = F.matvec(v)
        u = F.loss * 1.0 + Jv * 0.1
        lhs, rhs = F.dot(u, Jv), F.dot(F.matvec_T(u), v)
        assert abs(lhs - rhs) <= 1e-2 * max(abs(lhs), abs(rhs))


def _quirk_gaussians(model, F):
    """Gaussians whose parameters enter one of the reference's two NON-adjoint code paths in some view of the fixture:
    (1) opacity > 0.99: alpha = min(0.99, o G) can clamp -- zero tangent in renderCUDAJvpSplit, live gradient in
        backward.cu:569;
    (2) |x/z| > 1.3 tan(fovx) or |y/z| > 1.3 tan(fovy) in camera space: computeCov2D clamps t.x / t.y
        (forward.cu:84-90), the backward pass treats the clamped coordinate as independent of t.z
        (backward.cu:300-312)."""
    import math
    xyz = model._xyz.detach()
    flag = torch.sigmoid(model._opacity.detach()).view(-1) > 0.99
    ones = torch.ones_like(xyz[:, :1])
    for vc in F._views:
        t = torch.cat((xyz, ones), dim=1) @ vc.cam.world_view_transform          # row-vector convention
        tx, ty = math.tan(vc.cam.FoVx * 0.5), math.tan(vc.cam.FoVy * 0.5)
        z = t[:, 2]
        flag |= (z > 0.2) & (((t[:, 0] / z).abs() > 1.3 * tx) | ((t[:, 1] / z).abs() > 1.3 * ty))
    return flag


def test_adjoint_identity_at_full_size_without_the_reference_quirks(full):
    """how much of the 1e-2 above is the documented quirks?  With the tangent restricted to the
    Gaussians that touch neither quirk, <u, J v> == <J^T u, v> holds to 1e-4 at the full per-view size -- the
    tolerance of the reference's own tests/test_matvec.py -- so the whole gap of the unrestricted identity is the two
    quirks the kernels reproduce on purpose."""
    model, F = full
    with torch.no_grad():
        quirk = _quirk_gaussians(model, F)
        frac = float(quirk.float().mean())
        assert 0.0 < frac < 0.2                      # the restriction keeps most of the scene
        v = _state(model, 5)
        for t in v.tensors()[:6]:
            t[quirk] = 0
        Jv = F.matvec(v)
        u = F.loss * 1.0 + Jv * 0.1
        JTu = F.matvec_T(u)
        lhs, rhs = F.dot(u, Jv), F.dot(JTu, v)
        assert abs(lhs - rhs) <= 1e-4 * max(abs(lhs), abs(rhs)), (lhs, rhs, frac)
        # and the gap of the unrestricted identity sits on the quirk Gaussians' rows
        w = _state(model, 5)
        for t in w.tensors()[:6]:
            t[~quirk] = 0
        Jw = F.matvec(w)
        gap = abs(F.dot(u, Jw) - F.dot(JTu, w))
        print(f"quirk Gaussians: {100 * frac:.2f} %; identity gap on their rows {gap:.4e}, on the others {abs(lhs - rhs):.4e}")
