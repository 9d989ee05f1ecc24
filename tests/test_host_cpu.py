"""CPU tests of the host-side mirror of the reference's solver interface: generalized vectors, CG / CGLS, the
generic (autograd) mode of LinearSolverFunctions on a toy differentiable loss, and the view-sharded multi-process
path over gloo (world_size 2)."""
import functools
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

import scene_synth as ss
from scene.gaussian_model import GaussianModel
from solver.conjugate_gradient import cgls_damped, conjugate_gradient
from solver.gaussian_model_state import (GaussianModelDampMatrix, GaussianModelParamGroupMask,
                                         GaussianModelSplatMask, GaussianModelState)
from solver.loss_image_state import BatchLossImageState, MultiBatchLossImageState

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _model(P=40, deg=2, seed=0):
    g = ss.make_gaussians(P, sh_degree=deg, seed=seed)
    return GaussianModel.from_activated(g, device="cpu", num_cameras=2)


def _rand_state(m, seed):
    s = GaussianModelState.from_gaussians(m)
    gen = torch.Generator().manual_seed(seed)
    s.packed().copy_(torch.randn(s.length, generator=gen))
    return s


def test_model_activations_roundtrip():
    g = ss.make_gaussians(50, sh_degree=3, seed=1)
    m = GaussianModel.from_activated(g, device="cpu")
    assert np.allclose(m.get_scaling.detach().numpy(), g["scales"], rtol=1e-5)
    assert np.allclose(m.get_opacity.detach().numpy(), g["opacities"], atol=1e-6)
    assert np.allclose(m.get_rotation.detach().numpy(), g["rotations"], atol=1e-6)
    assert np.array_equal(m.get_features.detach().numpy(), g["shs"])
    assert m.get_covariance().shape == (50, 6)


def test_state_is_packed_and_matches_per_tensor_semantics():
    m = _model()
    a, b = _rand_state(m, 1), _rand_state(m, 2)
    P = m._xyz.shape[0]
    assert a.length == P * (3 + 3 * 9 + 3 + 4 + 1) + 2 * 12
    assert a.xyz_grad.data_ptr() == a.flat.data_ptr()          # views into one buffer
    c = a + b * 2.0 - a * 0.5
    for x, y, z in zip(c.tensors(), a.tensors(), b.tensors()):
        assert torch.allclose(x, y + 2.0 * z - 0.5 * y)
    ref = sum(float((x * y).sum()) for x, y in zip(a.tensors(), b.tensors()))
    assert abs(a.dot(b, 1.0) - ref) < 1e-3 * abs(ref) + 1e-4
    assert abs(a.dot(b, 3.0) - 3 * ref) < 1e-3 * abs(ref) + 1e-4
    D = GaussianModelDampMatrix(1.0, 2.0, 3.0, 4.0, 5.0, 6.0, 7.0)
    refD = sum(d * float((x * y).sum()) for d, x, y in zip(D.values(), a.tensors(), b.tensors()))
    assert abs(a.dot(b, D) - refD) < 1e-3 * abs(refD) + 1e-4
    aD = a * D
    for d, x, y in zip(D.values(), aD.tensors(), a.tensors()):
        assert torch.allclose(x, d * y)
    assert torch.allclose((a * (-D)).packed(), -(a * D).packed())
    # 1-d round trip + index_to_desc
    flat = a.as_1d_tensor()
    z = GaussianModelState.from_gaussians(m)
    z.load_1d_tensor(flat)
    assert torch.equal(z.packed(), a.packed())
    assert a.index_to_desc(0) == ("xyz_grad", (0, 0))
    assert a.index_to_desc(3 * P + 4) == ("features_dc_grad", (1, 0, 1))
    # loose-tensor constructor (reference signature) still works
    loose = GaussianModelState(*[t.clone() for t in a.tensors()])
    assert torch.allclose((loose + b).packed(), (a + b).packed())


def test_masks_zero_in_constructor_only():
    m = _model()
    a = _rand_state(m, 3)
    pm = GaussianModelParamGroupMask(mask_xyz=True, mask_opacity=True)
    sm = GaussianModelSplatMask(torch.arange(40) % 2 == 0)
    b = GaussianModelState(*[t.clone() for t in a.tensors()], param_mask=pm, splat_mask=sm)
    assert float(b.xyz_grad.abs().sum()) == 0 and float(b.opacity_grad.abs().sum()) == 0
    assert float(b.scaling_grad[0::2].abs().sum()) == 0 and float(b.scaling_grad[1::2].abs().sum()) > 0
    # arithmetic results are not re-masked (reference :203-250)
    c = b + a
    assert float(c.xyz_grad.abs().sum()) > 0
    for p in (m._xyz, m._scaling):
        p.grad = torch.ones_like(p)
    gstate = GaussianModelState.from_gaussians_grad(m, param_mask=pm)
    assert float(gstate.xyz_grad.abs().sum()) == 0 and float(gstate.scaling_grad.sum()) == m._scaling.numel()


def test_loss_image_state_aliasing_and_dot():
    t = torch.randn(2, 3, 4, 5)
    u = torch.randn(2, 3, 4, 5)
    e = torch.zeros(0)
    a = BatchLossImageState(t, t, e, [(4, 5), (4, 5)], False)       # disable_ssim aliasing
    b = BatchLossImageState(u, u.clone(), e, [(4, 5), (4, 5)], False)
    assert abs(float(a.loss_scalar) - 2 * float((t * t).sum())) < 1e-3
    assert abs(a.dot(b, 1.0) - 2 * float((t * u).sum())) < 1e-3
    assert a.length == 2 * 2 * 3 * 4 * 5
    c = a - b * 0.5
    assert torch.allclose(c.Ll1_per_pixel, t - 0.5 * u)
    s = a + a
    assert s.ssim_loss_per_pixel is s.Ll1_per_pixel                  # aliasing preserved -> one pass downstream
    M = MultiBatchLossImageState([a, a])
    assert abs(M.dot(M, 1.0) - 4 * float((t * t).sum())) < 1e-2
    a.check_invariant()
    # backward through an aliased residual == two backward passes of the reference
    x = torch.randn(2, 3, 4, 5, requires_grad=True)
    r = x * 3.0
    L = BatchLossImageState(r, r, torch.zeros(0, requires_grad=True), [(4, 5)] * 2, False)
    L.backward(b)
    assert torch.allclose(x.grad, 3.0 * (u + u))


def test_conjugate_gradient_spd():
    gen = torch.Generator().manual_seed(0)
    Q = torch.randn(30, 30, generator=gen, dtype=torch.float64)
    A = Q @ Q.T + 30 * torch.eye(30, dtype=torch.float64)
    b = torch.randn(30, generator=gen, dtype=torch.float64)
    x = conjugate_gradient(lambda v: A @ v, torch.add, torch.sub, lambda u, v: float(u @ v), lambda a, v: a * v, b,
                           torch.zeros(30, dtype=torch.float64), tol=1e-12, max_iter=200)
    assert torch.allclose(A @ x, b, atol=1e-8)


def test_cgls_damped_matches_normal_equations(capsys):
    gen = torch.Generator().manual_seed(1)
    A = torch.randn(60, 20, generator=gen, dtype=torch.float64)
    b = torch.randn(60, generator=gen, dtype=torch.float64)
    lam = 0.7
    dot = lambda u, v, d=1.0: float(u @ v) * d
    x = cgls_damped(lambda v: A @ v, lambda u: A.T @ u, dot, lambda a, x, y: a * x + y, b,
                    torch.zeros(20, dtype=torch.float64), damp=lam, max_iter=60, restart_iter=30, tol=1e-14)
    ref = torch.linalg.solve(A.T @ A + lam * torch.eye(20, dtype=torch.float64), A.T @ b)
    assert torch.allclose(x, ref, atol=1e-8)
    x2 = cgls_damped(lambda v: A @ v, lambda u: A.T @ u, dot, lambda a, x, y: a * x + y, b,
                     torch.zeros(20, dtype=torch.float64), damp=lam, max_iter=60, restart_iter=30, tol=1e-14,
                     recompute_residual=False)
    assert torch.allclose(x2, ref, atol=1e-8)
    capsys.readouterr()


def test_generic_mode_linear_solver_functions_on_toy_loss():
    """Generic mode = the reference's composition (fwAD J v, autograd J^T u): adjoint identity and
    normal_matvec == J^T (J v) on a toy differentiable loss (no rasterizer, CPU)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from toy_loss import make_toy
    from solver.solver_functions import LinearSolverFunctions
    m, cams, loss_func = make_toy()
    F = LinearSolverFunctions(loss_func, m, cams, batch_size=2)
    assert not F.cached
    F.evaluate_loss()
    v = _rand_state(m, 5)
    Jv = F.matvec(v)
    u = Jv * 0.3 + F.loss * 1.0
    JTu = F.matvec_T(u)
    lhs, rhs = F.dot(u, Jv), F.dot(JTu, v)
    assert abs(lhs - rhs) <= 1e-4 * max(abs(lhs), abs(rhs))
    n1 = F.normal_matvec(v)
    n2 = F.matvec_T(F.matvec(v))
    assert torch.allclose(n1.packed(), n2.packed(), rtol=1e-5, atol=1e-6)
    n3 = F.normal_matvec(v, damp=0.5)
    assert torch.allclose(n3.packed(), n2.packed() + 0.5 * v.packed(), rtol=1e-5, atol=1e-6)
    # LM step: CGLS on min ||J s + r||^2 + lam ||s||^2 decreases the linearised objective
    b = -1 * F.loss
    s = cgls_damped(F.matvec, F.matvec_T, F.dot, F.saxpy, b, F.get_initial_solution(), damp=1e-3, max_iter=4,
                    restart_iter=4)
    lin = F.saxpy(-1.0, F.matvec(s), b)
    assert F.dot(lin, lin) < F.dot(b, b)


@pytest.mark.timeout(300)
def test_view_sharding_over_gloo_world2():
    """world_size 2, gloo: views sharded over ranks, one allreduce per matvec_T / normal_matvec, scalar allreduce
    for residual-space dots -- results equal the single-process ones."""
    script = os.path.join(ROOT, "tests", "dist_worker.py")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29517")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29517", script],
                         capture_output=True, text=True, env=env, timeout=280)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "DIST_OK" in out.stdout


def test_ply_checkpoint_roundtrip_and_reference_layout(tmp_path):
    """save_ply / load_ply (reference scene/gaussian_model.py:315-404): header and column order as plyfile writes them
    for the reference (binary little endian, float properties x y z nx ny nz f_dc_* f_rest_* opacity scale_* rot_*,
    SH channel-major), exact round trip, and a big-endian / double file of the same content loads identically."""
    import scene_synth as ss
    from scene.gaussian_model import GaussianModel
    g = ss.make_gaussians(257, sh_degree=2, seed=21)
    m = GaussianModel.from_activated(g, device="cpu")
    path = str(tmp_path / "point_cloud" / "iteration_7" / "point_cloud.ply")
    m.save_ply(path)
    raw = open(path, "rb").read()
    head, body = raw.split(b"end_header\n", 1)
    lines = head.decode().splitlines()
    assert lines[:3] == ["ply", "format binary_little_endian 1.0", "element vertex 257"]
    names = [ln.split()[2] for ln in lines[3:]]
    assert all(ln.startswith("property float ") for ln in lines[3:])
    assert names == (["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"] + [f"f_rest_{i}" for i in range(24)]
                     + ["opacity", "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3"])
    cols = np.frombuffer(body, dtype="<f4").reshape(257, len(names))
    # channel-major SH: f_rest_0..7 are the 8 red coefficients (features_rest[:, :, 0])
    assert np.array_equal(cols[:, 9:17], m._features_rest.detach().numpy()[:, :, 0])
    assert np.array_equal(cols[:, 3:6], np.zeros((257, 3), np.float32))
    m2 = GaussianModel(2)
    m2.load_ply(path, device="cpu")
    for n in ("_xyz", "_features_dc", "_features_rest", "_scaling", "_rotation", "_opacity"):
        assert torch.equal(getattr(m2, n).detach(), getattr(m, n).detach()), n
    assert m2.active_sh_degree == 2
    # the same content as big-endian doubles with a comment line and a second element after the vertices
    be = str(tmp_path / "be.ply")
    with open(be, "wb") as f:
        f.write(("ply\nformat binary_big_endian 1.0\ncomment made by a test\nelement vertex 257\n"
                 + "".join(f"property double {n}\n" for n in names)
                 + "element face 0\nproperty list uchar int vertex_indices\nend_header\n").encode())
        f.write(cols.astype(">f8").tobytes())
    m3 = GaussianModel(2)
    m3.load_ply(be, device="cpu", requires_grad=False)
    assert torch.equal(m3._features_rest, m._features_rest.detach()) and torch.equal(m3._rotation, m._rotation.detach())


def test_render_chunks_merge_whole_batches_of_one_size():
    """loss_value() renders chunks that are unions of WHOLE batches (the alpha-mask rule of the reference is per
    batch, solver/batch_training_loss.py:61-67) of ONE image size (a padded batch's residual depends on the
    padding), at most RENDER_CHUNK views each -- host logic only."""
    from types import SimpleNamespace
    from solver.solver_functions import LinearSolverFunctions

    def chunks(sizes, batch_size):
        cams = [SimpleNamespace(image_height=h, image_width=w) for h, w in sizes]
        F = LinearSolverFunctions(lambda **k: None, SimpleNamespace(), cams, batch_size=batch_size, cached=False,
                                  shard_views=False)
        return list(F._render_chunks())

    a, b = (96, 128), (56, 80)
    assert chunks([a] * 5, 1) == [(0, (1, 2, 3, 4, 5))]
    assert chunks([a] * 40, 3)[0] == (0, (3, 6, 9, 12, 15)) and chunks([a] * 40, 3)[-1] == (30, (33, 36, 39, 40))
    assert chunks([a, a, b, a, a], 1) == [(0, (1, 2)), (2, (3,)), (3, (4, 5))]
    assert chunks([a, b, a, a], 2) == [(0, (2,)), (2, (4,))]          # a mixed-size batch stays on its own
    assert chunks([a] * 50, 20) == [(0, (20,)), (20, (40,)), (40, (50,))]   # batches larger than the chunk
    assert chunks([], 4) == []
    covered = [v for s, bounds in chunks([a] * 7 + [b] * 9 + [a] * 3, 2) for v in range(s, bounds[-1])]
    assert covered == list(range(19))


def test_balanced_view_assignment():
    """Greedy (LPT) view -> rank assignment on per-view instance counts: a partition, at most ceil(V / N)
    views per rank, deterministic, and better balanced than round robin on a skewed rig."""
    from solver.solver_functions import balanced_view_assignment
    costs = [100, 1, 1, 1, 90, 1, 1, 1, 80, 1, 1, 1, 70, 1, 1, 1]       # round robin would put all heavy views on rank 0
    parts = balanced_view_assignment(costs, 4)
    assert sorted(i for p in parts for i in p) == list(range(16))
    assert all(len(p) == 4 and p == sorted(p) for p in parts)
    loads = [sum(costs[i] for i in p) for p in parts]
    rr = [sum(costs[r::4]) for r in range(4)]
    assert max(loads) - min(loads) < max(rr) - min(rr) and max(loads) <= 103
    assert balanced_view_assignment(costs, 4) == parts
    # ragged: 5 views over 4 ranks, one rank gets two, nobody more; more ranks than views leaves ranks empty
    p5 = balanced_view_assignment([5, 4, 3, 2, 1], 4)
    assert sorted(len(p) for p in p5) == [1, 1, 1, 2] and sorted(i for p in p5 for i in p) == [0, 1, 2, 3, 4]
    p2 = balanced_view_assignment([7, 7], 4)
    assert sorted(len(p) for p in p2) == [0, 0, 1, 1]
