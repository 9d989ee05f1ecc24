"""torchrun worker for tests/test_host_cpu.py::test_view_sharding_over_gloo_world2 (gloo, CPU)."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "gaussian-splatting-lm_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

from solver.gaussian_model_state import GaussianModelState   # noqa: E402
from solver.solver_functions import LinearSolverFunctions    # noqa: E402
from toy_loss import make_toy                                 # noqa: E402


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    m, cams, loss_func = make_toy()
    v = GaussianModelState.from_gaussians(m)
    v.packed().copy_(torch.randn(v.length, generator=torch.Generator().manual_seed(9)))
    sharded = LinearSolverFunctions(loss_func, m, cams, batch_size=1)
    assert sharded.B == len(cams) // world and sharded.dist is not None
    single = LinearSolverFunctions(loss_func, m, cams, batch_size=2, shard_views=False)
    for F in (sharded, single):
        F.evaluate_loss()
    assert abs(float(sharded.loss_scalar) - float(single.loss_scalar)) < 1e-5 * float(single.loss_scalar)
    a, b = sharded.normal_matvec(v), single.normal_matvec(v)
    assert torch.allclose(a.packed(), b.packed(), rtol=1e-5, atol=1e-6), (a.packed() - b.packed()).abs().max()
    Jv_s, Jv_1 = sharded.matvec(v), single.matvec(v)
    d_s, d_1 = sharded.dot(Jv_s, Jv_s), single.dot(Jv_1, Jv_1)
    assert abs(d_s - d_1) < 1e-5 * abs(d_1)
    # every rank holds the same replicated result
    chk = a.packed().clone()
    dist.broadcast(chk, 0)
    assert torch.equal(chk, a.packed())
    # host staging helpers: 1/N slices + all_gather reproduce the vector; slices of the result tile the host copy
    host = torch.arange(v.length, dtype=torch.float32) * 0.5
    w = GaussianModelState.from_gaussians(m)
    moved = sharded.state_from_host(host, w)
    assert torch.equal(w.packed(), host) and moved <= 4 * ((v.length + world - 1) // world)
    back = torch.full((v.length,), -1.0)
    sharded.state_to_host(w, back)
    per = (v.length + world - 1) // world
    lo, hi = rank * per, min((rank + 1) * per, v.length)
    assert torch.equal(back[lo:hi], host[lo:hi]) and (back[:lo] == -1).all() and (back[hi:] == -1).all()
    # a rank that owns NO view (more ranks than views) must still enter every collective -- loss scalars,
    # residual-space dots and the parameter-space allreduce -- instead of deadlocking the others
    lonely = LinearSolverFunctions(loss_func, m, cams[:1], batch_size=1)
    assert lonely.B == (1 if rank == 0 else 0)
    one = LinearSolverFunctions(loss_func, m, cams[:1], batch_size=1, shard_views=False)
    for F in (lonely, one):
        F.evaluate_loss()
    assert abs(float(lonely.loss_scalar) - float(one.loss_scalar)) < 1e-5 * float(one.loss_scalar)
    assert abs(float(lonely.Ll1_scalar) - float(one.Ll1_scalar)) < 1e-5 * float(one.Ll1_scalar)
    a1, b1 = lonely.normal_matvec(v), one.normal_matvec(v)
    assert torch.allclose(a1.packed(), b1.packed(), rtol=1e-5, atol=1e-6)
    J1, J2 = lonely.matvec(v), one.matvec(v)
    assert abs(lonely.dot(J1, J1) - one.dot(J2, J2)) < 1e-5 * abs(one.dot(J2, J2))
    dist.barrier()
    if rank == 0:
        print("DIST_OK")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
