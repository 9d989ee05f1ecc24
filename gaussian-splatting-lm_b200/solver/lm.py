"""The Levenberg-Marquardt inner loop of the reference driver (train_jvp.py:237-288) on top of the fused
Gauss-Newton product: linearise, solve the damped normal equations with CG, back-track on validation views.

    reference                                                     here
    cgls_damped(matvec, matvec_T, ..., b=-r, x0=0, damp=D)        cg_normal_equations(F, damp=D)
      min ||J s + r||^2 + s^T D s, 2 J v + 1-2 J^T u sweeps/iter    (J^T J + D) s = J^T(-r), ONE fused J^T J v/iter
    line search: alpha = 2, 1, ..., 1/16 on 50 validation views     line_search(...) -- same schedule

CGLS and CG on the normal equations generate the same iterates in exact arithmetic (Bjorck 1996, 7.4); the
reference's CGLS remains available unchanged in solver.conjugate_gradient.cgls_damped.

Device-resident: every scalar of the CG recurrence (alpha, beta, <r, r>, p^T A p) is a 0-dim CUDA tensor,
convergence is handled by masking the step instead of branching -- so a CG solve issues NO host synchronisation
(the reference's dot() does one .item() per inner product, solver/solver_functions.py:125-130) and the iteration is
captured in a CUDA graph (cg_normal_equations(graph=True)).
"""
import torch

from solver.gaussian_model_state import GaussianModelDampMatrix, GaussianModelState


def _damp_mul(v, damp):
    """D v as a packed tensor (damp: None, float or GaussianModelDampMatrix)."""
    if damp is None:
        return None
    return (v * damp).packed()


def cg_normal_equations(F, damp=None, max_iter=2, tol=1e-10, rhs=None, x0=None, callback=None, graph=False):
    """Solve (J^T J + D) s = J^T(-r) at F's current linearisation with `max_iter` CG iterations.

    F: LinearSolverFunctions in cached mode with evaluate_loss() already called (the linearisation).
    Returns (s, info) with info = {"residual_norms": 1-D device tensor [max_iter + 1]} -- nothing is read back.
    A converged solve (||r|| < tol ||r0||) keeps iterating with zero steps, so the launch sequence is static.

    p^T A p is taken as dot(p, A p) on the (masked, damped) product: a fixed-order reduction, so the CG step sizes
    are bit-reproducible run to run.  (The fused tile kernel can also return v^T J^T W J v through out_vJJv, but it
    accumulates that scalar with one atomicAdd per tile -- order-dependent in the last bits; it stays available as a
    diagnostic, F.normal_matvec(return_quad=True).)

    graph=True: iteration 1 runs eagerly (it doubles as the warm-up), iteration 2 is captured into a CUDA graph
    (every launch of the iteration: activation JVP, tangent preprocess / fused tile kernel / adjoint per chunk of
    views, the NCCL allreduce, activation VJP, the CG vector updates) and iterations 2..max_iter are replays: one
    cudaGraphLaunch each instead of ~30 launches through Python and ctypes.  Same arithmetic, same bits."""
    if rhs is None:
        rhs = F.matvec_T(-1 * F.loss)
    b = rhs.packed()
    x = torch.zeros_like(b) if x0 is None else x0.packed().clone()
    like = rhs
    wrap = lambda flat: GaussianModelState.from_flat(flat, like)
    if x0 is None:
        r = b.clone()
    else:
        r = b - F.normal_matvec(wrap(x), damp=damp).packed()
    p = r.clone()
    rs = torch.dot(r, r)
    rs0 = rs.clone()
    norms = torch.zeros(max_iter + 1, dtype=torch.float32, device=b.device)
    norms[0] = rs.sqrt()
    tiny = torch.finfo(torch.float32).tiny
    pv = wrap(p)            # views of the static buffer p

    def iteration():
        """One CG iteration, in place on the static buffers x, r, p, rs."""
        Apf = F.normal_matvec(pv, damp=damp).packed()                    # one fused sweep over all views
        pAp = torch.dot(p, Apf)
        live = (rs > (tol * tol) * rs0) & (pAp > tiny)
        alpha = torch.where(live, rs / pAp.clamp_min(tiny), torch.zeros_like(rs))
        # one pass per vector update (the scalar stays on the device and is broadcast inside the kernel): 13 sweeps
        # over the 59 P floats per iteration instead of 20 with temporaries
        x.addcmul_(alpha, p)                          # x += alpha p
        r.addcmul_(alpha, Apf, value=-1.0)            # r -= alpha A p
        rs_new = torch.dot(r, r)
        beta = torch.where(live, rs_new / rs.clamp_min(tiny), torch.zeros_like(rs))
        torch.addcmul(r, beta, p, out=p)              # p = r + beta p
        rs.copy_(rs_new)

    g = None
    for k in range(max_iter):
        # (multi-GPU products end in a symmetric-memory barrier + multimem kernel or an NCCL call on the caller's
        # stream; their capture is not exercised on hardware here, so the graph is a single-GPU feature for now)
        if graph and k == 1 and max_iter > 2 and callback is None and getattr(F, "dist", None) is None:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                iteration()
        if g is not None:
            g.replay()
        else:
            iteration()
        norms[k + 1] = rs.sqrt()
        if callback is not None:
            callback(wrap(x), wrap(r), k + 1)
    info = {"residual_norms": norms, "graphed": g is not None}
    if g is not None:
        # the graph's private pool owns the buffers of the captured iteration; results live in x / r / p (ours)
        torch.cuda.current_stream().synchronize()
        del g
    return wrap(x), info


def line_search(gaussians, s, val_loss, alpha0=2.0, halvings=6):
    """train_jvp.py:262-279: try alpha0, alpha0/2, ... on the validation loss, keep the best, leave the model at
    theta + best_alpha * s.  val_loss() -> scalar loss at the model's current parameters.
    Returns (best_alpha, best_loss, [(alpha, loss), ...])."""
    alpha, best_alpha, best_loss = alpha0, alpha0, float("inf")
    trace = []
    with torch.no_grad():
        gaussians.update_step(alpha * s)
        for _ in range(halvings):
            v = float(val_loss())
            trace.append((alpha, v))
            if v < best_loss:
                best_loss, best_alpha = v, alpha
            new_alpha = alpha * 0.5
            gaussians.update_step((new_alpha - alpha) * s)
            alpha = new_alpha
        gaussians.update_step((best_alpha - alpha) * s)
    return best_alpha, best_loss, trace


def lm_step(loss_func, gaussians, train_cams, val_cams=None, damp=None, param_mask=None, splat_mask=None,
            cg_iters=2, batch_size=20, alpha0=2.0, halvings=6, verbose=False):
    """One LM step as train_jvp.py:237-288 does it: linearise on train_cams, solve for the update, back-track on
    val_cams (default: the training views).  Returns a dict with the losses, the step and the CG residual norms."""
    from solver.solver_functions import LinearSolverFunctions
    if damp is None:
        damp = GaussianModelDampMatrix(5e2, 5e-2, 5e-2, 5e-2, 5e-2, 5e-2, 1e1)        # train_jvp.py:229-235
    F = LinearSolverFunctions(loss_func, gaussians, train_cams, batch_size=batch_size, param_mask=param_mask,
                              splat_mask=splat_mask)
    with torch.no_grad():
        F.evaluate_loss()
        start_loss = float(F.loss_scalar)
        s, info = cg_normal_equations(F, damp=damp, max_iter=cg_iters)
    del F
    V = LinearSolverFunctions(loss_func, gaussians, val_cams if val_cams is not None else train_cams,
                              batch_size=batch_size)

    def val_loss():
        return V.loss_value()          # render-only passes: no linearisation cache for points that are only compared

    best_alpha, best_loss, trace = line_search(gaussians, s, val_loss, alpha0, halvings)
    final = float(val_loss())
    if verbose:
        print(f"LM step: train loss {start_loss:.6f}; line search {trace}; alpha* = {best_alpha}; val loss {final:.6f}")
    return {"start_loss": start_loss, "val_loss": final, "best_alpha": best_alpha, "trace": trace, "step": s,
            "cg_residual_norms": info["residual_norms"]}
