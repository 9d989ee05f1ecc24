"""Matrix-free Krylov solvers with the call signatures of solver/conjugate_gradient.py of the reference:
conjugate_gradient (:3-49) for A x = b with a symmetric operator, and cgls_damped (:51-127), restarted CGLS for
min ||A x - b||^2 + x^T D x.  Backend-agnostic: every vector operation arrives as a callable; scalars are Python
floats (or anything supporting float arithmetic)."""
import math


def conjugate_gradient(matvec, vec_add, vec_sub, dot, scalar_mul, b, x0, tol=1e-10, atol=0.0, max_iter=1000,
                       callback=None, verbose=False):
    """Standard CG. matvec(x) -> A x. Stops when ||r|| < max(tol * ||r0||, atol). Returns x."""
    x = x0
    r = vec_sub(b, matvec(x))
    p = r
    rs = dot(r, r)
    norm0 = rs ** 0.5
    if norm0 < atol:
        return x
    for k in range(max_iter):
        Ap = matvec(p)
        alpha = rs / dot(p, Ap)
        x = vec_add(x, scalar_mul(alpha, p))
        r = vec_sub(r, scalar_mul(alpha, Ap))
        rs_next = dot(r, r)
        norm_r = rs_next ** 0.5
        if verbose:
            print(f"[Iter {k + 1}] Residual norm: {norm_r:.2e}")
        if callback:
            callback(x, r, k + 1)
        if norm_r < max(tol * norm0, atol):
            break
        p = vec_add(r, scalar_mul(rs_next / rs, p))
        rs = rs_next
    return x


def cgls_damped(matvec, matvec_T, dot, saxpy, b, x0, damp=0.0, tol=1e-10, atol=0.0, max_iter=1000, restart_iter=5,
                callback=None, verbose=False, recompute_residual=True):
    """Restarted damped CGLS. matvec(x) -> A x, matvec_T(y) -> A^T y, saxpy(a, x, y) -> a x + y (a may be a
    damping object), dot(x, y[, damp]).

    Behaviour follows the reference step by step, including its safeguards: stop when delta < 1e-20, when the
    objective ||b - A x||^2 + <x, x>_D increases, or when gamma < max(tol * sqrt(gamma_prev), atol).
    recompute_residual=True evaluates that objective with a fresh A x every iteration as the reference does
    (:103-105, one extra J v sweep); False uses the recurrence residual r, which equals b - A x in exact
    arithmetic and saves the sweep."""
    x = x0
    done = 0
    last_res = math.inf
    stop = False
    while done < max_iter and not stop:
        if verbose:
            print(f"Restarting CG iteration {done + 1}...")
        r = saxpy(-1.0, matvec(x), b)
        s = saxpy(-damp, x, matvec_T(r))
        p = s
        gamma = dot(s, s)
        for _ in range(restart_iter):
            q = matvec(p)
            delta = dot(q, q) + dot(p, p, damp)
            if delta < 1e-20:
                print("Early termination: delta is too small.")
                stop = True
                break
            alpha = gamma / delta
            x = saxpy(alpha, p, x)
            r = saxpy(-alpha, q, r)
            s = saxpy(-damp, x, matvec_T(r))
            gamma_prev, gamma = gamma, dot(s, s)
            p = saxpy(gamma / gamma_prev, p, s)
            cur_r = saxpy(-1.0, matvec(x), b) if recompute_residual else r
            res = dot(cur_r, cur_r) + dot(x, x, damp)
            if verbose or recompute_residual:
                print(f"[Iter {done + 1}] res: {res:.2e}")
            if callback:
                callback(x, cur_r, done + 1)
            if res > last_res:
                print("Warning: Residual norm increased!")
                stop = True
                break
            last_res = res
            if gamma < max(tol * (gamma_prev ** 0.5), atol):
                if verbose:
                    print(f"Convergence achieved at iteration {done + 1}.")
                stop = True
                break
            done += 1
            if done >= max_iter:
                stop = True
                break
    return x
