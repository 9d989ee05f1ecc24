#!/usr/bin/env python
"""LM inner-loop benchmark (BASELINE config B "LM/CG step on 1xB200"): whole Levenberg-Marquardt
steps as train_jvp.py:237-288 runs them -- linearise the training views, right-hand side J^T(-r), `cg_iters` CG
iterations on the damped normal equations (one fused J^T J v each), back-tracking line search alpha = 2 ... 1/16 on
the validation views -- through solver.lm on synthetic data.  Targets are renders of the unperturbed scene, the model
starts from a perturbed copy, so the loss has something to descend.

    python tools/bench_lm.py [--config B|C|D] [--views V] [--val-views N] [--cg-iters 2] [--steps 3] [--ssim]
Prints one JSON line: ms per LM step and its split (wall clock around device synchronisations, per phase)."""
import argparse
import functools
import json
import os
import sys
import time
from types import SimpleNamespace

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gaussian-splatting-lm_b200"))

import torch  # noqa: E402

import scene_synth as ss  # noqa: E402
from gaussian_renderer import render  # noqa: E402
from scene.gaussian_model import GaussianModel  # noqa: E402
from solver.batch_training_loss import batch_training_loss  # noqa: E402
from solver.gaussian_model_state import GaussianModelDampMatrix, GaussianModelParamGroupMask  # noqa: E402
from solver.lm import cg_normal_equations, line_search  # noqa: E402
from solver.solver_functions import LinearSolverFunctions  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="B", choices=["A", "B", "C", "D"])
    ap.add_argument("--views", type=int, default=0)
    ap.add_argument("--points", type=int, default=0)
    ap.add_argument("--val-views", type=int, default=0, help="validation views of the line search (0 = all views)")
    ap.add_argument("--cg-iters", type=int, default=2)        # train_jvp.py:250 max_iter=2
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--ssim", action="store_true", help="full L1 + SSIM residual instead of disable_ssim=True")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    g, cams_np = ss.make_config(a.config, sh_degree=3, views=a.views or None, P=a.points or None)
    P, V = g["means3D"].shape[0], len(cams_np)
    W, H = cams_np[0].image_width, cams_np[0].image_height
    model = GaussianModel.from_activated(g, device=dev, num_cameras=V, requires_grad=False)
    pipe = SimpleNamespace(debug=False, antialiasing=False, compute_cov3D_python=False, convert_SHs_python=False)
    bg = torch.zeros(3, device=dev)
    cams = [ss.torch_camera(c, dev, uid=i) for i, c in enumerate(cams_np)]
    gen = torch.Generator().manual_seed(11)
    with torch.no_grad():
        for c in cams:
            c.original_image = render(c, model, pipe, bg)["render"].clamp(0, 1)
        model._features_dc.data += 0.2 * torch.randn(model._features_dc.shape, generator=gen).to(dev)
        model._opacity.data += 0.3 * torch.randn(model._opacity.shape, generator=gen).to(dev)
        model._scaling.data += 0.05 * torch.randn(model._scaling.shape, generator=gen).to(dev)
    loss_func = functools.partial(batch_training_loss, iteration=1, opt=SimpleNamespace(lambda_dssim=0.2), pipe=pipe,
                                  bg=bg, train_test_exp=False, depth_l1_weight=lambda it: 0.0,
                                  disable_ssim=not a.ssim)
    damp = GaussianModelDampMatrix(5e2, 5e-2, 5e-2, 5e-2, 5e-2, 5e-2, 1e1)          # train_jvp.py:229-235
    pm = GaussianModelParamGroupMask(mask_xyz=True)                                 # train_jvp.py:221-227
    val = cams[:a.val_views] if a.val_views else cams
    bs = 16 if a.ssim else 1

    def tick():
        torch.cuda.synchronize()
        return time.perf_counter()

    rows = []
    for step in range(a.steps + 1):                       # step 0 = warm-up (allocator, sort scratch)
        t0 = tick()
        with torch.no_grad():
            F = LinearSolverFunctions(loss_func, model, cams, batch_size=bs, param_mask=pm)
            F.evaluate_loss()
            start = float(F.loss_scalar)
            t1 = tick()
            rhs = F.matvec_T(-1 * F.loss)
            t2 = tick()
            s, info = cg_normal_equations(F, damp=damp, max_iter=a.cg_iters, rhs=rhs)
            t3 = tick()
            del F
            Vf = LinearSolverFunctions(loss_func, model, val, batch_size=bs)

            def val_loss():
                return Vf.loss_value()

            best_alpha, best_loss, trace = line_search(model, s, val_loss)
            t4 = tick()
            del Vf
        if step:
            rows.append({"linearize_ms": (t1 - t0) * 1e3, "rhs_ms": (t2 - t1) * 1e3, "cg_ms": (t3 - t2) * 1e3,
                         "line_search_ms": (t4 - t3) * 1e3, "total_ms": (t4 - t0) * 1e3, "start_loss": start,
                         "val_loss": best_loss, "alpha": best_alpha})
    mean = lambda k: sum(r[k] for r in rows) / len(rows)
    print(json.dumps({
        "workload": f"config {a.config}: {P} Gaussians, {V} training views at {W}x{H}, {len(val)} validation views, "
                    f"{a.cg_iters} CG iterations, line search over 6 step sizes, "
                    f"{'L1 + SSIM' if a.ssim else 'disable_ssim'} residual, xyz masked",
        "lm_steps": len(rows), "ms_per_lm_step": mean("total_ms"),
        "split_ms": {k: round(mean(k), 2) for k in ("linearize_ms", "rhs_ms", "cg_ms", "line_search_ms")},
        "ms_per_cg_iteration": mean("cg_ms") / max(a.cg_iters, 1),
        "loss_trace": [(round(r["start_loss"], 6), round(r["val_loss"], 6), r["alpha"]) for r in rows]}))


if __name__ == "__main__":
    main()
