"""torchrun worker for tests/test_gpu_multi.py: the CACHED CUDA path (linearisation cache, fused J^T J v, NCCL
collectives) with the views sharded over WORLD_SIZE GPUs.  Every rank writes its results to <out>/rank<r>.pt; the
single-process test compares them with its own unsharded run."""
import functools
import os
import sys
from types import SimpleNamespace

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "gaussian-splatting-lm_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import scene_synth as ss                                        # noqa: E402


def build(dev, P=100_000, V=16, W=800, H=800, deg=3, ssim=False):
    """Config B of BASELINE.json (100k Gaussians, 16 views at 800x800) with random targets and alpha masks.
    EVERY view carries a mask: the reference zeroes a mask-less view when another view of its BATCH has one
    (solver/batch_training_loss.py:61-67), which would make the result depend on how views are grouped."""
    from scene.gaussian_model import GaussianModel
    from solver.batch_training_loss import batch_training_loss
    g = ss.make_gaussians(P, sh_degree=deg, seed=1001)
    model = GaussianModel.from_activated(g, device=dev, num_cameras=V, requires_grad=False)
    gen = torch.Generator(device="cpu").manual_seed(5)
    cams = []
    for i, c in enumerate(ss.hemisphere_rig(V, W, H)):
        gt = torch.rand(3, H, W, generator=gen).to(dev)
        am = (torch.rand(1, H, W, generator=gen) > 0.1).float().to(dev)
        cams.append(ss.torch_camera(c, dev, original_image=gt, alpha_mask=am, uid=i))
    pipe = SimpleNamespace(debug=False, antialiasing=False, compute_cov3D_python=False, convert_SHs_python=False)
    loss_func = functools.partial(batch_training_loss, iteration=1, opt=SimpleNamespace(lambda_dssim=0.2), pipe=pipe,
                                  bg=torch.zeros(3, device=dev), train_test_exp=False,
                                  depth_l1_weight=lambda it: 0.0, disable_ssim=not ssim)
    return model, cams, loss_func


def products(F, model):
    """What the LM step takes from LinearSolverFunctions, as CPU tensors."""
    from solver.gaussian_model_state import GaussianModelDampMatrix, GaussianModelState
    from solver.lm import cg_normal_equations
    F.evaluate_loss()
    v = GaussianModelState.from_gaussians(model)
    v.packed().copy_(torch.randn(v.length, generator=torch.Generator().manual_seed(9)).to(v.packed().device))
    v.exposure_grad.zero_()
    D = GaussianModelDampMatrix(5e2, 5e-2, 5e-2, 5e-2, 5e-2, 5e-2, 1e1)
    out = {"loss": torch.as_tensor(float(F.loss_scalar))}
    Av, quad = F.normal_matvec(v, damp=D, return_quad=True)
    out["normal_matvec"] = Av.packed().cpu()
    out["quad"] = quad.cpu()
    out["matvec_T"] = F.matvec_T(-1 * F.loss).packed().cpu()
    Jv = F.matvec(v)
    out["JvJv"] = torch.as_tensor(F.dot(Jv, Jv))
    s, info = cg_normal_equations(F, damp=D, max_iter=3)
    out["cg_step"] = s.packed().cpu()
    out["cg_norms"] = info["residual_norms"].cpu()
    out["loss_value"] = torch.as_tensor(float(F.loss_value()))
    return out


def one_view_products(dev, model, cams, loss_func):
    """J^T J v twice in a row and J^T u over ONE view: on two ranks, rank 1 owns no view and must still enter every
    collective (its accumulator contributes zeros); a smaller-P product would reuse the same symmetric buffer."""
    from solver.gaussian_model_state import GaussianModelState
    from solver.solver_functions import LinearSolverFunctions
    F1 = LinearSolverFunctions(loss_func, model, cams[:1], batch_size=4)
    F1.evaluate_loss()
    v = GaussianModelState.from_gaussians(model)
    v.packed().copy_(torch.randn(v.length, generator=torch.Generator().manual_seed(13)).to(v.packed().device))
    v.exposure_grad.zero_()
    first = F1.normal_matvec(v).packed().clone()
    second = F1.normal_matvec(v).packed().clone()        # back to back on the shared accumulator
    return {"one_view_n_local": torch.as_tensor(F1.B), "one_view_jtjv": first.cpu(), "one_view_jtjv_again": second.cpu(),
            "one_view_jtu": F1.matvec_T(-1 * F1.loss).packed().cpu()}


def lm_outcome(dev, ssim):
    """One whole LM step (train_jvp.py:237-288: linearise, 2 CG iterations, 6-point line search) on a fresh copy of the
    scene: what a sharded run must reproduce."""
    from solver.gaussian_model_state import GaussianModelParamGroupMask
    from solver.lm import lm_step
    model, cams, loss_func = build(dev, ssim=ssim)
    with torch.no_grad():
        gen = torch.Generator().manual_seed(4)
        model._features_dc.data += 0.2 * torch.randn(model._features_dc.shape, generator=gen).to(dev)
        model._opacity.data += 0.3 * torch.randn(model._opacity.shape, generator=gen).to(dev)
    r = lm_step(loss_func, model, cams[:8], val_cams=cams[8:], param_mask=GaussianModelParamGroupMask(mask_xyz=True),
                cg_iters=2, batch_size=4)
    return {"lm_start": torch.as_tensor(r["start_loss"]), "lm_val": torch.as_tensor(r["val_loss"]),
            "lm_alpha": torch.as_tensor(r["best_alpha"]), "lm_step_norm": torch.linalg.vector_norm(r["step"].packed()).cpu(),
            "lm_trace": torch.tensor([v for _, v in r["trace"]])}


def main():
    out_dir = sys.argv[1]
    ssim = len(sys.argv) > 2 and sys.argv[2] == "ssim"
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    import datetime
    dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=120))
    rank, world = dist.get_rank(), dist.get_world_size()
    from solver.solver_functions import LinearSolverFunctions
    model, cams, loss_func = build(dev, ssim=ssim)
    F = LinearSolverFunctions(loss_func, model, cams, batch_size=4)
    assert F.cached and F.dist is not None and F.B == len(cams) // world
    res = products(F, model)
    # which reduction ended the products: the multimem kernel over the symmetric accumulator (GSLM_PEER_REDUCE=1, or
    # auto from 3 ranks up) or NCCL
    res["peer_used"] = torch.as_tensor(getattr(F, "_peer", None) is not None)
    if res["peer_used"]:
        from solver import solver_functions as sf
        assert F._flat_acc.data_ptr() == sf._PEER_CACHE[dev.index].t.data_ptr()
    res.update(one_view_products(dev, model, cams, loss_func))
    res.update(lm_outcome(dev, ssim))
    res["view_ids"] = torch.tensor(F.view_ids)
    res["view_costs"] = torch.tensor(getattr(F, "view_costs", []))
    torch.save(res, os.path.join(out_dir, f"rank{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
