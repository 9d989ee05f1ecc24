"""Builds tests/golden/loss_fixture.npz by RUNNING the reference's own Python loss functions (CPU torch, fp32).

Run in the build container only (imports from /root/reference, which does not exist on the GPU box):
    python tests/golden/make_loss_golden.py
Reference functions executed, unmodified:
  utils/loss_utils.py:91-117              ssim_per_pixel / _ssim_per_pixel
  solver/batch_training_loss.py:10-31     compute_batch_loss_block
`solver/batch_training_loss.py` also imports gaussian_renderer.batch_render (needs the CUDA rasterizer extension, not
importable here); that one unrelated import is satisfied with an empty stand-in module.  Derivatives are what PyTorch
derives from the reference functions: forward AD for the tangent, autograd for the pull-back.
Cases: (a) two 37x45 images, no alpha mask; (b) 40x33 images with an alpha mask, the second image zero-padded from
29x33 (the batch tensor layout of batch_training_loss.py:54-66); (c) disable_ssim=True.
"""
import importlib
import math
import os
import sys
import types

import numpy as np
import torch
import torch.autograd.forward_ad as fwAD

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "loss_fixture.npz")


def reference_functions():
    sys.path.insert(0, REF)
    stub_pkg = types.ModuleType("gaussian_renderer")
    stub_pkg.__path__ = []
    stub = types.ModuleType("gaussian_renderer.batch_render")
    stub.batch_render = None
    sys.modules["gaussian_renderer"] = stub_pkg
    sys.modules["gaussian_renderer.batch_render"] = stub
    lu = importlib.import_module("utils.loss_utils")
    btl = importlib.import_module("solver.batch_training_loss")
    return lu.ssim_per_pixel, btl.compute_batch_loss_block


def case(block, gen, B, H, W, sizes, with_mask, lam, disable_ssim):
    img = torch.rand((B, 3, H, W), generator=gen)
    gt = torch.rand((B, 3, H, W), generator=gen)
    # a patch where the render already equals the target, and one where it is far off
    img[0, :, 5:12, 7:15] = gt[0, :, 5:12, 7:15]
    img[-1, :, 20:26, 3:9] = 1.0 - gt[-1, :, 20:26, 3:9]
    mask = None
    if with_mask:
        mask = (torch.rand((B, 1, H, W), generator=gen) > 0.2).float()
    for i, (h, w) in enumerate(sizes):           # zero padding outside each view's own size
        for t in (img, gt) + ((mask,) if mask is not None else ()):
            t[i, :, h:, :] = 0
            t[i, :, :, w:] = 0
    alphas = torch.tensor([math.sqrt((1 - lam) / (3 * h * w)) for h, w in sizes]).view(B, 1, 1, 1)
    betas = torch.tensor([math.sqrt(lam / (3 * h * w)) for h, w in sizes]).view(B, 1, 1, 1)
    t_img = torch.randn((B, 3, H, W), generator=gen)
    u1 = torch.randn((B, 3, H, W), generator=gen)
    u2 = torch.randn((B, 3, H, W), generator=gen)
    r1, r2 = block(img, mask, gt, alphas, betas, False, disable_ssim)
    with fwAD.dual_level():
        d1, d2 = block(fwAD.make_dual(img, t_img), mask, gt, alphas, betas, False, disable_ssim)
        t1, t2 = fwAD.unpack_dual(d1).tangent, fwAD.unpack_dual(d2).tangent
    x = img.clone().requires_grad_(True)
    a1, a2 = block(x, mask, gt, alphas, betas, False, disable_ssim)
    ((a1 * u1).sum() + (a2 * u2).sum()).backward()
    out = dict(img=img, gt=gt, alphas=alphas.view(-1), betas=betas.view(-1), t_img=t_img, u1=u1, u2=u2, r1=r1, r2=r2,
               t1=t1, t2=t2, gx=x.grad, disable_ssim=torch.tensor(int(disable_ssim)))
    if mask is not None:
        out["mask"] = mask
    return {k: v.detach().numpy() for k, v in out.items()}


def main():
    ssim_pp, block = reference_functions()
    gen = torch.Generator().manual_seed(20261104)
    d = {}
    for name, args in (("a", (2, 37, 45, [(37, 45), (37, 45)], False, 0.2, False)),
                       ("b", (2, 40, 33, [(40, 33), (29, 33)], True, 0.2, False)),
                       ("c", (2, 24, 24, [(24, 24), (24, 24)], True, 0.2, True))):
        for k, v in case(block, gen, *args).items():
            d[f"{name}_{k}"] = v
    x = torch.rand((1, 3, 30, 41), generator=gen)
    y = torch.rand((1, 3, 30, 41), generator=gen)
    d["ssim_x"], d["ssim_y"], d["ssim_map"] = x.numpy(), y.numpy(), ssim_pp(x, y).numpy()
    np.savez_compressed(OUT, **d)
    print("wrote", OUT, os.path.getsize(OUT), "bytes;", len(d), "arrays")


if __name__ == "__main__":
    main()
