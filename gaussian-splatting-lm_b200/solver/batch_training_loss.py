"""The per-pixel residual vector r(theta) of a batch of views, with the signature of the reference's
solver/batch_training_loss.py:33-105 (and compute_batch_loss_block :10-31).

disable_ssim=True (what the LM driver uses, train_jvp.py:212): both residual blocks are the SAME tensor
`images * alpha_mask - gt` (so ||r||^2 counts every pixel twice and J^T J carries a factor 2).
Otherwise: sqrt((1-lambda)/n) * sqrt(|image - gt| + 1e-6) and sqrt(lambda/n) * sqrt(|1 - SSIM| + 1e-6), n = 3 H W.
On CUDA tensors that second form is ONE kernel of libgslm_b200.so (csrc/loss_blocks.cu) with its own forward-mode and
reverse-mode rules (`_LossBlocks`), where the reference runs 5 grouped 11x11 convolutions plus ~20 elementwise
kernels through autograd / forward AD."""
import math

import torch

from diff_gaussian_rasterization import _C
from diff_gaussian_rasterization.utils import get_tangent, has_tangent
from gaussian_renderer.batch_render import batch_render
from solver.loss_image_state import BatchLossImageState
from solver.loss_utils import l1_loss_per_pixel, ssim_per_pixel


class _LossBlocks(torch.autograd.Function):
    """(Ll1_per_pixel, ssim_loss_per_pixel) of compute_batch_loss_block with disable_ssim=False, differentiable
    in `images` in both modes (the tangent travels explicitly, like _RasterizeGaussians)."""

    @staticmethod
    def forward(ctx, images, alpha_masks, gt_images, per_image_alphas, per_image_betas, images_tangent):
        lb = _C.LossBatch(images, alpha_masks, gt_images, per_image_alphas, per_image_betas)
        r1, r2 = lb.blocks(want_lin=True)
        if images_tangent is not None:
            ctx.save_for_forward(*lb.jvp(images_tangent))
        ctx.loss_batch = lb
        return r1, r2

    @staticmethod
    def backward(ctx, u1, u2):
        return (ctx.loss_batch.vjp(u1, u2),) + (None,) * 5

    @staticmethod
    def jvp(ctx, *_tangents):
        t1, t2 = ctx.saved_tensors
        return t1, t2


def compute_batch_loss_block(images, alpha_masks, gt_images, per_image_alphas, per_image_betas,
                             FUSED_SSIM_AVAILABLE=False, disable_ssim=False):
    if disable_ssim:
        if alpha_masks is not None:
            images = images * alpha_masks
        r = images - gt_images
        return r, r
    if images.is_cuda:
        return _LossBlocks.apply(images, alpha_masks, gt_images, per_image_alphas, per_image_betas,
                                 get_tangent(images) if has_tangent(images) else None)
    # host tensors only (CPU unit tests of the host logic): the reference's own torch composition
    if alpha_masks is not None:
        images = images * alpha_masks
    l1 = l1_loss_per_pixel(images, gt_images)
    ssim_loss = (1.0 - ssim_per_pixel(images, gt_images)).abs()
    return per_image_alphas * torch.sqrt(l1 + 1e-6), per_image_betas * torch.sqrt(ssim_loss + 1e-6)


def gather_targets(viewpoint_cams, like):
    """gt images [B,3,Hmax,Wmax] (zero padded) and alpha masks [B,1,Hmax,Wmax] or None, on like.device."""
    B = len(viewpoint_cams)
    gt = torch.zeros(like.shape, dtype=like.dtype, device=like.device)
    for i, vc in enumerate(viewpoint_cams):
        H, W = vc.image_height, vc.image_width
        gt[i, :, :H, :W] = vc.original_image.to(like.device)
    masks = None
    if any(getattr(vc, "alpha_mask", None) is not None for vc in viewpoint_cams):
        masks = torch.zeros((B, 1) + tuple(like.shape[-2:]), device=like.device, dtype=like.dtype)
        for i, vc in enumerate(viewpoint_cams):
            if vc.alpha_mask is not None:
                H, W = vc.image_height, vc.image_width
                masks[i, :, :H, :W] = vc.alpha_mask.to(like.device)
    return gt, masks


def batch_training_loss(iteration, opt, viewpoint_cams, gaussians, pipe, bg, train_test_exp, depth_l1_weight,
                        batch_stats=None, SPARSE_ADAM_AVAILABLE=False, FUSED_SSIM_AVAILABLE=False,
                        disable_ssim=False):
    B = len(viewpoint_cams)
    sizes_list = [(vc.image_height, vc.image_width) for vc in viewpoint_cams]
    pkg = batch_render(viewpoint_cams, gaussians, pipe, bg, use_trained_exp=train_test_exp,
                       separate_sh=SPARSE_ADAM_AVAILABLE)
    images = pkg["render"]
    if batch_stats is not None:
        batch_stats["viewspace_point_tensor"] = pkg["viewspace_points"]
        batch_stats["visibility_filter"] = pkg["visibility_filter"]
        batch_stats["max_radii"] = pkg["max_radii"]
        batch_stats["viewcount"] = pkg.get("viewcount", None)
    gt_images, alpha_masks = gather_targets(viewpoint_cams, images)
    lam = opt.lambda_dssim
    alphas = torch.tensor([math.sqrt((1.0 - lam) / (3 * H * W)) for H, W in sizes_list], dtype=images.dtype,
                          device=images.device).view(B, 1, 1, 1)
    betas = torch.tensor([math.sqrt(lam / (3 * H * W)) for H, W in sizes_list], dtype=images.dtype,
                         device=images.device).view(B, 1, 1, 1)
    l1_block, ssim_block = compute_batch_loss_block(images, alpha_masks, gt_images, alphas, betas,
                                                    FUSED_SSIM_AVAILABLE, disable_ssim)
    has_depth = any(getattr(vc, "depth_reliable", False) for vc in viewpoint_cams)
    if depth_l1_weight(iteration) > 0 and has_depth:
        raise NotImplementedError("Ll1depth_pure_per_pixel is not implemented in this version.")   # as the reference
    depth_block = torch.zeros((0,), dtype=l1_block.dtype, device=l1_block.device, requires_grad=True)
    return BatchLossImageState(l1_block, ssim_block, depth_block, sizes_list, has_depth)
