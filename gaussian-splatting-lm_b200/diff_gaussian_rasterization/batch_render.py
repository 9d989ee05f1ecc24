"""Multi-view rasterizer with the surface of DGR/diff_gaussian_rasterization/batch_render.py:
BatchGaussianRasterizationSettings, BatchGaussianRasterizer, batch_rasterize_gaussians,
_BatchRasterizeGaussians (forward / backward / jvp), BinningBufferManager, allocate_batch_buffers.

Outputs as the reference (:409-447): colors [B,3,Hmax,Wmax], radii [B,P] int32, invdepths [B,1,Hmax,Wmax].
Differences that do not change results:
  * views of different sizes are supported (zero padding) -- the reference's `colors[i] = color` (:193) only
    works for equal sizes;
  * the backward pass accumulates the per-view gradients inside the kernel (GslmGrads.accumulate) instead of
    allocating eight zero-filled tensors per view and summing in Python (:330-346);
  * no torch.cuda.empty_cache() per view (:205,268).
"""
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _C
from .utils import get_tangent, has_tangent


class BatchGaussianRasterizationSettings(NamedTuple):
    batch_size: int
    image_heights: list
    image_widths: list
    tanfovxs: list
    tanfovys: list
    bg: torch.Tensor
    scale_modifier: float
    viewmatrices: list
    projmatrices: list
    sh_degree: int
    camposes: list
    prefiltered: bool
    debug: bool
    antialiasing: bool


def allocate_batch_buffers(B, size, device="cuda", align=128):
    size = ((size + align - 1) // align) * align
    return torch.empty((B, size), device=device, dtype=torch.uint8)


class BinningBufferManager:
    """Keeps the per-view binning buffers of a batch (reference :92-122 packs them into one growing tensor with
    1.2x head-room and copies each buffer in; here each view's tensor is simply retained -- no copy)."""

    def __init__(self, batch_size, grow_factor=1.2, align=128, device="cuda"):
        self.batch_size = batch_size
        self.grow_factor = grow_factor
        self.align = align
        self.max_size = 0
        self.buffers = []

    def store_buffer(self, binningBuffer, batch_index):
        self.max_size = max(self.max_size, binningBuffer.numel())
        if batch_index == len(self.buffers):
            self.buffers.append(binningBuffer)
        else:
            self.buffers[batch_index] = binningBuffer

    def get(self, batch_index):
        return self.buffers[batch_index]


def _settings_tangent(bs):
    return bs._replace(tanfovxs=get_tangent(bs.tanfovxs), tanfovys=get_tangent(bs.tanfovys), bg=get_tangent(bs.bg),
                       scale_modifier=get_tangent(bs.scale_modifier), viewmatrices=get_tangent(bs.viewmatrices),
                       projmatrices=get_tangent(bs.projmatrices), camposes=get_tangent(bs.camposes))


def batch_rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                              batch_raster_settings):
    bs = batch_raster_settings
    inputs = (means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp)
    jvp = any(has_tangent(x) for x in inputs + (bs.bg, bs.viewmatrices, bs.projmatrices, bs.camposes))
    if not jvp:
        return _BatchRasterizeGaussians.apply(*inputs, bs, False, *([None] * 9))
    tangents = tuple(get_tangent(x) for x in inputs)
    return _BatchRasterizeGaussians.apply(*inputs, bs, True, *tangents, _settings_tangent(bs))


class _BatchRasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                batch_raster_settings, jvp, means3D_tangent, means2D_tangent, sh_tangent, colors_precomp_tangent,
                opacities_tangent, scales_tangent, rotations_tangent, cov3Ds_precomp_tangent,
                batch_raster_settings_tangent):
        bs = batch_raster_settings
        B, P = bs.batch_size, means3D.shape[0]
        dev = means3D.device
        Hm, Wm = max(bs.image_heights), max(bs.image_widths)
        ragged = any(h != Hm for h in bs.image_heights) or any(w != Wm for w in bs.image_widths)
        mk = torch.zeros if ragged else torch.empty
        colors = mk((B, 3, Hm, Wm), dtype=torch.float32, device=dev)
        invdepths = mk((B, 1, Hm, Wm), dtype=torch.float32, device=dev)
        radii = torch.empty((B, P), dtype=torch.int32, device=dev)
        colors_t = invdepths_t = None
        if jvp:
            colors_t = mk((B, 3, Hm, Wm), dtype=torch.float32, device=dev)
            invdepths_t = mk((B, 1, Hm, Wm), dtype=torch.float32, device=dev)
        geoms, imgs, nums = [], [], []
        binning = BinningBufferManager(B, device=dev)
        bt = batch_raster_settings_tangent
        for i in range(B):
            H, W = bs.image_heights[i], bs.image_widths[i]
            if not jvp:
                (n, color, rad, geom, binb, img, invd) = _C.rasterize_gaussians(
                    bs.bg, means3D, colors_precomp, opacities, scales, rotations, bs.scale_modifier, cov3Ds_precomp,
                    bs.viewmatrices[i], bs.projmatrices[i], bs.tanfovxs[i], bs.tanfovys[i], H, W, sh, bs.sh_degree,
                    bs.camposes[i], bs.prefiltered, bs.antialiasing, bs.debug)
            else:
                (n, color, rad, geom, binb, img, invd, color_t, invd_t) = _C.rasterize_gaussians_jvp(
                    bs.bg, bt.bg, means3D, means3D_tangent, colors_precomp, colors_precomp_tangent, opacities,
                    opacities_tangent, scales, scales_tangent, rotations, rotations_tangent, bs.scale_modifier,
                    bt.scale_modifier, cov3Ds_precomp, cov3Ds_precomp_tangent, bs.viewmatrices[i], bt.viewmatrices[i],
                    bs.projmatrices[i], bt.projmatrices[i], bs.tanfovxs[i], bt.tanfovxs[i], bs.tanfovys[i],
                    bt.tanfovys[i], H, W, sh, sh_tangent, bs.sh_degree, bs.camposes[i], bt.camposes[i],
                    bs.prefiltered, bs.antialiasing, bs.debug)
                colors_t[i, :, :H, :W] = color_t
                invdepths_t[i, :, :H, :W] = invd_t
            colors[i, :, :H, :W] = color
            invdepths[i, :, :H, :W] = invd
            radii[i] = rad
            geoms.append(geom)
            imgs.append(img)
            nums.append(n)
            binning.store_buffer(binb, i)
        if jvp:
            ctx.save_for_forward(colors_t, invdepths_t)
        ctx.batch_raster_settings = bs
        ctx.num_rendered = nums
        ctx.buffers = (geoms, binning, imgs)
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, opacities)
        ctx.mark_non_differentiable(radii)
        return colors, radii, invdepths

    @staticmethod
    def backward(ctx, grad_out_colors, _, grad_out_depths):
        bs = ctx.batch_raster_settings
        colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, opacities = ctx.saved_tensors
        geoms, binning, imgs = ctx.buffers
        P, M, dev = means3D.shape[0], (sh.shape[1] if sh.numel() else 0), means3D.device
        z = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)
        acc = (z(P, 3), z(P, 3), z(P, 1), z(P, 3), z(P, 6), z(P, M, 3), z(P, 3), z(P, 4))
        for i in range(bs.batch_size):
            H, W = bs.image_heights[i], bs.image_widths[i]
            gd = grad_out_depths[i, :, :H, :W] if grad_out_depths is not None else None
            _C.rasterize_gaussians_backward(
                bs.bg, means3D, radii[i], colors_precomp, opacities, scales, rotations, bs.scale_modifier,
                cov3Ds_precomp, bs.viewmatrices[i], bs.projmatrices[i], bs.tanfovxs[i], bs.tanfovys[i],
                grad_out_colors[i, :, :H, :W], gd, sh, bs.sh_degree, bs.camposes[i], geoms[i], ctx.num_rendered[i],
                binning.get(i), imgs[i], bs.antialiasing, bs.debug, accumulate_into=acc)
        (g_means2D, g_colors, g_opac, g_means3D, g_cov, g_sh, g_scales, g_rot) = acc

        def shaped(g, like):
            return g.reshape(like.shape) if like.numel() else None

        return (g_means3D, g_means2D, shaped(g_sh, sh), shaped(g_colors, colors_precomp), shaped(g_opac, opacities),
                shaped(g_scales, scales), shaped(g_rot, rotations), shaped(g_cov, cov3Ds_precomp)) + (None,) * 11

    @staticmethod
    def jvp(ctx, *_tangents):
        colors_t, invdepths_t = ctx.saved_tensors
        return colors_t, None, invdepths_t


class BatchGaussianRasterizer(nn.Module):
    def __init__(self, batch_raster_settings):
        super().__init__()
        self.batch_raster_settings = batch_raster_settings

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None, dc=None):
        if dc is not None:   # `separate_sh=True` callers (gaussian_renderer/batch_render.py:92): superset of the reference
            shs = dc if shs is None else torch.cat((dc, shs), dim=1)
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        empty = torch.Tensor([])
        return batch_rasterize_gaussians(
            means3D, means2D,
            empty if shs is None else shs,
            empty if colors_precomp is None else colors_precomp,
            opacities,
            empty if scales is None else scales,
            empty if rotations is None else rotations,
            empty if cov3D_precomp is None else cov3D_precomp,
            self.batch_raster_settings)
