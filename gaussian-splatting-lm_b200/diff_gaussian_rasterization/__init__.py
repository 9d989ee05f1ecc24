"""B200-native drop-in for the reference package `diff_gaussian_rasterization`
(DGR/diff_gaussian_rasterization/__init__.py): same public names, argument order, return
shapes and exception behaviour --

    GaussianRasterizationSettings, GaussianRasterizer (forward / markVisible),
    rasterize_gaussians, _RasterizeGaussians (forward / backward / jvp)

-- on top of the sm_100a kernels in libgslm_b200.so (see ._C).  Forward-mode tangents are
detected on the inputs exactly as the reference does (__init__.py:24-99) and the output
tangents are attached through autograd.Function.jvp + save_for_forward (:260-284).
"""
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _C
from .utils import get_tangent, has_tangent

_N_IN = 8   # means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    antialiasing: bool


def _settings_tangent(rs):
    return rs._replace(tanfovx=get_tangent(rs.tanfovx), tanfovy=get_tangent(rs.tanfovy), bg=get_tangent(rs.bg),
                       scale_modifier=get_tangent(rs.scale_modifier), viewmatrix=get_tangent(rs.viewmatrix),
                       projmatrix=get_tangent(rs.projmatrix), campos=get_tangent(rs.campos))


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    inputs = (means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp)
    rs = raster_settings
    jvp = any(has_tangent(x) for x in inputs + (rs.bg, rs.viewmatrix, rs.projmatrix, rs.campos))
    if not jvp:
        return _RasterizeGaussians.apply(*inputs, rs, False, *([None] * (_N_IN + 1)))
    tangents = tuple(get_tangent(x) for x in inputs)
    return _RasterizeGaussians.apply(*inputs, rs, True, *tangents, _settings_tangent(rs))


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings, jvp, means3D_tangent, means2D_tangent, sh_tangent, colors_precomp_tangent,
                opacities_tangent, scales_tangent, rotations_tangent, cov3Ds_precomp_tangent,
                raster_settings_tangent):
        rs = raster_settings
        if not jvp:
            (num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer,
             invdepths) = _C.rasterize_gaussians(
                rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, sh,
                rs.sh_degree, rs.campos, rs.prefiltered, rs.antialiasing, rs.debug)
        else:
            rt = raster_settings_tangent
            (num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer, invdepths, color_grad,
             invdepths_grad) = _C.rasterize_gaussians_jvp(
                rs.bg, rt.bg, means3D, means3D_tangent, colors_precomp, colors_precomp_tangent, opacities,
                opacities_tangent, scales, scales_tangent, rotations, rotations_tangent, rs.scale_modifier,
                rt.scale_modifier, cov3Ds_precomp, cov3Ds_precomp_tangent, rs.viewmatrix, rt.viewmatrix,
                rs.projmatrix, rt.projmatrix, rs.tanfovx, rt.tanfovx, rs.tanfovy, rt.tanfovy, rs.image_height,
                rs.image_width, sh, sh_tangent, rs.sh_degree, rs.campos, rt.campos, rs.prefiltered,
                rs.antialiasing, rs.debug)
            ctx.save_for_forward(color_grad, invdepths_grad)
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, opacities,
                              geomBuffer, binningBuffer, imgBuffer)
        ctx.mark_non_differentiable(radii)
        return color, radii, invdepths

    @staticmethod
    def backward(ctx, grad_out_color, _, grad_out_depth):
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, opacities, geomBuffer,
         binningBuffer, imgBuffer) = ctx.saved_tensors
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh, grad_scales,
         grad_rotations) = _C.rasterize_gaussians_backward(
            rs.bg, means3D, radii, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
            rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_out_color, grad_out_depth, sh, rs.sh_degree,
            rs.campos, geomBuffer, ctx.num_rendered, binningBuffer, imgBuffer, rs.antialiasing, rs.debug)

        def shaped(g, like):
            return g.reshape(like.shape) if like.numel() else None

        return (grad_means3D, grad_means2D, shaped(grad_sh, sh), shaped(grad_colors_precomp, colors_precomp),
                shaped(grad_opacities, opacities), shaped(grad_scales, scales), shaped(grad_rotations, rotations),
                shaped(grad_cov3Ds_precomp, cov3Ds_precomp)) + (None,) * (2 + _N_IN + 1)

    @staticmethod
    def jvp(ctx, *_tangents):
        color_grad, invdepths_grad = ctx.saved_tensors
        return color_grad, None, invdepths_grad


def _check_exclusive(shs, colors_precomp, scales, rotations, cov3D_precomp):
    if (shs is None) == (colors_precomp is None):
        raise Exception('Please provide excatly one of either SHs or precomputed colors!')
    if ((scales is None or rotations is None) and cov3D_precomp is None) or \
            ((scales is not None or rotations is not None) and cov3D_precomp is not None):
        raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        _check_exclusive(shs, colors_precomp, scales, rotations, cov3D_precomp)
        empty = torch.Tensor([])
        return rasterize_gaussians(
            means3D, means2D,
            empty if shs is None else shs,
            empty if colors_precomp is None else colors_precomp,
            opacities,
            empty if scales is None else scales,
            empty if rotations is None else rotations,
            empty if cov3D_precomp is None else cov3D_precomp,
            self.raster_settings)
