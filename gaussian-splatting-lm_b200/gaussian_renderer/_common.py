"""Shared pieces of gaussian_renderer.render / batch_render: which Gaussian tensors are handed to the rasterizer
(gaussian_renderer/__init__.py:56-87 and gaussian_renderer/batch_render.py:54-85 of the reference)."""
import torch


def eval_sh_colors(pc, camera_center):
    """convert_SHs_python path: SH -> RGB in torch (utils/sh_utils.py:57-112 basis), clamp_min(+0.5, 0)."""
    from .sh import eval_sh
    feats = pc.get_features
    shs_view = feats.transpose(1, 2).view(-1, 3, (pc.max_sh_degree + 1) ** 2)
    dir_pp = pc.get_xyz - camera_center.view(1, 3)
    dir_pp = dir_pp / dir_pp.norm(dim=1, keepdim=True)
    return torch.clamp_min(eval_sh(pc.active_sh_degree, shs_view, dir_pp) + 0.5, 0.0)


def gaussian_inputs(pc, pipe, scaling_modifier, separate_sh, override_color, camera_center=None):
    kw = dict(means3D=pc.get_xyz, opacities=pc.get_opacity, shs=None, colors_precomp=None, scales=None,
              rotations=None, cov3D_precomp=None)
    if getattr(pipe, "compute_cov3D_python", False):
        kw["cov3D_precomp"] = pc.get_covariance(scaling_modifier)
    else:
        kw["scales"], kw["rotations"] = pc.get_scaling, pc.get_rotation
    if override_color is not None:
        kw["colors_precomp"] = override_color
    elif getattr(pipe, "convert_SHs_python", False):
        if camera_center is None:
            raise NotImplementedError("convert_SHs_python needs one camera centre (not available in batch_render)")
        kw["colors_precomp"] = eval_sh_colors(pc, camera_center)
    elif separate_sh:
        kw["dc"], kw["shs"] = pc.get_features_dc, pc.get_features_rest
    else:
        kw["shs"] = pc.get_features
    return kw
