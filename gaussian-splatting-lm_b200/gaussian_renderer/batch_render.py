"""gaussian_renderer.batch_render.batch_render with the signature and return dict of the reference
(gaussian_renderer/batch_render.py:8-135): keys render [B,3,H,W] (clamped to [0,1]), viewspace_points,
visibility_filter, max_radii, depth, viewcount."""
import math

import torch

from diff_gaussian_rasterization.batch_render import BatchGaussianRasterizationSettings, BatchGaussianRasterizer

from ._common import gaussian_inputs


def batch_render(viewpoint_cameras, pc, pipe, bg_color, scaling_modifier=1.0, separate_sh=False, override_color=None,
                 use_trained_exp=False):
    """Render a list of views. Background tensor (bg_color) must be on the GPU."""
    screenspace_points = torch.zeros_like(pc.get_xyz, dtype=pc.get_xyz.dtype, requires_grad=True) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    raster_settings = BatchGaussianRasterizationSettings(
        batch_size=len(viewpoint_cameras),
        image_heights=[int(c.image_height) for c in viewpoint_cameras],
        image_widths=[int(c.image_width) for c in viewpoint_cameras],
        tanfovxs=[math.tan(c.FoVx * 0.5) for c in viewpoint_cameras],
        tanfovys=[math.tan(c.FoVy * 0.5) for c in viewpoint_cameras],
        bg=bg_color, scale_modifier=scaling_modifier,
        viewmatrices=[c.world_view_transform for c in viewpoint_cameras],
        projmatrices=[c.full_proj_transform for c in viewpoint_cameras],
        sh_degree=pc.active_sh_degree, camposes=[c.camera_center for c in viewpoint_cameras],
        prefiltered=False, debug=pipe.debug, antialiasing=pipe.antialiasing)
    rasterizer = BatchGaussianRasterizer(batch_raster_settings=raster_settings)
    kw = gaussian_inputs(pc, pipe, scaling_modifier, separate_sh, override_color)
    rendered_image, radii, depth_image = rasterizer(means2D=screenspace_points, **kw)
    if use_trained_exp:
        raise NotImplementedError("Batch exposure not implemented yet")   # as the reference (:112)
    rendered_image = rendered_image.clamp(0, 1)
    max_radii = radii.max(dim=0).values
    return {"render": rendered_image, "viewspace_points": screenspace_points,
            "visibility_filter": (max_radii > 0).nonzero(), "max_radii": max_radii, "depth": depth_image,
            "viewcount": (radii > 0).sum(dim=0)}
