"""gaussian_renderer.render with the signature and return dict of the reference
(gaussian_renderer/__init__.py:18-128): keys render, viewspace_points, visibility_filter, radii, depth."""
import math

import torch

from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer

from ._common import gaussian_inputs


def render(viewpoint_camera, pc, pipe, bg_color, scaling_modifier=1.0, separate_sh=False, override_color=None,
           use_trained_exp=False):
    """Render one view. Background tensor (bg_color) must be on the GPU."""
    # zero tensor whose gradient receives the screen-space (2D mean) gradients
    screenspace_points = torch.zeros_like(pc.get_xyz, dtype=pc.get_xyz.dtype, requires_grad=True) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    raster_settings = GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width),
        tanfovx=math.tan(viewpoint_camera.FoVx * 0.5), tanfovy=math.tan(viewpoint_camera.FoVy * 0.5), bg=bg_color,
        scale_modifier=scaling_modifier, viewmatrix=viewpoint_camera.world_view_transform,
        projmatrix=viewpoint_camera.full_proj_transform, sh_degree=pc.active_sh_degree,
        campos=viewpoint_camera.camera_center, prefiltered=False, debug=pipe.debug,
        antialiasing=pipe.antialiasing)
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)
    kw = gaussian_inputs(pc, pipe, scaling_modifier, separate_sh, override_color, viewpoint_camera.camera_center)
    dc = kw.pop("dc", None)
    if dc is not None:   # the single-view rasterizer takes one SH tensor
        kw["shs"] = torch.cat((dc, kw["shs"]), dim=1)
    rendered_image, radii, depth_image = rasterizer(means2D=screenspace_points, **kw)
    if use_trained_exp:
        exposure = pc.get_exposure_from_name(viewpoint_camera.image_name)
        rendered_image = torch.matmul(rendered_image.permute(1, 2, 0), exposure[:3, :3]).permute(2, 0, 1) + \
            exposure[:3, 3, None, None]
    rendered_image = rendered_image.clamp(0, 1)
    return {"render": rendered_image, "viewspace_points": screenspace_points,
            "visibility_filter": (radii > 0).nonzero(), "radii": radii, "depth": depth_image}
