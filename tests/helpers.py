"""Shared helpers of the test-suite: scene -> torch tensors, thin wrappers over the product's `_C` surface,
and error metrics."""
import math

import numpy as np
import torch


def relerr(a, b):
    """max |a-b| / max(|b|_max, tiny): the 'rel' of BASELINE.json's tolerances (scale = largest reference entry)."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    if a.size == 0:
        return 0.0
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def rel_l2(a, b):
    a = np.asarray(a, np.float64).ravel()
    b = np.asarray(b, np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def to_dev(g, dev="cuda"):
    return {k: (torch.from_numpy(np.ascontiguousarray(v)).to(dev) if isinstance(v, np.ndarray) else v)
            for k, v in g.items()}


def cam_tensors(cam, bg=(0.0, 0.0, 0.0), dev="cuda"):
    """camera namespace (scene_synth) -> kwargs of the _C calls."""
    return dict(bg=torch.tensor(bg, dtype=torch.float32, device=dev),
                viewmatrix=torch.from_numpy(cam.world_view_transform).to(dev),
                projmatrix=torch.from_numpy(cam.full_proj_transform).to(dev),
                campos=torch.from_numpy(cam.camera_center).to(dev),
                tan_fovx=math.tan(cam.FoVx * 0.5), tan_fovy=math.tan(cam.FoVy * 0.5),
                H=cam.image_height, W=cam.image_width)


def c_forward(_C, gd, ct, antialiasing=False, scale_modifier=1.0, debug=False, keep_sort_scratch=False):
    """keep_sort_scratch: keep the sort scratch of this call alive in _C.LAST_SORT_SCRATCH (sorted keys / full
    point_list for the bit-exactness checks)."""
    e = torch.empty(0, device=gd["means3D"].device)
    _C.KEEP_SORT_SCRATCH = bool(keep_sort_scratch)
    try:
        return _c_forward(_C, gd, ct, antialiasing, scale_modifier, debug, e)
    finally:
        _C.KEEP_SORT_SCRATCH = False


def _c_forward(_C, gd, ct, antialiasing, scale_modifier, debug, e):
    return _C.rasterize_gaussians(ct["bg"], gd["means3D"], gd.get("colors_precomp", e), gd["opacities"],
                                  gd.get("scales", e), gd.get("rotations", e), scale_modifier,
                                  gd.get("cov3D_precomp", e), ct["viewmatrix"], ct["projmatrix"], ct["tan_fovx"],
                                  ct["tan_fovy"], ct["H"], ct["W"], gd.get("shs", e), gd.get("sh_degree", 0),
                                  ct["campos"], False, antialiasing, debug)


def c_backward(_C, gd, ct, fwd, dL_dcolor, dL_dinvdepth=None, antialiasing=False, scale_modifier=1.0):
    e = torch.empty(0, device=gd["means3D"].device)
    R, color, radii, geom, binning, img, invd = fwd
    return _C.rasterize_gaussians_backward(ct["bg"], gd["means3D"], radii, gd.get("colors_precomp", e),
                                           gd["opacities"], gd.get("scales", e), gd.get("rotations", e),
                                           scale_modifier, gd.get("cov3D_precomp", e), ct["viewmatrix"],
                                           ct["projmatrix"], ct["tan_fovx"], ct["tan_fovy"], dL_dcolor,
                                           dL_dinvdepth if dL_dinvdepth is not None else e, gd.get("shs", e),
                                           gd.get("sh_degree", 0), ct["campos"], geom, R, binning, img,
                                           antialiasing, False)


def gauss_args(gd, scale_modifier=1.0):
    e = torch.empty(0, device=gd["means3D"].device)
    return (gd["means3D"], gd.get("colors_precomp", e), gd["opacities"], gd.get("scales", e), gd.get("rotations", e),
            scale_modifier, gd.get("cov3D_precomp", e), gd.get("shs", e), gd.get("sh_degree", 0))


def tangent_args(td):
    g = lambda k: td.get(k)
    return (g("means3D"), g("colors_precomp"), g("opacities"), g("scales"), g("rotations"), g("cov3D_precomp"),
            g("shs"))


def cam_args(ct, antialiasing=False):
    return (ct["bg"], ct["viewmatrix"], ct["projmatrix"], ct["tan_fovx"], ct["tan_fovy"], ct["H"], ct["W"],
            ct["campos"], antialiasing, False)
