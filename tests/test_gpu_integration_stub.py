"""The ctypes stub INTEGRATION.md shows a maintainer of the reference (Option B) is executed as written (only the
library path is made absolute) and must give the shipped binding's result: documentation that cannot rot."""
import os
import re

import pytest
import torch

import helpers as h
import scene_synth as ss

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_integration_md_stub_runs_and_matches_the_binding():
    from diff_gaussian_rasterization import _C
    md = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    blocks = re.findall(r"```python\n(.*?)```", md, flags=re.S)
    stub = next(b for b in blocks if "def rasterize_gaussians(" in b and "CDLL" in b)
    lib = os.path.join(ROOT, "gaussian-splatting-lm_b200", "libgslm_b200.so")
    ns = {}
    exec(stub.replace('"libgslm_b200.so"', repr(lib)), ns)
    g = ss.make_gaussians(3000, sh_degree=2, seed=5, scale_mult=0.8)
    cam = ss.hemisphere_rig(3, 144, 80)[1]
    gd, ct = h.to_dev(g), h.cam_tensors(cam, bg=(0.2, 0.1, 0.0))
    e = torch.empty(0, device="cuda")
    args = (ct["bg"], gd["means3D"], e, gd["opacities"], gd["scales"], gd["rotations"], 1.0, e, ct["viewmatrix"],
            ct["projmatrix"], ct["tan_fovx"], ct["tan_fovy"], ct["H"], ct["W"], gd["shs"], gd["sh_degree"], ct["campos"],
            False, False, False)
    (n, n_cached), color, radii, geom, binning, img, invd = ns["rasterize_gaussians"](*args)
    R, color2, radii2, _, _, _, invd2 = _C.rasterize_gaussians(*args)
    assert n == int(R) and n_cached == R.cached and n > n_cached > 0
    assert torch.equal(color, color2) and torch.equal(radii, radii2) and torch.equal(invd, invd2)
    assert geom.numel() > 0 and binning.numel() > 0 and img.numel() > 0
