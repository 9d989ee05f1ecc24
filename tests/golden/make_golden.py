"""Builds tests/golden/dgr_fixture.npz from the reference's OWN test fixtures.

Run in the build container only (reads /root/reference, which does not exist on the GPU box):
    python tests/golden/make_golden.py
Sources (all under /root/reference/submodules/diff-gaussian-rasterization/tests/):
  data/args.pth                     exact fp32 inputs of one rasterize_gaussians call (P=17073, 640x480, SH deg 0)
  render_data/*_data.csv            preprocessCUDA outputs dumped by the reference (6 significant digits)
  render_data/*_grad.csv            preprocessCUDAJvp tangents for the camera tangents of forward_tests.cu:398-409
  render_data/point_offsets.csv, point_list.csv, img_ranges.csv   exact integer binning results
Nothing is computed here: the file is a re-packing (npz, compressed) of reference-produced numbers.
"""
import os
import numpy as np
import torch

REF = "/root/reference/submodules/diff-gaussian-rasterization/tests"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "dgr_fixture.npz")


def csv(name, dtype=np.float64):
    return np.loadtxt(os.path.join(REF, "render_data", name), dtype=dtype, ndmin=2)


def main():
    a = torch.load(os.path.join(REF, "data", "args.pth"), map_location="cpu", weights_only=False)
    a = tuple(x.detach() if isinstance(x, torch.Tensor) else x for x in a)
    # argument order of _C.rasterize_gaussians (diff_gaussian_rasterization/__init__.py:133-154)
    d = dict(
        bg=a[0].numpy(), means3D=a[1].numpy(), opacities=a[3].numpy(), scales=a[4].numpy(), rotations=a[5].numpy(),
        scale_modifier=np.float32(a[6]), viewmatrix=a[8].numpy().reshape(16), projmatrix=a[9].numpy().reshape(16),
        projmatrix_raw=a[10].numpy().reshape(16), tan_fovx=np.float64(a[11]), tan_fovy=np.float64(a[12]),
        H=np.int32(a[13]), W=np.int32(a[14]), shs=a[15].numpy(), sh_degree=np.int32(a[16]),
        campos_raw=a[17].numpy().reshape(16), prefiltered=np.bool_(a[18]), debug=np.bool_(a[19]),
    )
    tile_grid = csv("tile_grid.csv", np.int64).reshape(-1)
    T = int(tile_grid[0] * tile_grid[1])
    d.update(
        tile_grid=tile_grid.astype(np.int32),
        means2D=csv("means2D_data.csv").astype(np.float32), conic_opacity=csv("conic_opacity_data.csv").astype(np.float32),
        rgb=csv("rgb_data.csv").astype(np.float32), depth=csv("depth_data.csv").astype(np.float32).reshape(-1),
        t_means2D=csv("means2D_grad.csv").astype(np.float32), t_conic_opacity=csv("conic_opacity_grad.csv").astype(np.float32),
        t_rgb=csv("rgb_grad.csv").astype(np.float32), t_depth=csv("depth_grad.csv").astype(np.float32).reshape(-1),
        point_offsets=csv("point_offsets.csv", np.int64).reshape(-1).astype(np.uint32),
        point_list=csv("point_list.csv", np.int64).reshape(-1).astype(np.uint32),
        # the reference over-allocates ranges per pixel (rasterizer_impl.cu:200); only the first T rows are tiles
        ranges=csv("img_ranges.csv", np.int64)[:T].astype(np.uint32),
    )
    np.savez_compressed(OUT, **d)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
