"""One forward pass configuration for profiling: python tools/fwd_one.py [P] [culling 0|1] [iters]"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("gaussian-splatting-lm_b200", "tests", "oracle"):
    sys.path.insert(0, os.path.join(ROOT, p))
import torch
import scene_synth as ss, helpers as h
from diff_gaussian_rasterization import _C as C
P = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
C.TILE_CULLING = bool(int(sys.argv[2])) if len(sys.argv) > 2 else True
N = int(sys.argv[3]) if len(sys.argv) > 3 else 3
g = ss.make_gaussians(P, sh_degree=3, seed=1004)
cam = ss.hemisphere_rig(4, 1920, 1080)[1]
gd, ct = h.to_dev(g), h.cam_tensors(cam)
for _ in range(N):
    fwd = h.c_forward(C, gd, ct)
torch.cuda.synchronize()
print(P, C.TILE_CULLING, int(fwd[0]), fwd[0].cached)
