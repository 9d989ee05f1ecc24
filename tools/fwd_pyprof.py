import sys, os, time, cProfile, pstats
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("gaussian-splatting-lm_b200", "tests", "oracle"):
    sys.path.insert(0, os.path.join(ROOT, p))
import torch
import scene_synth as ss, helpers as h
from diff_gaussian_rasterization import _C as G
P = 1000000
g = ss.make_gaussians(P, sh_degree=3, seed=1004)
cam = ss.hemisphere_rig(4, 1920, 1080)[1]
gd, ct = h.to_dev(g), h.cam_tensors(cam)
for _ in range(3): f = h.c_forward(G, gd, ct)
torch.cuda.synchronize()
orig = G._Alloc._alloc
times = []
def timed_alloc(self, ctx, which, nbytes):
    t0 = time.perf_counter(); r = orig(self, ctx, which, nbytes); times.append((int(which), int(nbytes), (time.perf_counter() - t0) * 1e3)); return r
G._Alloc._alloc = timed_alloc
pr = cProfile.Profile(); pr.enable()
t0 = time.perf_counter()
for _ in range(10): f = h.c_forward(G, gd, ct)
torch.cuda.synchronize()
print("wall ms/fwd", (time.perf_counter() - t0) / 10 * 1e3)
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(14)
for t in times[:12]: print(t)
print(torch.cuda.memory_stats()["num_device_alloc"], torch.cuda.memory_stats()["num_device_free"])
