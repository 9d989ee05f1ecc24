import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("gaussian-splatting-lm_b200", "tests", "oracle"):
    sys.path.insert(0, os.path.join(ROOT, p))
import torch
import scene_synth as ss, helpers as h
from diff_gaussian_rasterization import _C as C
import itertools
for P, cull in itertools.product((100000, 1000000, 3000000), (False, True)):
    C.TILE_CULLING = cull
    g = ss.make_gaussians(P, sh_degree=3, seed=1004)
    cam = ss.hemisphere_rig(4, 1920, 1080)[1]
    gd, ct = h.to_dev(g), h.cam_tensors(cam)
    for _ in range(3):
        fwd = h.c_forward(C, gd, ct)
    torch.cuda.synchronize()
    C.timing_enable(True); C.timing_collect()
    t0 = time.perf_counter()
    N = 5
    for _ in range(N):
        fwd = h.c_forward(C, gd, ct)
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / N * 1e3
    st = C.timing_collect(); C.timing_enable(False)
    print(P, "culling" if cull else "reference binning", "R", int(fwd[0]), "wall ms/fwd", round(wall, 3), {k: round(ms / N, 3) for k, (ms, c) in st.items() if c}, "sum", round(sum(ms for ms, c in st.values()) / N, 3), flush=True)
    print(torch.cuda.memory_stats()["num_alloc_retries"], torch.cuda.memory_reserved() / 1e9)
    del fwd; torch.cuda.empty_cache()
