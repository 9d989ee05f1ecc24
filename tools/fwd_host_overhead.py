"""Where does the forward call's wall time go beyond its kernels?  (a) the Python binding as shipped,
(b) the C entry point alone with a pre-allocated arena allocator (no torch.empty in the callback)."""
import sys, os, time, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("gaussian-splatting-lm_b200", "tests", "oracle"):
    sys.path.insert(0, os.path.join(ROOT, p))
import torch
import scene_synth as ss, helpers as h
from diff_gaussian_rasterization import _C as G
P = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
g = ss.make_gaussians(P, sh_degree=3, seed=1004)
cam = ss.hemisphere_rig(4, 1920, 1080)[1]
gd, ct = h.to_dev(g), h.cam_tensors(cam)
def wall(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
print("binding as shipped      ms/fwd", round(wall(lambda: h.c_forward(G, gd, ct)), 3))
L = G.lib()
e = torch.empty(0, device="cuda")
gs, k1 = G._gaussians(gd["means3D"], e, gd["opacities"], gd["scales"], gd["rotations"], 1.0, e, gd["shs"], 3)
cm, k2 = G._camera(ct["bg"], ct["viewmatrix"], ct["projmatrix"], ct["tan_fovx"], ct["tan_fovy"], ct["H"], ct["W"], ct["campos"], False, False, False)
arena = torch.empty(12 << 30, dtype=torch.uint8, device="cuda")
off = [0]
def alloc(_ctx, which, nbytes):
    p = arena.data_ptr() + off[0]
    off[0] += (int(nbytes) + 255) // 256 * 256
    return p
cb = G.ALLOC_FN(alloc)
color = torch.empty(3, ct["H"], ct["W"], device="cuda"); invd = torch.empty(1, ct["H"], ct["W"], device="cuda")
radii = torch.zeros(P, dtype=torch.int32, device="cuda")
n, nc = C.c_int32(0), C.c_int32(0)
def raw():
    off[0] = 0
    G._check(L.gslm_forward(C.addressof(gs), C.addressof(cm), cb, None, color.data_ptr(), invd.data_ptr(), radii.data_ptr(),
                            C.addressof(n), C.addressof(nc), G._stream()))
print("C entry, arena allocator ms/fwd", round(wall(raw), 3))
G.timing_enable(True); G.timing_collect()
for _ in range(5): raw()
st = G.timing_collect(); G.timing_enable(False)
print("kernel sum ms/fwd", round(sum(ms for ms, c in st.values()) / 5, 3), {k: round(ms / 5, 3) for k, (ms, c) in st.items() if c})
