"""Hit-mask statistics of the compact cache at config C (1 view): how dense are the masks, how many 4-row bands
(= the float2 pixel pairs of one lane) of a (tile, instance) are empty."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("gaussian-splatting-lm_b200", "tests", "oracle"):
    sys.path.insert(0, os.path.join(ROOT, p))
import torch, numpy as np
import scene_synth as ss, helpers as h
from diff_gaussian_rasterization import _C as C
g, cams = ss.make_config("C", sh_degree=3, views=2)
gd = h.to_dev(g)
ct = h.cam_tensors(cams[1])
R, color, radii, geom, binning, img, invd = h.c_forward(C, gd, ct)
P = g["means3D"].shape[0]
v = C.buffer_view(P, ct["W"], ct["H"], R, geom, binning, img)
m = v["masks"]  # (Rc, 8) int32
Rc = m.shape[0]
def popc(x):
    x = x.to(torch.int64) & 0xffffffff
    c = torch.zeros_like(x)
    for i in range(32):
        c += (x >> i) & 1
    return c
pc = popc(m)  # (Rc, 8)
tot = pc.sum(1)
print("R", int(R), "Rc", Rc, "mean hits per instance", float(tot.float().mean()), "of 256")
band = (m[:, 0::2] | m[:, 1::2])  # (Rc,4)
print("empty 4-row bands:", float((band == 0).float().mean()))
word = (m == 0).float().mean()
print("empty 2-row words:", float(word))
hist = torch.histc(tot.float(), bins=8, min=0, max=256)
print("hist hits/instance (bins of 32):", (hist / Rc).tolist())
# column occupancy: any hit in column x (bits x and 16+x of any word)
allw = m[:, 0]
for k in range(1, 8):
    allw = allw | m[:, k]
cols = ((allw & 0xffff) | ((allw >> 16) & 0xffff))
print("mean occupied columns:", float(popc(cols).float().mean()), "of 16")
rows_occ = (m != 0).sum(1).float().mean()
print("mean occupied 2-row words:", float(rows_occ), "of 8")
