import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("gaussian-splatting-lm_b200", "tests", "oracle"):
    sys.path.insert(0, os.path.join(ROOT, p))
import torch, numpy as np
import scene_synth as ss, helpers as h
from diff_gaussian_rasterization import _C as C
g = ss.make_gaussians(3000, sh_degree=2, seed=21, scale_mult=0.8)
t = ss.random_tangent(g, seed=22)
gd, td = h.to_dev(g), h.to_dev(t)
cams = ss.hemisphere_rig(10, 112, 80) + ss.hemisphere_rig(9, 96, 64)
views = []
for i, cam in enumerate(cams):
    ct = h.cam_tensors(cam, bg=(0.1, 0.0, 0.2))
    R, color, radii, geom, binning, img, invd = h.c_forward(C, gd, ct)
    w = None if i % 3 == 0 else (torch.rand(color.shape) > 0.3).float().cuda() * 2.0
    views.append((h.cam_args(ct), radii, geom, R, binning, img, w))
P, M = g["means3D"].shape[0], g["shs"].shape[1]
zz = lambda *s: torch.zeros(s, device="cuda")
mk = lambda: (zz(P, 3), None, zz(P, 1), zz(P, 3), None, zz(P, M, 3), zz(P, 3), zz(P, 4))
def single(nv):
    o = mk()
    for v in views[:nv]:
        C.jtjv(h.gauss_args(gd), h.tangent_args(td), *v, o)
    return o
def batch(nv):
    o = mk()
    plan = C.JtjvBatchPlan(views[:nv], torch.device("cuda"))
    C.jtjv_batch(h.gauss_args(gd), h.tangent_args(td), plan, o)
    return o
names = ["means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rot"]
for nv in (1, 2, 3, 19):
    a, a2, b, b2 = single(nv), single(nv), batch(nv), batch(nv)
    for i, n in enumerate(names):
        if a[i] is None: continue
        print(nv, n, "single==single", torch.equal(a[i], a2[i]), "batch==batch", torch.equal(b[i], b2[i]),
              "single==batch", torch.equal(a[i], b[i]), float((a[i] - b[i]).abs().max()), float(a[i].abs().max()))
plan = C.JtjvBatchPlan(views, torch.device("cuda"))
ref = single(19)
out2 = tuple(None if o is None else torch.full_like(o, 7.0) for o in ref)
C.jtjv_batch(h.gauss_args(gd), h.tangent_args(td), plan, out2, accumulate=False)
for i, n in enumerate(names):
    if ref[i] is None: continue
    d = (out2[i] - ref[i]).abs()
    print("overwrite", n, torch.equal(out2[i], ref[i]), float(d.max()), int((d > 0).sum()), d.flatten().argmax().item())
