#!/usr/bin/env python
"""Residual-side microbenchmark: the loss-block kernels (csrc/loss_blocks.cu) at BASELINE config C image size against
the reference's own code path for the same quantity -- its torch composition (5 grouped 11x11 F.conv2d + elementwise,
utils/loss_utils.py:98-117, solver/batch_training_loss.py:10-31) run on the same GPU through autograd / forward AD.

    python tools/bench_loss.py [--views 16] [--width 1920] [--height 1080] [--iters 10]
Algorithmic bytes per pixel-channel (fp32; alpha mask adds 4/3): blocks 8 in + 24 out, JVP 28 in + 8 out,
VJP 32 in + 4 out.  Prints one JSON line."""
import argparse
import json
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gaussian-splatting-lm_b200"))

import torch  # noqa: E402
import torch.autograd.forward_ad as fwAD  # noqa: E402
import torch.nn.functional as F  # noqa: E402

from diff_gaussian_rasterization import _C as C  # noqa: E402


def timed(fn, iters, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def reference_composition(images, masks, gt, al, be):
    """The reference's formula with its own 2-D window (loss_utils.py:53-57,98-117)."""
    g = torch.tensor([math.exp(-(x - 5) ** 2 / 4.5) for x in range(11)])
    g = (g / g.sum()).unsqueeze(1)
    w = g.mm(g.t()).float()[None, None].expand(3, 1, 11, 11).contiguous().to(images.device)
    x = images * masks
    mu1, mu2 = F.conv2d(x, w, padding=5, groups=3), F.conv2d(gt, w, padding=5, groups=3)
    s1 = F.conv2d(x * x, w, padding=5, groups=3) - mu1.pow(2)
    s2 = F.conv2d(gt * gt, w, padding=5, groups=3) - mu2.pow(2)
    s12 = F.conv2d(x * gt, w, padding=5, groups=3) - mu1 * mu2
    m = ((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1.pow(2) + mu2.pow(2) + 1e-4) * (s1 + s2 + 9e-4))
    return al * torch.sqrt((x - gt).abs() + 1e-6), be * torch.sqrt((1.0 - m).abs() + 1e-6)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--views", type=int, default=16)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--no-torch", action="store_true", help="skip the torch composition legs")
    a = ap.parse_args()
    B, H, W = a.views, a.height, a.width
    gen = torch.Generator(device="cuda").manual_seed(1)
    img = torch.rand((B, 3, H, W), device="cuda", generator=gen)
    gt = torch.rand((B, 3, H, W), device="cuda", generator=gen)
    masks = (torch.rand((B, 1, H, W), device="cuda", generator=gen) > 0.1).float()
    n = 3 * H * W
    al = torch.full((B, 1, 1, 1), math.sqrt(0.8 / n), device="cuda")
    be = torch.full((B, 1, 1, 1), math.sqrt(0.2 / n), device="cuda")
    t, u1, u2 = (torch.randn((B, 3, H, W), device="cuda", generator=gen) for _ in range(3))
    lb = C.LossBatch(img, masks, gt, al, be)
    lb.blocks()
    elems = B * n
    mb = 4.0 / 3.0
    out = {"views": B, "H": H, "W": W}
    for name, fn, nbytes in (("blocks", lambda: lb.blocks(), 32 + mb), ("jvp", lambda: lb.jvp(t), 36 + mb),
                             ("vjp", lambda: lb.vjp(u1, u2), 36 + mb)):
        ms = timed(fn, a.iters)
        out[f"ours_{name}_ms"] = ms
        out[f"ours_{name}_GBps"] = elems * nbytes / ms / 1e6

    def ref_fwd():
        with torch.no_grad():
            reference_composition(img, masks, gt, al, be)

    def ref_jvp():
        with torch.no_grad(), fwAD.dual_level():
            reference_composition(fwAD.make_dual(img, t), masks, gt, al, be)

    def ref_vjp():
        x = img.detach().requires_grad_(True)
        r1, r2 = reference_composition(x, masks, gt, al, be)
        torch.autograd.backward((r1, r2), (u1, u2))

    for name, fn in (() if a.no_torch else (("blocks", ref_fwd), ("jvp", ref_jvp), ("vjp", ref_vjp))):
        out[f"torch_{name}_ms"] = timed(fn, max(3, a.iters // 3), warm=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
