#!/usr/bin/env python
"""Config E of BASELINE.json: rasterizer forward / backward microbenchmark (ms per view) at 1080p, ours against the
REFERENCE'S OWN kernels (oracle/_ref/libgsref.so: unmodified forward.cu / backward.cu / rasterizer_impl.cu built for
sm_100a) on the same box, same inputs.  The reference's JVP kernels cannot be built, so the J v and
fused J^T J v lines have no reference column.  Test / bench infrastructure: this is one of the places allowed to load
oracle/.

    python tools/bench_render.py [--points 100000,1000000] [--degrees 0,3] [--aa 0,1] [--iters 5]
Prints one JSON line per case and a table."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("gaussian-splatting-lm_b200", "tests", "oracle"):
    sys.path.insert(0, os.path.join(ROOT, p))

import torch  # noqa: E402

import gs_ref  # noqa: E402
import helpers as h  # noqa: E402
import scene_synth as ss  # noqa: E402
from diff_gaussian_rasterization import _C as C  # noqa: E402


def timed(fn, iters, warm=2):
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", default="100000,1000000,3000000")
    ap.add_argument("--degrees", default="0,3")
    ap.add_argument("--aa", default="0,1")
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--tile-culling", action="store_true",
                    help="GslmCamera::tile_culling = 1 (same image / cache bits, shorter lists) for our forward pass")
    a = ap.parse_args()
    C.TILE_CULLING = a.tile_culling
    rows = []
    have_ref = gs_ref.available()
    for P in [int(x) for x in a.points.split(",")]:
        for deg in [int(x) for x in a.degrees.split(",")]:
            for aa in [bool(int(x)) for x in a.aa.split(",")]:
                g = ss.make_gaussians(P, sh_degree=deg, seed=1004)
                cam = ss.hemisphere_rig(4, a.width, a.height)[1]
                gd, ct = h.to_dev(g), h.cam_tensors(cam)
                t = h.to_dev(ss.random_tangent(g, seed=7))
                fwd = h.c_forward(C, gd, ct, antialiasing=aa)
                R, color, radii, geom, binning, img, invd = fwd
                u = torch.randn_like(color)
                row = {"P": P, "sh_degree": deg, "antialiasing": aa, "tile_culling": a.tile_culling, "W": a.width,
                       "H": a.height, "R": int(R),
                       "Rc": int(R.cached)}
                row["ours_fwd_ms"] = timed(lambda: h.c_forward(C, gd, ct, antialiasing=aa), a.iters)
                # the same forward through gslm_forward_batch (what LinearSolverFunctions.linearize issues): 16 views of
                # the rig per call, ms per view -- no per-view host synchronisation, views on 4 internal streams
                rig = ss.hemisphere_rig(16, a.width, a.height)
                e = torch.empty(0, device="cuda")
                views = []
                for cm in rig:
                    c2 = h.cam_tensors(cm)
                    views.append((c2["viewmatrix"], c2["projmatrix"], c2["tan_fovx"], c2["tan_fovy"], c2["H"], c2["W"], c2["campos"]))
                fb = lambda: C.forward_batch(ct["bg"], gd["means3D"], e, gd["opacities"], gd["scales"], gd["rotations"], 1.0,
                                             e, views, gd["shs"], deg, aa, False, tile_culling=a.tile_culling)
                row["ours_fwd_batch_ms_per_view"] = timed(fb, max(2, a.iters // 2)) / len(views)
                row["ours_bwd_ms"] = timed(lambda: h.c_backward(C, gd, ct, fwd, u, antialiasing=aa), a.iters)
                ga, ta, ca = h.gauss_args(gd), h.tangent_args(t), h.cam_args(ct, antialiasing=aa)
                row["ours_jvp_cached_ms"] = timed(lambda: C.jvp_cached(ga, ta, ca, radii, geom, R, binning, img), a.iters)
                M = g["shs"].shape[1]
                z = lambda *s: torch.zeros(s, device="cuda")
                outs = (None, None, z(P, 1), z(P, 3), None, z(P, M, 3), z(P, 3), z(P, 4))
                plan = C.JtjvBatchPlan([(ca, radii, geom, R, binning, img, None)], torch.device("cuda"))
                row["ours_jtjv_cached_ms"] = timed(lambda: C.jtjv_batch(ga, ta, plan, outs), a.iters)
                if have_ref:
                    gr = dict(gd)
                    row["ref_fwd_ms"] = timed(lambda: gs_ref.forward(gr, ct, antialiasing=aa), a.iters)
                    Rr, cr, ir, rr = gs_ref.forward(gr, ct, antialiasing=aa)
                    row["ref_bwd_ms"] = timed(lambda: gs_ref.backward(gr, ct, rr, u, antialiasing=aa), a.iters)
                    if gs_ref.has_jvp():
                        # the reference's own forward-mode pass (forwardJvp = primal + tangent in one sweep, built with
                        # the functional float_grad.h stand-in of oracle/ref_shim): what its matvec costs per view
                        tg = {k: t[k] for k in ("means3D", "opacities", "scales", "rotations", "shs")}
                        row["ref_fwd_jvp_ms"] = timed(lambda: gs_ref.forward_jvp(gr, ct, tg=tg, antialiasing=aa), max(2, a.iters // 2))
                        row["ref_jtjv_composed_ms"] = row["ref_fwd_jvp_ms"] + row["ref_bwd_ms"]
                        row["jtjv_speedup"] = row["ref_jtjv_composed_ms"] / row["ours_jtjv_cached_ms"]
                    row["fwd_speedup"] = row["ref_fwd_ms"] / row["ours_fwd_ms"]
                    row["fwd_batch_speedup"] = row["ref_fwd_ms"] / row["ours_fwd_batch_ms_per_view"]
                    row["bwd_speedup"] = row["ref_bwd_ms"] / row["ours_bwd_ms"]
                    row["fwd_bwd_speedup"] = (row["ref_fwd_ms"] + row["ref_bwd_ms"]) / (row["ours_fwd_ms"] + row["ours_bwd_ms"])
                print(json.dumps(row), flush=True)
                rows.append(row)
                del fwd, geom, binning, img, plan, outs
                torch.cuda.empty_cache()
    print(f"\n{'P':>9} {'deg':>3} {'aa':>2} {'R':>10} {'Rc':>9} | {'fwd':>7} {'fwd/16':>7} {'bwd':>7} {'Jv':>7} {'JtJv':>7} | {'ref fwd':>8} {'ref bwd':>8} | fwd+bwd  fwd(batch) speed-up | ref fwdJvp  JtJv speed-up")
    for r in rows:
        print(f"{r['P']:9d} {r['sh_degree']:3d} {int(r['antialiasing']):2d} {r['R']:10d} {r['Rc']:9d} | "
              f"{r['ours_fwd_ms']:7.3f} {r['ours_fwd_batch_ms_per_view']:7.3f} {r['ours_bwd_ms']:7.3f} {r['ours_jvp_cached_ms']:7.3f} {r['ours_jtjv_cached_ms']:7.3f} | "
              f"{r.get('ref_fwd_ms', float('nan')):8.3f} {r.get('ref_bwd_ms', float('nan')):8.3f} | {r.get('fwd_bwd_speedup', float('nan')):5.2f}x   {r.get('fwd_batch_speedup', float('nan')):5.2f}x | {r.get('ref_fwd_jvp_ms', float('nan')):8.3f}   {r.get('jtjv_speedup', float('nan')):6.1f}x")


if __name__ == "__main__":
    main()
