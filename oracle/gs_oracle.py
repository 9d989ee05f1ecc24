"""ctypes front-end of the CPU oracle (oracle/libgs_oracle.so).

TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module; nothing under
gaussian-splatting-lm_b200/ (the product) does.

The numerics live in gs_oracle.hpp / gs_oracle.cpp (each function cites the
reference file:line it restates).  This file only marshals numpy arrays.

Parity status (see gs_oracle.hpp): PINNED for preprocess primal, preprocess JVP
with camera tangents and binning by the reference's own golden vectors
(tests/golden/dgr_fixture.npz, made by tests/golden/make_golden.py);
images / gradients / parameter tangents / J^T J v are "parity unpinned"
(self-consistency only: FD, adjoint identity, float64 instantiation).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    so = os.path.join(_HERE, "libgs_oracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("gs_oracle.cpp", "gs_oracle.hpp")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "libgs_oracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.gso_higher_msb.restype = C.c_uint32
        _LIB.gso_scan.restype = C.c_uint32
        _LIB.gso_num_threads.restype = C.c_int
    return _LIB


def num_threads():
    return lib().gso_num_threads()


def set_num_threads(n):
    lib().gso_set_num_threads(C.c_int(int(n)))


def _fp(dt):
    return C.c_float if dt == np.float32 else C.c_double


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _arr(x, dt, shape=None):
    if x is None:
        return None
    a = np.ascontiguousarray(np.asarray(x, dtype=dt))
    if shape is not None:
        a = a.reshape(shape)
    return a


def _make_struct(fields):
    class S(C.Structure):
        _fields_ = fields
    return S


def _pre_struct(dt):
    f = _fp(dt)
    vp = C.c_void_p
    return _make_struct(
        [(n, C.c_int) for n in ("P", "D", "M", "W", "H", "antialiasing", "jvp")]
        + [(n, vp) for n in ("means3D", "scales", "rotations", "opacities", "shs", "cov3D_precomp", "colors_precomp",
                             "viewmatrix", "projmatrix", "campos")]
        + [(n, f) for n in ("scale_modifier", "tan_fovx", "tan_fovy")]
        + [(n, vp) for n in ("t_means3D", "t_scales", "t_rotations", "t_opacities", "t_shs", "t_cov3D_precomp",
                             "t_colors_precomp", "t_viewmatrix", "t_projmatrix", "t_campos")]
        + [(n, f) for n in ("t_scale_modifier", "t_tan_fovx", "t_tan_fovy")]
        + [(n, vp) for n in ("radii", "tiles_touched", "xy", "depths", "cov3D", "rgb", "conic_opacity", "clamped",
                             "t_xy", "t_depths", "t_cov3D", "t_rgb", "t_conic_opacity")])


def preprocess(means3D, opacities, viewmatrix, projmatrix, campos, W, H, tan_fovx, tan_fovy, *, scales=None,
               rotations=None, shs=None, sh_degree=0, cov3D_precomp=None, colors_precomp=None, scale_modifier=1.0,
               antialiasing=False, tangents=None, dtype=np.float32):
    """forward.cu:154-273 (+ forward_impl.h:208-367 when `tangents` is a dict).

    tangents keys (all optional): means3D, scales, rotations, opacities, shs, cov3D_precomp, colors_precomp,
    viewmatrix, projmatrix, campos, scale_modifier, tan_fovx, tan_fovy.
    Returns a dict of per-Gaussian arrays; tangent outputs are prefixed `t_`.
    """
    dt = np.dtype(dtype).type
    means3D = _arr(means3D, dt)
    P = means3D.shape[0]
    M = 0 if shs is None else np.asarray(shs).reshape(P, -1, 3).shape[1]
    S = _pre_struct(dt)
    a = S()
    keep = []

    def put(name, x, shape=None):
        arr = _arr(x, dt, shape)
        keep.append(arr)
        setattr(a, name, _ptr(arr))

    a.P, a.D, a.M, a.W, a.H = P, int(sh_degree), M, int(W), int(H)
    a.antialiasing = int(bool(antialiasing))
    a.jvp = int(tangents is not None)
    put("means3D", means3D)
    put("scales", scales)
    put("rotations", rotations)
    put("opacities", opacities, (P,))
    put("shs", shs)
    put("cov3D_precomp", cov3D_precomp)
    put("colors_precomp", colors_precomp)
    put("viewmatrix", viewmatrix, (16,))
    put("projmatrix", projmatrix, (16,))
    put("campos", np.asarray(campos).reshape(-1)[:3])
    a.scale_modifier, a.tan_fovx, a.tan_fovy = float(scale_modifier), float(tan_fovx), float(tan_fovy)
    t = tangents or {}
    put("t_means3D", t.get("means3D"))
    put("t_scales", t.get("scales"))
    put("t_rotations", t.get("rotations"))
    put("t_opacities", t.get("opacities"))
    put("t_shs", t.get("shs"))
    put("t_cov3D_precomp", t.get("cov3D_precomp"))
    put("t_colors_precomp", t.get("colors_precomp"))
    put("t_viewmatrix", t.get("viewmatrix"))
    put("t_projmatrix", t.get("projmatrix"))
    tc = t.get("campos")
    put("t_campos", None if tc is None else np.asarray(tc).reshape(-1)[:3])
    a.t_scale_modifier = float(t.get("scale_modifier", 0.0))
    a.t_tan_fovx = float(t.get("tan_fovx", 0.0))
    a.t_tan_fovy = float(t.get("tan_fovy", 0.0))
    out = {
        "radii": np.zeros(P, np.int32), "tiles_touched": np.zeros(P, np.uint32),
        "xy": np.zeros((P, 2), dt), "depths": np.zeros(P, dt), "cov3D": np.zeros((P, 6), dt),
        "rgb": np.zeros((P, 3), dt), "conic_opacity": np.zeros((P, 4), dt), "clamped": np.zeros((P, 3), np.uint8),
    }
    if tangents is not None:
        for k in ("xy", "depths", "cov3D", "rgb", "conic_opacity"):
            out["t_" + k] = np.zeros_like(out[k])
    for k, v in out.items():
        setattr(a, k, _ptr(v))
    fn = lib().gso_preprocess_f32 if dt == np.float32 else lib().gso_preprocess_f64
    fn(C.byref(a))
    return out


def mark_visible(means3D, viewmatrix):
    m = _arr(means3D, np.float32)
    V = _arr(viewmatrix, np.float32, (16,))
    out = np.zeros(m.shape[0], np.uint8)
    lib().gso_mark_visible_f32(C.c_int(m.shape[0]), _ptr(m), _ptr(V), _ptr(out))
    return out.astype(bool)


def higher_msb(n):
    return int(lib().gso_higher_msb(C.c_uint32(int(n))))


def binning(xy, depths, radii, tiles_touched, W, H):
    """rasterizer_impl.cu:321-364. Returns dict(point_offsets, R, keys, point_list, ranges[T,2])."""
    xy = _arr(xy, np.float32)
    depths = _arr(depths, np.float32)
    radii = _arr(radii, np.int32)
    tt = _arr(tiles_touched, np.uint32)
    P = radii.shape[0]
    off = np.zeros(P, np.uint32)
    R = int(lib().gso_scan(C.c_int(P), _ptr(tt), _ptr(off)))
    gx, gy = (W + 15) // 16, (H + 15) // 16
    keys = np.zeros(R, np.uint64)
    pl = np.zeros(R, np.uint32)
    ranges = np.zeros((gx * gy, 2), np.uint32)
    lib().gso_binning(C.c_int(P), C.c_int(W), C.c_int(H), _ptr(xy), _ptr(depths), _ptr(radii), _ptr(off),
                      C.c_uint32(R), _ptr(keys), _ptr(pl), _ptr(ranges))
    return {"point_offsets": off, "R": R, "keys": keys, "point_list": pl, "ranges": ranges}


def _blend_struct(dt):
    vp = C.c_void_p
    return _make_struct([(n, C.c_int) for n in ("W", "H", "jvp")] + [(n, vp) for n in (
        "ranges", "point_list", "xy", "rgb", "conic_opacity", "depths", "bg",
        "t_xy", "t_rgb", "t_conic_opacity", "t_depths", "t_bg",
        "out_color", "out_invdepth", "final_T", "n_contrib", "t_out_color", "t_out_invdepth", "t_final_T")])


def blend(ranges, point_list, xy, rgb, conic_opacity, depths, bg, W, H, tangents=None, dtype=np.float32):
    """forward.cu:278-402 (+ forward_impl.h:373-539 with `tangents` = dict(xy, rgb, conic_opacity, depths, bg))."""
    dt = np.dtype(dtype).type
    a = _blend_struct(dt)()
    keep = []

    def put(name, x, d=dt):
        arr = _arr(x, d)
        keep.append(arr)
        setattr(a, name, _ptr(arr))

    a.W, a.H, a.jvp = int(W), int(H), int(tangents is not None)
    put("ranges", ranges, np.uint32)
    put("point_list", point_list, np.uint32)
    put("xy", xy)
    put("rgb", rgb)
    put("conic_opacity", conic_opacity)
    put("depths", depths)
    put("bg", bg)
    t = tangents or {}
    for k in ("xy", "rgb", "conic_opacity", "depths", "bg"):
        put("t_" + k, t.get(k))
    out = {"out_color": np.zeros((3, H, W), dt), "out_invdepth": np.zeros((1, H, W), dt),
           "final_T": np.zeros((H, W), dt), "n_contrib": np.zeros((H, W), np.uint32)}
    if tangents is not None:
        out["t_out_color"] = np.zeros((3, H, W), dt)
        out["t_out_invdepth"] = np.zeros((1, H, W), dt)
        out["t_final_T"] = np.zeros((H, W), dt)
    for k, v in out.items():
        setattr(a, k, _ptr(v))
    fn = lib().gso_blend_f32 if dt == np.float32 else lib().gso_blend_f64
    fn(C.byref(a))
    return out


def _blend_bwd_struct():
    vp = C.c_void_p
    return _make_struct([(n, C.c_int) for n in ("P", "W", "H")] + [(n, vp) for n in (
        "ranges", "point_list", "bg", "xy", "conic_opacity", "rgb", "depths", "final_T", "n_contrib",
        "dL_dpix", "dL_dinvdepth", "dL_dmean2D", "dL_dconic", "dL_dopacity", "dL_dcolors", "dL_dinvdepths")])


def blend_bwd(ranges, point_list, bg, xy, conic_opacity, rgb, depths, final_T, n_contrib, dL_dpix, dL_dinvdepth,
              W, H, dtype=np.float32):
    """backward.cu:452-638."""
    dt = np.dtype(dtype).type
    P = np.asarray(xy).shape[0]
    a = _blend_bwd_struct()()
    keep = []

    def put(name, x, d=dt):
        arr = _arr(x, d)
        keep.append(arr)
        setattr(a, name, _ptr(arr))

    a.P, a.W, a.H = P, int(W), int(H)
    put("ranges", ranges, np.uint32)
    put("point_list", point_list, np.uint32)
    put("bg", bg)
    put("xy", xy)
    put("conic_opacity", conic_opacity)
    put("rgb", rgb)
    put("depths", depths)
    put("final_T", final_T)
    put("n_contrib", n_contrib, np.uint32)
    put("dL_dpix", dL_dpix)
    put("dL_dinvdepth", dL_dinvdepth)
    out = {"dL_dmean2D": np.zeros((P, 3), dt), "dL_dconic": np.zeros((P, 4), dt), "dL_dopacity": np.zeros(P, dt),
           "dL_dcolors": np.zeros((P, 3), dt), "dL_dinvdepths": np.zeros(P, dt)}
    for k, v in out.items():
        setattr(a, k, _ptr(v))
    fn = lib().gso_blend_bwd_f32 if dt == np.float32 else lib().gso_blend_bwd_f64
    fn(C.byref(a))
    return out


def _pre_bwd_struct(dt):
    f = _fp(dt)
    vp = C.c_void_p
    return _make_struct(
        [(n, C.c_int) for n in ("P", "D", "M", "W", "H", "antialiasing")]
        + [(n, vp) for n in ("means3D", "radii", "shs", "clamped", "opacities", "scales", "rotations")]
        + [("scale_modifier", f), ("cov3D", vp), ("viewmatrix", vp), ("projmatrix", vp), ("campos", vp),
           ("tan_fovx", f), ("tan_fovy", f)]
        + [(n, vp) for n in ("dL_dmean2D", "dL_dconic", "dL_dinvdepth", "dL_dopacity", "dL_dcolor", "dL_dmean3D",
                             "dL_dcov3D", "dL_dsh", "dL_dscale", "dL_drot")])


def preprocess_bwd(means3D, radii, shs, sh_degree, clamped, opacities, scales, rotations, scale_modifier, cov3D,
                   viewmatrix, projmatrix, campos, W, H, tan_fovx, tan_fovy, dL_dmean2D, dL_dconic, dL_dinvdepth,
                   dL_dopacity, dL_dcolor, antialiasing=False, dtype=np.float32):
    """backward.cu:147-326,398-449,23-142,330-393.  Returns dict of grads (dL_dopacity is rescaled when AA)."""
    dt = np.dtype(dtype).type
    means3D = _arr(means3D, dt)
    P = means3D.shape[0]
    M = 0 if shs is None else np.asarray(shs).reshape(P, -1, 3).shape[1]
    a = _pre_bwd_struct(dt)()
    keep = []

    def put(name, x, d=dt, shape=None):
        arr = _arr(x, d, shape)
        keep.append(arr)
        setattr(a, name, _ptr(arr))
        return arr

    a.P, a.D, a.M, a.W, a.H, a.antialiasing = P, int(sh_degree), M, int(W), int(H), int(bool(antialiasing))
    put("means3D", means3D)
    put("radii", radii, np.int32)
    put("shs", shs)
    put("clamped", clamped, np.uint8)
    put("opacities", opacities, dt, (P,))
    put("scales", scales)
    put("rotations", rotations)
    a.scale_modifier = float(scale_modifier)
    put("cov3D", cov3D)
    put("viewmatrix", viewmatrix, dt, (16,))
    put("projmatrix", projmatrix, dt, (16,))
    put("campos", np.asarray(campos).reshape(-1)[:3])
    a.tan_fovx, a.tan_fovy = float(tan_fovx), float(tan_fovy)
    put("dL_dmean2D", dL_dmean2D)
    put("dL_dconic", dL_dconic)
    put("dL_dinvdepth", dL_dinvdepth)
    dop = np.array(np.asarray(dL_dopacity, dtype=dt).reshape(P), copy=True)
    a.dL_dopacity = _ptr(dop)
    put("dL_dcolor", dL_dcolor)
    out = {"dL_dmean3D": np.zeros((P, 3), dt), "dL_dcov3D": np.zeros((P, 6), dt),
           "dL_dsh": np.zeros((P, max(M, 1), 3), dt), "dL_dscale": np.zeros((P, 3), dt),
           "dL_drot": np.zeros((P, 4), dt)}
    for k, v in out.items():
        setattr(a, k, _ptr(v))
    fn = lib().gso_preprocess_bwd_f32 if dt == np.float32 else lib().gso_preprocess_bwd_f64
    fn(C.byref(a))
    out["dL_dopacity"] = dop
    if M == 0:
        out["dL_dsh"] = np.zeros((P, 0, 3), dt)
    return out


# ---------------------------------------------------------------------------
# Whole-pipeline compositions (rasterizer_impl.cu:239-391 forward,
# rasterizer_jvp_impl.h:125-325 forwardJvp, rasterizer_impl.cu:395-500 backward)
# ---------------------------------------------------------------------------
def render_forward(g, cam, tangents=None, dtype=np.float32):
    """g: dict(means3D, opacities, scales, rotations, shs, sh_degree[, cov3D_precomp, colors_precomp, scale_modifier])
    cam: dict(W, H, tan_fovx, tan_fovy, viewmatrix, projmatrix, campos, bg[, antialiasing])
    tangents: None or dict over Gaussian/camera inputs (+ 'bg').  Returns dict with color, invdepth, radii and the
    intermediate state needed by render_backward (and t_color / t_invdepth when tangents is given)."""
    W, H = int(cam["W"]), int(cam["H"])
    pre = preprocess(g["means3D"], g["opacities"], cam["viewmatrix"], cam["projmatrix"], cam["campos"], W, H,
                     cam["tan_fovx"], cam["tan_fovy"], scales=g.get("scales"), rotations=g.get("rotations"),
                     shs=g.get("shs"), sh_degree=g.get("sh_degree", 0), cov3D_precomp=g.get("cov3D_precomp"),
                     colors_precomp=g.get("colors_precomp"), scale_modifier=g.get("scale_modifier", 1.0),
                     antialiasing=cam.get("antialiasing", False), tangents=tangents, dtype=dtype)
    # binning always on the float32 view of the geometry (integer work)
    b = binning(pre["xy"].astype(np.float32), pre["depths"].astype(np.float32), pre["radii"], pre["tiles_touched"],
                W, H)
    bt = None
    if tangents is not None:
        bt = {"xy": pre["t_xy"], "rgb": pre["t_rgb"], "conic_opacity": pre["t_conic_opacity"],
              "depths": pre["t_depths"], "bg": tangents.get("bg")}
    bl = blend(b["ranges"], b["point_list"], pre["xy"], pre["rgb"], pre["conic_opacity"], pre["depths"], cam["bg"],
               W, H, tangents=bt, dtype=dtype)
    out = {"color": bl["out_color"], "invdepth": bl["out_invdepth"], "radii": pre["radii"], "pre": pre, "bin": b,
           "final_T": bl["final_T"], "n_contrib": bl["n_contrib"], "num_rendered": b["R"]}
    if tangents is not None:
        out["t_color"] = bl["t_out_color"]
        out["t_invdepth"] = bl["t_out_invdepth"]
    return out


def render_backward(g, cam, st, dL_dcolor, dL_dinvdepth=None, dtype=np.float32):
    """Returns dict(dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations)
    -- the 8-tuple of rasterize_points.cu:300-399 as a dict."""
    W, H = int(cam["W"]), int(cam["H"])
    pre, b = st["pre"], st["bin"]
    bb = blend_bwd(b["ranges"], b["point_list"], cam["bg"], pre["xy"], pre["conic_opacity"], pre["rgb"],
                   pre["depths"], st["final_T"], st["n_contrib"], dL_dcolor, dL_dinvdepth, W, H, dtype=dtype)
    has_sh = g.get("colors_precomp") is None
    has_sr = g.get("cov3D_precomp") is None
    pb = preprocess_bwd(g["means3D"], pre["radii"], g.get("shs") if has_sh else None, g.get("sh_degree", 0),
                        pre["clamped"], g["opacities"], g.get("scales") if has_sr else None,
                        g.get("rotations") if has_sr else None, g.get("scale_modifier", 1.0), pre["cov3D"],
                        cam["viewmatrix"], cam["projmatrix"], cam["campos"], W, H, cam["tan_fovx"], cam["tan_fovy"],
                        bb["dL_dmean2D"], bb["dL_dconic"], bb["dL_dinvdepths"] if dL_dinvdepth is not None else None,
                        bb["dL_dopacity"], bb["dL_dcolors"], antialiasing=cam.get("antialiasing", False), dtype=dtype)
    return {"dL_dmeans2D": bb["dL_dmean2D"], "dL_dcolors": bb["dL_dcolors"], "dL_dopacity": pb["dL_dopacity"],
            "dL_dmeans3D": pb["dL_dmean3D"], "dL_dcov3D": pb["dL_dcov3D"], "dL_dsh": pb["dL_dsh"],
            "dL_dscales": pb["dL_dscale"], "dL_drotations": pb["dL_drot"], "dL_dconic": bb["dL_dconic"]}
