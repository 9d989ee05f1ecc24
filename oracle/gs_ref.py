"""ctypes front-end of oracle/_ref/libgsref.so -- the REFERENCE's own primal / VJP / binning CUDA kernels
(unmodified forward.cu, backward.cu, rasterizer_impl.cu) built for sm_100a by oracle/Makefile (target `ref`).

TEST INFRASTRUCTURE ONLY: GPU-side parity oracle, bit-exactness oracle for the tile sort / ranges, and the timing
baseline "kernel to beat".  Only tests/ and bench scripts import it; the product never does.  Needs a GPU and the
prebuilt library (it cannot be rebuilt on the GPU box: /root/reference does not exist there)."""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
PATH = os.path.join(_HERE, "_ref", "libgsref.so")
_LIB = None


def available():
    return os.path.exists(PATH)


def variant(name):
    """A second instance of this module bound to oracle/_ref/libgsref_<name>.so, or None when that library was not
    built.  "shfix": the reference with the one-line SH-coefficient indexing repair in its forward-mode colour
    evaluation (oracle/Makefile explains it), the oracle for JVPs at SH degree >= 1."""
    import importlib.util
    path = os.path.join(_HERE, "_ref", f"libgsref_{name}.so")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location(f"gs_ref_{name}", os.path.abspath(__file__))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    m.PATH = path
    return m


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(PATH)
        _LIB.gsref_last_error.restype = C.c_char_p
        vp, f, i = C.c_void_p, C.c_float, C.c_int
        _LIB.gsref_forward.argtypes = [i, i, i, vp, i, i, vp, vp, vp, vp, vp, f, vp, vp, vp, vp, vp, f, f, i, i, vp, vp,
                                       vp, i]
        _LIB.gsref_backward.argtypes = [i, i, i, vp, i, i, vp, vp, vp, vp, vp, f, vp, vp, vp, vp, vp, f, f, vp, vp, vp,
                                        vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, i]
        _LIB.gsref_state.argtypes = [vp] * 8
    return _LIB


def _p(t):
    return None if t is None or t.numel() == 0 else t.data_ptr()


def forward(g, cam, antialiasing=False, scale_modifier=1.0):
    """g: dict of CUDA tensors (means3D, opacities, scales, rotations, shs, sh_degree[, colors_precomp,
    cov3D_precomp]); cam: dict(bg, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, H, W).
    Returns (R, color[3,H,W], invdepth[1,H,W], radii[P])."""
    L = lib()
    dev = g["means3D"].device
    P = g["means3D"].shape[0]
    shs = g.get("shs")
    M = shs.shape[1] if shs is not None and shs.numel() else 0
    H, W = cam["H"], cam["W"]
    color = torch.zeros(3, H, W, device=dev)
    invd = torch.zeros(1, H, W, device=dev)
    radii = torch.zeros(P, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    R = L.gsref_forward(P, int(g.get("sh_degree", 0)), M, _p(cam["bg"]), W, H, _p(g["means3D"]), _p(shs),
                        _p(g.get("colors_precomp")), _p(g["opacities"]), _p(g.get("scales")), scale_modifier,
                        _p(g.get("rotations")), _p(g.get("cov3D_precomp")), _p(cam["viewmatrix"]),
                        _p(cam["projmatrix"]), _p(cam["campos"]), cam["tan_fovx"], cam["tan_fovy"], 0,
                        int(antialiasing), _p(color), _p(invd), _p(radii), 0)
    if R < 0:
        raise RuntimeError("gsref_forward: " + L.gsref_last_error().decode())
    return R, color, invd, radii


def state(P, W, H, R, dev="cuda"):
    T = ((W + 15) // 16) * ((H + 15) // 16)
    out = {"point_list": torch.zeros(max(R, 1), dtype=torch.int32, device=dev)[:R],
           "ranges": torch.zeros(T, 2, dtype=torch.int32, device=dev),
           "n_contrib": torch.zeros(H, W, dtype=torch.int32, device=dev),
           "final_T": torch.zeros(H, W, device=dev), "tiles_touched": torch.zeros(P, dtype=torch.int32, device=dev),
           "point_offsets": torch.zeros(P, dtype=torch.int32, device=dev), "depths": torch.zeros(P, device=dev),
           "means2D": torch.zeros(P, 2, device=dev)}
    rc = lib().gsref_state(_p(out["point_list"]), _p(out["ranges"]), _p(out["n_contrib"]), _p(out["final_T"]),
                           _p(out["tiles_touched"]), _p(out["point_offsets"]), _p(out["depths"]), _p(out["means2D"]))
    if rc:
        raise RuntimeError("gsref_state: " + lib().gsref_last_error().decode())
    return out


def backward(g, cam, radii, dL_dcolor, dL_dinvdepth=None, antialiasing=False, scale_modifier=1.0):
    """Gradients of the LAST forward call. Returns the dict of rasterize_points.cu:300-399's 8-tuple."""
    L = lib()
    dev = g["means3D"].device
    P = g["means3D"].shape[0]
    shs = g.get("shs")
    M = shs.shape[1] if shs is not None and shs.numel() else 0
    H, W = cam["H"], cam["W"]
    z = lambda *s: torch.zeros(s, device=dev)
    o = {"dL_dmeans2D": z(P, 3), "dL_dconic": z(P, 2, 2), "dL_dopacity": z(P, 1), "dL_dcolors": z(P, 3),
         "dL_dinvdepth": z(P, 1), "dL_dmeans3D": z(P, 3), "dL_dcov3D": z(P, 6), "dL_dsh": z(P, max(M, 1), 3),
         "dL_dscales": z(P, 3), "dL_drotations": z(P, 4)}
    dcol = dL_dcolor.contiguous()
    dinv = dL_dinvdepth.contiguous() if dL_dinvdepth is not None else None
    torch.cuda.synchronize()
    rc = L.gsref_backward(P, int(g.get("sh_degree", 0)), M, _p(cam["bg"]), W, H, _p(g["means3D"]), _p(shs),
                          _p(g.get("colors_precomp")), _p(g["opacities"]), _p(g.get("scales")), scale_modifier,
                          _p(g.get("rotations")), _p(g.get("cov3D_precomp")), _p(cam["viewmatrix"]),
                          _p(cam["projmatrix"]), _p(cam["campos"]), cam["tan_fovx"], cam["tan_fovy"], _p(radii),
                          _p(dcol), _p(dinv), _p(o["dL_dmeans2D"]), _p(o["dL_dconic"]), _p(o["dL_dopacity"]),
                          _p(o["dL_dcolors"]),
                          # like rasterize_points.cu:348-357: no per-Gaussian inverse-depth sink without an inverse-depth
                          # cotangent (the kernel keys its whole inverse-depth branch on this pointer, backward.cu:597)
                          _p(o["dL_dinvdepth"]) if dinv is not None else None, _p(o["dL_dmeans3D"]), _p(o["dL_dcov3D"]),
                          _p(o["dL_dsh"]), _p(o["dL_dscales"]), _p(o["dL_drotations"]), int(antialiasing), 0)
    if rc:
        raise RuntimeError("gsref_backward: " + L.gsref_last_error().decode())
    torch.cuda.synchronize()
    if M == 0:
        o["dL_dsh"] = z(P, 0, 3)
    return o


def has_jvp():
    """True when the library was built with the reference's forward-mode kernels (functional float_grad.h stand-in)."""
    L = lib()
    return hasattr(L, "gsref_has_jvp") and bool(L.gsref_has_jvp())


def forward_jvp(g, cam, tg=None, tcam=None, antialiasing=False, scale_modifier=1.0):
    """The reference's own forwardJvp (preprocessCUDAJvp + renderCUDAJvpSplit), driven like
    RasterizeGaussiansJvpCUDA (rasterize_points.cu:143-298).  tg: dict of Gaussian-parameter tangents (CUDA tensors,
    same keys as g; missing = zero), tcam: dict of camera-side tangents (bg, viewmatrix, projmatrix, campos tensors;
    tan_fovx, tan_fovy, scale_modifier floats).  Returns dict(R, color, invdepth, radii, color_t, invdepth_t) plus
    the per-Gaussian tangent records t_depths, t_means2D, t_conic_opacity, t_rgb and primal conic_opacity, rgb."""
    L = lib()
    vp, f, i = C.c_void_p, C.c_float, C.c_int
    L.gsref_forward_jvp.argtypes = [i, i, i, vp, vp, i, i, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, f, f, vp, vp, vp,
                                    vp, vp, vp, vp, vp, vp, vp, f, f, f, f, i, i, vp, vp, vp, vp, vp, i]
    L.gsref_jvp_state.argtypes = [vp] * 7
    tg, tcam = tg or {}, tcam or {}
    dev = g["means3D"].device
    P = g["means3D"].shape[0]
    shs = g.get("shs")
    M = shs.shape[1] if shs is not None and shs.numel() else 0
    H, W = cam["H"], cam["W"]
    keep = []

    def pair(x, t):
        """(value ptr, tangent ptr): a zero tangent of the same size when none is given; (None, None) for absent inputs"""
        if x is None or x.numel() == 0:
            return None, None
        x = x.contiguous().float()
        t = torch.zeros_like(x) if t is None or t.numel() == 0 else t.contiguous().float().reshape(x.shape)
        keep.extend((x, t))
        return x.data_ptr(), t.data_ptr()

    out = {"color": torch.zeros(3, H, W, device=dev), "color_t": torch.zeros(3, H, W, device=dev),
           "invdepth": torch.zeros(1, H, W, device=dev), "invdepth_t": torch.zeros(1, H, W, device=dev),
           "radii": torch.zeros(P, dtype=torch.int32, device=dev)}
    bg, m3, sh = pair(cam["bg"], tcam.get("bg")), pair(g["means3D"], tg.get("means3D")), pair(shs, tg.get("shs"))
    col, op = pair(g.get("colors_precomp"), tg.get("colors_precomp")), pair(g["opacities"], tg.get("opacities"))
    sc, rot = pair(g.get("scales"), tg.get("scales")), pair(g.get("rotations"), tg.get("rotations"))
    cov = pair(g.get("cov3D_precomp"), tg.get("cov3D_precomp"))
    vm, pm = pair(cam["viewmatrix"], tcam.get("viewmatrix")), pair(cam["projmatrix"], tcam.get("projmatrix"))
    cp = pair(cam["campos"], tcam.get("campos"))
    torch.cuda.synchronize()
    R = L.gsref_forward_jvp(P, int(g.get("sh_degree", 0)), M, *bg, W, H, *m3, *sh, *col, *op, *sc, scale_modifier,
                            float(tcam.get("scale_modifier", 0.0)), *rot, *cov, *vm, *pm, *cp, cam["tan_fovx"],
                            float(tcam.get("tan_fovx", 0.0)), cam["tan_fovy"], float(tcam.get("tan_fovy", 0.0)), 0,
                            int(antialiasing), _p(out["color"]), _p(out["color_t"]), _p(out["invdepth"]),
                            _p(out["invdepth_t"]), _p(out["radii"]), 0)
    if R < 0:
        raise RuntimeError("gsref_forward_jvp: " + L.gsref_last_error().decode())
    out["R"] = R
    z = lambda *s: torch.zeros(s, device=dev)
    out.update(t_depths=z(P), t_means2D=z(P, 2), t_cov3D=z(P, 6), t_conic_opacity=z(P, 4), t_rgb=z(P, 3),
               conic_opacity=z(P, 4), rgb=z(P, 3))
    if L.gsref_jvp_state(_p(out["t_depths"]), _p(out["t_means2D"]), _p(out["t_cov3D"]), _p(out["t_conic_opacity"]),
                         _p(out["t_rgb"]), _p(out["conic_opacity"]), _p(out["rgb"])):
        raise RuntimeError("gsref_jvp_state: " + L.gsref_last_error().decode())
    out["depths"] = state(P, W, H, R, dev)["depths"]
    torch.cuda.synchronize()
    return out
