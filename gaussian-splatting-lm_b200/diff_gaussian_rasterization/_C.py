"""ctypes binding of libgslm_b200.so (include/gslm_b200.h) with the call signatures of the
reference extension module `diff_gaussian_rasterization._C`
(DGR/ext.cpp:15-20, DGR/rasterize_points.cu).  Host plumbing only: torch supplies device
memory and the current stream; every kernel lives in the shared library.

There is NO fallback: if the library is missing or a tensor is not a CUDA tensor the call
raises.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(os.path.dirname(_HERE), "libgslm_b200.so")
_lib = None


class GslmGaussians(C.Structure):
    _fields_ = [("P", C.c_int32), ("sh_degree", C.c_int32), ("M", C.c_int32), ("scale_modifier", C.c_float),
                ("means3D", C.c_void_p), ("opacities", C.c_void_p), ("scales", C.c_void_p), ("rotations", C.c_void_p),
                ("cov3D_precomp", C.c_void_p), ("shs", C.c_void_p), ("colors_precomp", C.c_void_p)]


class GslmCamera(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("tan_fovx", C.c_float), ("tan_fovy", C.c_float),
                ("bg", C.c_void_p), ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p), ("campos", C.c_void_p),
                ("prefiltered", C.c_int32), ("antialiasing", C.c_int32), ("debug", C.c_int32),
                ("tile_culling", C.c_int32)]


class GslmCameraTangent(C.Structure):
    _fields_ = [("tan_fovx", C.c_float), ("tan_fovy", C.c_float), ("scale_modifier", C.c_float),
                ("bg", C.c_void_p), ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p), ("campos", C.c_void_p)]


class GslmGrads(C.Structure):
    _fields_ = [("dL_dmeans2D", C.c_void_p), ("dL_dcolors", C.c_void_p), ("dL_dopacity", C.c_void_p),
                ("dL_dmeans3D", C.c_void_p), ("dL_dcov3D", C.c_void_p), ("dL_dsh", C.c_void_p),
                ("dL_dscales", C.c_void_p), ("dL_drotations", C.c_void_p), ("accumulate", C.c_int32)]


class GslmLossBatch(C.Structure):
    _fields_ = [("B", C.c_int32), ("H", C.c_int32), ("W", C.c_int32), ("images", C.c_void_p),
                ("alpha_masks", C.c_void_p), ("gt", C.c_void_p), ("alphas", C.c_void_p), ("betas", C.c_void_p),
                ("tangent_masks", C.c_void_p)]


class GslmActivations(C.Structure):
    _fields_ = [("P", C.c_int32), ("M", C.c_int32), ("scales", C.c_void_p), ("rotations", C.c_void_p),
                ("rot_norm", C.c_void_p), ("opacities", C.c_void_p)]


class GslmRawVector(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("xyz", "features_dc", "features_rest", "scaling", "rotation", "opacity")]


class GslmActVector(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("means3D", "shs", "scales", "rotations", "opacities")]


class GslmResidualEpilogue(C.Structure):
    _fields_ = [("gt", C.c_void_p), ("alpha_mask", C.c_void_p), ("out_image", C.c_void_p), ("out_mask", C.c_void_p),
                ("out_weight", C.c_void_p), ("gt_plane", C.c_int64), ("gt_row", C.c_int64), ("alpha_row", C.c_int64),
                ("image_plane", C.c_int64), ("image_row", C.c_int64), ("mask_plane", C.c_int64), ("mask_row", C.c_int64),
                ("weight_scale", C.c_float), ("apply_alpha_mask", C.c_int32), ("subtract_gt", C.c_int32)]


class GslmBufferView(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("point_list", "tile_keys_sorted", "ranges", "n_contrib", "final_T",
                                          "tiles_touched", "point_offsets", "depths", "means2D", "conic_opacity",
                                          "rgb", "cov3D", "clamped", "n_contrib_c", "cstart", "records", "masks", "cid",
                                          "g_off", "g_list")]


class NumRendered(int):
    """The reference's `num_rendered` (R = sum of tiles_touched) as a plain int for API compatibility, carrying
    the size of the compact linearisation cache in `.cached` (Rc <= R) for the derivative entry points."""
    cached = 0

    def __new__(cls, rendered, cached):
        o = super().__new__(cls, rendered)
        o.cached = int(cached)
        return o


def _cached(R, imageBuffer=None, W=None, H=None):
    """Rc of a forward call: from the NumRendered it returned, else (a caller that rebuilt a plain int) read
    back from the image buffer (cstart[T]) -- one 4-byte device read."""
    if isinstance(R, NumRendered):
        return R.cached
    if imageBuffer is None or imageBuffer.numel() == 0:
        return 0
    v = GslmBufferView()
    _check(lib().gslm_buffer_view(0, int(W), int(H), 0, 0, None, None, imageBuffer.data_ptr(), None, C.addressof(v)))
    T = ((W + 15) // 16) * ((H + 15) // 16)
    off = v.cstart - imageBuffer.data_ptr() + 4 * T
    return int(imageBuffer[off:off + 4].view(torch.int32).item())


ALLOC_FN = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_int32, C.c_size_t)
SLICE_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_int32, C.c_int32, C.c_int32)

EXPORTS = ("gslm_last_error", "gslm_abi_version", "gslm_sizeof", "gslm_camera_offsetof", "gslm_geom_bytes", "gslm_image_bytes", "gslm_binning_bytes", "gslm_binning_temp_bytes",
           "gslm_mark_visible", "gslm_forward", "gslm_render", "gslm_render_batch", "gslm_forward_batch",
           "gslm_count_instances", "gslm_fp32_probe_scratch_bytes", "gslm_fp32_peak_probe", "gslm_forward_jvp", "gslm_backward_scratch_bytes", "gslm_backward",
           "gslm_jvp_scratch_bytes", "gslm_jvp_cached", "gslm_jvp_cached_noalloc", "gslm_jtjv_scratch_bytes", "gslm_jtjv",
           "gslm_jtjv_batch_scratch_bytes", "gslm_jtjv_batch", "gslm_jtjv_batch_sliced", "gslm_jvp_batch_scratch_bytes",
           "gslm_jvp_batch", "gslm_backward_batch_scratch_bytes", "gslm_backward_batch", "gslm_backward_batch_sliced",
           "gslm_activation_vjp_range", "gslm_loss_lin_bytes", "gslm_loss_blocks",
           "gslm_loss_blocks_jvp", "gslm_loss_blocks_vjp", "gslm_activation_jvp", "gslm_activation_vjp",
           "gslm_densification_stats", "gslm_multimem_allreduce_f32",
           "gslm_buffer_view", "gslm_tangent_records", "gslm_timing_num_stages",
           "gslm_timing_stage_name", "gslm_timing_enable", "gslm_timing_collect")


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise RuntimeError(f"{_LIB_PATH} not built: run `python __graft_entry__.py build` "
                               "(there is no CPU or PyTorch fallback for the rasterizer)")
        L = C.CDLL(_LIB_PATH)
        L.gslm_last_error.restype = C.c_char_p
        for n in ("gslm_geom_bytes", "gslm_image_bytes", "gslm_binning_bytes", "gslm_binning_temp_bytes",
                  "gslm_backward_scratch_bytes",
                  "gslm_jvp_scratch_bytes", "gslm_jtjv_scratch_bytes", "gslm_jtjv_batch_scratch_bytes",
                  "gslm_jvp_batch_scratch_bytes", "gslm_backward_batch_scratch_bytes"):
            getattr(L, n).restype = C.c_size_t
        L.gslm_geom_bytes.argtypes = [C.c_int32]
        L.gslm_image_bytes.argtypes = [C.c_int32, C.c_int32]
        L.gslm_binning_bytes.argtypes = [C.c_int64]
        L.gslm_binning_temp_bytes.argtypes = [C.c_int32, C.c_int64]
        L.gslm_backward_scratch_bytes.argtypes = [C.c_int64]
        L.gslm_jvp_scratch_bytes.argtypes = [C.c_int32]
        L.gslm_jtjv_scratch_bytes.argtypes = [C.c_int32, C.c_int64]
        vp = C.c_void_p
        L.gslm_mark_visible.argtypes = [C.c_int32, vp, vp, vp, vp, vp]
        L.gslm_forward.argtypes = [vp, vp, ALLOC_FN, vp, vp, vp, vp, vp, vp, vp]
        L.gslm_render.argtypes = [vp, vp, ALLOC_FN, vp, vp, vp, vp, vp, vp]
        L.gslm_render_batch.argtypes = [vp, C.c_int32, vp, ALLOC_FN, vp, vp, vp, vp, vp, vp]
        L.gslm_forward_batch.argtypes = [vp, C.c_int32, vp, ALLOC_FN, vp, vp, vp, vp, vp, vp, vp, vp]
        L.gslm_count_instances.argtypes = [vp, C.c_int32, vp, ALLOC_FN, vp, vp, vp]
        L.gslm_fp32_probe_scratch_bytes.restype = C.c_size_t
        L.gslm_fp32_peak_probe.argtypes = [C.c_int32, vp, vp, vp, vp]
        L.gslm_forward_jvp.argtypes = [vp, vp, vp, vp, ALLOC_FN, vp, vp, vp, vp, vp, vp, vp, vp, vp]
        L.gslm_backward.argtypes = [vp, vp, vp, vp, vp, vp, C.c_int32, vp, vp, vp, vp, vp]
        L.gslm_jvp_cached.argtypes = [vp, vp, vp, vp, vp, vp, vp, C.c_int32, vp, vp, vp, vp]
        L.gslm_jvp_cached_noalloc.argtypes = [vp, vp, vp, vp, vp, vp, vp, C.c_int32, vp, vp, vp]
        L.gslm_jtjv.argtypes = [vp, vp, vp, vp, vp, vp, vp, C.c_int32, vp, vp, vp, vp, vp]
        L.gslm_jtjv_batch_scratch_bytes.argtypes = [C.c_int32, C.c_int32, vp]
        L.gslm_jtjv_batch.argtypes = [vp, vp, C.c_int32, vp, vp, vp, vp, vp, vp, vp, vp, C.c_size_t, vp, vp, vp]
        L.gslm_jtjv_batch_sliced.argtypes = [vp, vp, C.c_int32, vp, vp, vp, vp, vp, vp, vp, vp, C.c_size_t, vp, vp,
                                             C.c_int32, SLICE_FN, vp, vp]
        L.gslm_backward_batch_sliced.argtypes = [vp, C.c_int32, vp, vp, vp, vp, vp, vp, vp, vp, C.c_size_t, vp,
                                                 C.c_int32, SLICE_FN, vp, vp]
        L.gslm_activation_vjp_range.argtypes = [vp, vp, vp, vp, vp, C.c_int32, C.c_int32, vp]
        L.gslm_jvp_batch_scratch_bytes.argtypes = [C.c_int32, C.c_int32]
        L.gslm_jvp_batch.argtypes = [vp, vp, C.c_int32, vp, vp, vp, vp, vp, vp, C.c_size_t, vp, vp]
        L.gslm_backward_batch_scratch_bytes.argtypes = [C.c_int32, vp]
        L.gslm_backward_batch.argtypes = [vp, C.c_int32, vp, vp, vp, vp, vp, vp, vp, vp, C.c_size_t, vp, vp]
        L.gslm_loss_lin_bytes.restype = C.c_size_t
        L.gslm_loss_lin_bytes.argtypes = [C.c_int32, C.c_int32, C.c_int32]
        L.gslm_loss_blocks.argtypes = [vp, vp, vp, vp, vp]
        L.gslm_loss_blocks_jvp.argtypes = [vp, vp, vp, vp, vp, vp]
        L.gslm_loss_blocks_vjp.argtypes = [vp, vp, vp, vp, vp, C.c_int32, vp]
        L.gslm_activation_jvp.argtypes = [vp, vp, vp, vp]
        L.gslm_activation_vjp.argtypes = [vp, vp, vp, vp, vp, vp]
        L.gslm_multimem_allreduce_f32.argtypes = [vp, C.c_size_t, C.c_int32, C.c_int32, vp]
        L.gslm_densification_stats.argtypes = [C.c_int32, vp, vp, vp, vp, vp, vp]
        L.gslm_buffer_view.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_int64, vp, vp, vp, vp, vp]
        L.gslm_tangent_records.argtypes = [vp, vp, vp, vp, vp, vp, vp]
        L.gslm_timing_stage_name.restype = C.c_char_p
        L.gslm_timing_collect.argtypes = [vp, vp]
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise RuntimeError("gslm_b200: " + lib().gslm_last_error().decode())


def _ptr(t):
    """Device pointer of a float32/int32/uint8 CUDA tensor, None for empty / None."""
    if t is None or (isinstance(t, torch.Tensor) and t.numel() == 0):
        return None
    if not t.is_cuda:
        raise RuntimeError("gslm_b200: tensors must live on a CUDA device (no CPU path)")
    return t.data_ptr()


def _f32(t):
    if t is None:
        return None
    if not isinstance(t, torch.Tensor):
        return t
    if t.numel() == 0:
        return t
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _stream():
    return torch.cuda.current_stream().cuda_stream


class _Alloc:
    """The three opaque scratch tensors (geomBuffer, binningBuffer, imgBuffer) are torch uint8 tensors
    created through this callback -- the resizeFunctional of DGR/rasterize_points.cu:28-34."""

    def __init__(self, device):
        self.device = device
        self.bufs = {}
        self.cb = ALLOC_FN(self._alloc)

    def _alloc(self, _ctx, which, nbytes):
        which, nbytes = int(which), max(int(nbytes), 1)
        # batch calls (gslm_forward_batch) ask with kind | (index << 8): per-view buffers / the two sort scratch slots
        kind, index = (which & 0xff, which >> 8) if which >= 0 else (which, 0)
        if kind == 3:
            # sort scratch: dead when the call returns, so ONE grow-only buffer per (device, stream, slot) serves every
            # forward call (stream order makes the reuse safe); a fresh multi-GB torch.empty per view costs
            # milliseconds of allocator work and fragments the cache
            key = (self.device, _stream(), index)
            t = _SORT_SCRATCH.get(key)
            if t is None or t.numel() < nbytes:
                _SORT_SCRATCH.pop(key, None)
                t = None
                t = torch.empty(nbytes + nbytes // 4, dtype=torch.uint8, device=self.device)
                _SORT_SCRATCH[key] = t
        else:
            t = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        self.bufs[which] = t
        return t.data_ptr()

    def get(self, which):
        t = self.bufs.get(which)
        return t if t is not None else torch.empty(0, dtype=torch.uint8, device=self.device)


_SORT_SCRATCH = {}            # (device, stream) -> grow-only uint8 tensor


def _view_tuple(v):
    return v[0], v[1], float(v[2]), float(v[3]), int(v[4]), int(v[5]), v[6]


def _camera_array(background, views, antialiasing, debug, tile_culling):
    n = len(views)
    cams = (GslmCamera * max(n, 1))()
    keep = []
    for i, v in enumerate(views):
        vm, pm, tfx, tfy, H, W, campos = _view_tuple(v)
        c, k = _camera(background, vm, pm, tfx, tfy, H, W, campos, False, antialiasing, debug, tile_culling)
        C.memmove(C.addressof(cams[i]), C.addressof(c), C.sizeof(GslmCamera))
        keep.append(k)
    return cams, keep


def release_scratch():
    """Drop the cached sort scratch (e.g. before handing the memory to another phase)."""
    _SORT_SCRATCH.clear()


KEEP_SORT_SCRATCH = False     # tests: keep the sort scratch (sorted keys) of the last forward call alive
LAST_SORT_SCRATCH = None


def _after_forward(al):
    global LAST_SORT_SCRATCH
    if KEEP_SORT_SCRATCH:
        LAST_SORT_SCRATCH = al.bufs.get(3)
    # al -> cb -> bound method -> al is a reference cycle: without this the three buffers of every call would stay
    # alive until the cycle collector runs, and every forward would cudaMalloc fresh blocks (measured: 6 ms per call)
    al.cb = None


def _gaussians(means3D, colors_precomp, opacities, scales, rotations, scale_modifier, cov3D_precomp, sh, degree):
    keep = [_f32(means3D), _f32(colors_precomp), _f32(opacities), _f32(scales), _f32(rotations), _f32(cov3D_precomp),
            _f32(sh)]
    m, cp, op, sc, ro, cov, shs = keep
    if m.dim() != 2 or m.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")   # rasterize_points.cu:59-61
    g = GslmGaussians()
    g.P = m.shape[0]
    g.sh_degree = int(degree)
    g.M = int(shs.shape[1]) if (shs is not None and shs.numel() != 0) else 0
    g.scale_modifier = float(scale_modifier)
    g.means3D, g.opacities, g.scales, g.rotations = _ptr(m), _ptr(op), _ptr(sc), _ptr(ro)
    g.cov3D_precomp, g.shs, g.colors_precomp = _ptr(cov), _ptr(shs), _ptr(cp)
    return g, keep


def _tangent_gaussians(means3D, colors_precomp, opacities, scales, rotations, cov3D_precomp, sh):
    keep = [_f32(means3D), _f32(colors_precomp), _f32(opacities), _f32(scales), _f32(rotations), _f32(cov3D_precomp),
            _f32(sh)]
    m, cp, op, sc, ro, cov, shs = keep
    g = GslmGaussians()
    g.means3D, g.opacities, g.scales, g.rotations = _ptr(m), _ptr(op), _ptr(sc), _ptr(ro)
    g.cov3D_precomp, g.shs, g.colors_precomp = _ptr(cov), _ptr(shs), _ptr(cp)
    return g, keep


TILE_CULLING = False     # default of the reference-shaped entry points: bin like the reference (bit-exact point_list)


def _camera(bg, viewmatrix, projmatrix, tan_fovx, tan_fovy, H, W, campos, prefiltered, antialiasing, debug,
            tile_culling=None):
    keep = [_f32(bg), _f32(viewmatrix), _f32(projmatrix), _f32(campos)]
    c = GslmCamera()
    c.width, c.height = int(W), int(H)
    c.tan_fovx, c.tan_fovy = float(tan_fovx), float(tan_fovy)
    c.bg, c.viewmatrix, c.projmatrix, c.campos = [_ptr(k) for k in keep]
    c.prefiltered, c.antialiasing, c.debug = int(bool(prefiltered)), int(bool(antialiasing)), int(bool(debug))
    c.tile_culling = int(bool(TILE_CULLING if tile_culling is None else tile_culling))
    return c, keep


def _as_float(x):
    return float(x.item()) if isinstance(x, torch.Tensor) else float(x)


# ---------------------------------------------------------------------------------------
# the four functions of the reference module (DGR/ext.cpp:15-20)
# ---------------------------------------------------------------------------------------
def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                        viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                        prefiltered, antialiasing, debug, tile_culling=None):
    """DGR/rasterize_points.cu:36-141 -> (num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer, invdepth)"""
    L = lib()
    dev = means3D.device
    g, k1 = _gaussians(means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, sh, degree)
    cam, k2 = _camera(background, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, campos,
                      prefiltered, antialiasing, debug, tile_culling)
    P, H, W = g.P, int(image_height), int(image_width)
    out_color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
    out_invdepth = torch.empty((1, H, W), dtype=torch.float32, device=dev)
    radii = torch.zeros((P,), dtype=torch.int32, device=dev)
    if P == 0:   # rasterize_points.cu:105
        out_color.zero_(); out_invdepth.zero_()
        e = torch.empty(0, dtype=torch.uint8, device=dev)
        return NumRendered(0, 0), out_color, radii, e, e.clone(), e.clone(), out_invdepth
    al = _Alloc(dev)
    n, nc = C.c_int32(0), C.c_int32(0)
    _check(L.gslm_forward(C.addressof(g), C.addressof(cam), al.cb, None, out_color.data_ptr(),
                          out_invdepth.data_ptr(), radii.data_ptr(), C.addressof(n), C.addressof(nc), _stream()))
    _after_forward(al)
    return NumRendered(n.value, nc.value), out_color, radii, al.get(0), al.get(1), al.get(2), out_invdepth


def render_only(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos, prefiltered,
                antialiasing, debug, tile_culling=None):
    """The forward pass without the linearisation cache (gslm_render): (num_rendered, color, radii, invdepth).
    Same arguments as rasterize_gaussians; nothing that comes out of it can be differentiated."""
    L = lib()
    dev = means3D.device
    g, k1 = _gaussians(means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, sh, degree)
    cam, k2 = _camera(background, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, campos,
                      prefiltered, antialiasing, debug, tile_culling)
    P, H, W = g.P, int(image_height), int(image_width)
    out_color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
    out_invdepth = torch.empty((1, H, W), dtype=torch.float32, device=dev)
    radii = torch.zeros((P,), dtype=torch.int32, device=dev)
    if P == 0:
        out_color.zero_(); out_invdepth.zero_()
        return 0, out_color, radii, out_invdepth
    al = _Alloc(dev)
    n = C.c_int32(0)
    _check(L.gslm_render(C.addressof(g), C.addressof(cam), al.cb, None, out_color.data_ptr(),
                         out_invdepth.data_ptr(), radii.data_ptr(), C.addressof(n), _stream()))
    _after_forward(al)
    return n.value, out_color, radii, out_invdepth


def render_batch(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, views, sh,
                 degree, antialiasing, debug, out=None, tile_culling=None):
    """gslm_render_batch: render-only passes of one set of Gaussians from several cameras with one host
    synchronisation for the whole batch.  views: sequence of (viewmatrix, projmatrix, tan_fovx, tan_fovy, H, W,
    campos).  out: optional list of [3,H,W] float32 CUDA tensors to render into (e.g. slices of a stacked batch).
    Returns (num_rendered list, list of colour images)."""
    L = lib()
    dev = means3D.device
    g, k1 = _gaussians(means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, sh, degree)
    n = len(views)
    cams = (GslmCamera * max(n, 1))()
    keep = []
    for i, (vm, pm, tfx, tfy, H, W, campos) in enumerate(views):
        c, k = _camera(background, vm, pm, tfx, tfy, H, W, campos, False, antialiasing, debug, tile_culling)
        C.memmove(C.addressof(cams[i]), C.addressof(c), C.sizeof(GslmCamera))
        keep.append(k)
    if out is None:
        out = [torch.empty((3, int(v[4]), int(v[5])), dtype=torch.float32, device=dev) for v in views]
    for o, v in zip(out, views):
        if not (o.is_cuda and o.dtype == torch.float32 and o.is_contiguous() and o.numel() == 3 * int(v[4]) * int(v[5])):
            raise RuntimeError("gslm_b200: render_batch outputs must be contiguous float32 [3,H,W] CUDA tensors")
    if n == 0:
        return [], out
    if g.P == 0:
        for o in out:
            o.zero_()
        return [0] * n, out
    ptrs = (C.c_void_p * n)(*[o.data_ptr() for o in out])
    counts = (C.c_int32 * n)()
    al = _Alloc(dev)
    _check(L.gslm_render_batch(C.addressof(g), n, C.addressof(cams), al.cb, None, C.addressof(ptrs), None, None,
                               C.addressof(counts), _stream()))
    _after_forward(al)
    return list(counts), out


def count_instances(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, views, sh,
                    degree, antialiasing=False, tile_culling=None):
    """gslm_count_instances: num_rendered of every view (list of ints) without rendering; one host synchronisation."""
    L = lib()
    g, k1 = _gaussians(means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, sh, degree)
    n = len(views)
    if n == 0 or g.P == 0:
        return [0] * n
    cams, keep = _camera_array(background, views, antialiasing, False, tile_culling)
    counts = (C.c_int32 * n)()
    al = _Alloc(k1[0].device)
    _check(L.gslm_count_instances(C.addressof(g), n, C.addressof(cams), al.cb, None, C.addressof(counts), _stream()))
    al.cb = None
    return list(counts)


class ResidualSpec:
    """Per-view outputs of the fused residual epilogue of forward_batch (GslmResidualEpilogue): strided views into the
    solver's padded batch tensors.  image / mask: [3,H,W] views (any strides with unit column stride), weight:
    contiguous [3,H,W]; gt: [3,H,W] view or None; alpha_mask: [1,H,W] / [H,W] view or None."""
    __slots__ = ("gt", "alpha_mask", "image", "mask", "weight", "weight_scale", "apply_alpha_mask", "subtract_gt")

    def __init__(self, image=None, mask=None, weight=None, gt=None, alpha_mask=None, weight_scale=1.0,
                 apply_alpha_mask=True, subtract_gt=True):
        self.gt, self.alpha_mask, self.image, self.mask, self.weight = gt, alpha_mask, image, mask, weight
        self.weight_scale, self.apply_alpha_mask, self.subtract_gt = weight_scale, apply_alpha_mask, subtract_gt

    def c_struct(self, H, W):
        def chk3(t, name):
            if t is None:
                return None, 0, 0
            if not (t.is_cuda and t.dtype == torch.float32 and tuple(t.shape) == (3, H, W) and (W == 1 or t.stride(2) == 1)):
                raise RuntimeError(f"gslm_b200: epilogue {name} must be a float32 CUDA [3,{H},{W}] view with unit column stride")
            return t.data_ptr(), t.stride(0), t.stride(1)
        e = GslmResidualEpilogue()
        e.gt, e.gt_plane, e.gt_row = chk3(self.gt, "gt")
        e.out_image, e.image_plane, e.image_row = chk3(self.image, "image")
        e.out_mask, e.mask_plane, e.mask_row = chk3(self.mask, "mask")
        if self.weight is not None:
            if not (self.weight.is_contiguous() and tuple(self.weight.shape) == (3, H, W) and self.weight.is_cuda
                    and self.weight.dtype == torch.float32):
                raise RuntimeError("gslm_b200: epilogue weight must be a contiguous float32 CUDA [3,H,W] tensor")
            e.out_weight = self.weight.data_ptr()
        am = self.alpha_mask
        if am is not None:
            am = am.reshape(am.shape[-2], am.shape[-1]) if am.dim() == 3 else am
            if not (am.is_cuda and am.dtype == torch.float32 and tuple(am.shape) == (H, W) and (W == 1 or am.stride(1) == 1)):
                raise RuntimeError("gslm_b200: epilogue alpha_mask must be a float32 CUDA [H,W] view")
            e.alpha_mask, e.alpha_row = am.data_ptr(), am.stride(0)
        e.weight_scale = float(self.weight_scale)
        e.apply_alpha_mask, e.subtract_gt = int(bool(self.apply_alpha_mask)), int(bool(self.subtract_gt))
        return e

    def fill_from(self, color):
        """The same outputs from an image that did not come out of gslm_forward_batch: the scene without Gaussians,
        whose all-zero images (DGR/rasterize_points.cu:69,96) forward_batch produces view by view.  Same arithmetic as k_residual_epilogue."""
        m = ((color >= 0) & (color <= 1)).to(color.dtype)
        img = color.clamp(0, 1)
        if self.apply_alpha_mask and self.alpha_mask is not None:
            am = self.alpha_mask.reshape(1, color.shape[-2], color.shape[-1])
            img, m = img * am, m * am
        if self.gt is not None and self.subtract_gt:
            img = img - self.gt
        if self.image is not None:
            self.image.copy_(img)
        if self.mask is not None:
            self.mask.copy_(m)
        if self.weight is not None:
            self.weight.copy_(float(self.weight_scale) * m * m)


def forward_batch(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, views, sh,
                  degree, antialiasing, debug, residual=None, tile_culling=None, want_invdepth=False):
    """gslm_forward_batch: the LINEARISING forward pass (rasterize_gaussians) of several cameras in one library call.
    views: sequence of (viewmatrix, projmatrix, tan_fovx, tan_fovy, H, W, campos); residual: optional list of
    ResidualSpec (one per view).  Returns a list of per-view tuples with the layout of rasterize_gaussians:
    (num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer, invdepth-or-None)."""
    L = lib()
    dev = means3D.device
    g, k1 = _gaussians(means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, sh, degree)
    n = len(views)
    if n == 0:
        return []
    if g.P == 0:
        out = [rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                                   cov3D_precomp, v[0], v[1], v[2], v[3], v[4], v[5], sh, degree, v[6], False,
                                   antialiasing, debug, tile_culling) for v in views]
        for o, spec in zip(out, residual or ()):
            if spec is not None:
                spec.fill_from(o[1])        # the caller's residual / mask / weight tensors are torch.empty
        return out
    cams, keep = _camera_array(background, views, antialiasing, debug, tile_culling)
    sizes = [(int(v[4]), int(v[5])) for v in views]
    colors_out = [torch.empty((3, H, W), dtype=torch.float32, device=dev) for H, W in sizes]
    invd = [torch.empty((1, H, W), dtype=torch.float32, device=dev) for H, W in sizes] if want_invdepth else None
    radii = torch.zeros((n, g.P), dtype=torch.int32, device=dev)
    VP = C.c_void_p * n
    p_color = VP(*[t.data_ptr() for t in colors_out])
    p_invd = VP(*[t.data_ptr() for t in invd]) if invd else None
    p_radii = VP(*[radii[i].data_ptr() for i in range(n)])
    epi = None
    if residual is not None:
        epi = (GslmResidualEpilogue * n)()
        for i, (spec, (H, W)) in enumerate(zip(residual, sizes)):
            epi[i] = spec.c_struct(H, W) if spec is not None else GslmResidualEpilogue()
    nr, nc = (C.c_int32 * n)(), (C.c_int32 * n)()
    al = _Alloc(dev)
    _check(L.gslm_forward_batch(C.addressof(g), n, C.addressof(cams), al.cb, None, C.addressof(p_color),
                                C.addressof(p_invd) if p_invd is not None else None, C.addressof(p_radii),
                                C.addressof(epi) if epi is not None else None, C.addressof(nr), C.addressof(nc),
                                _stream()))
    _after_forward(al)
    out = []
    for i in range(n):
        out.append((NumRendered(nr[i], nc[i]), colors_out[i], radii[i], al.get(0 | (i << 8)), al.get(1 | (i << 8)),
                    al.get(2 | (i << 8)), invd[i] if invd else None))
    return out


def fp32_peak_probe(mode=0, device=None):
    """gslm_fp32_peak_probe -> (TFLOP/s, ms): FP32 throughput measured on the current device and stream."""
    L = lib()
    dev = device if device is not None else torch.device("cuda", torch.cuda.current_device())
    scratch = torch.empty(L.gslm_fp32_probe_scratch_bytes(), dtype=torch.uint8, device=dev)
    tf, ms = C.c_float(0), C.c_float(0)
    _check(L.gslm_fp32_peak_probe(int(mode), scratch.data_ptr(), C.addressof(tf), C.addressof(ms), _stream()))
    return float(tf.value), float(ms.value)


def rasterize_gaussians_jvp(background, background_grad, means3D, means3D_grad, colors, colors_grad, opacity,
                            opacity_grad, scales, scales_grad, rotations, rotations_grad, scale_modifier,
                            scale_modifier_grad, cov3D_precomp, cov3D_precomp_grad, viewmatrix, viewmatrix_grad,
                            projmatrix, projmatrix_grad, tan_fovx, tan_fovx_grad, tan_fovy, tan_fovy_grad,
                            image_height, image_width, sh, sh_grad, degree, campos, campos_grad, prefiltered,
                            antialiasing, debug):
    """DGR/rasterize_points.cu:143-298 -> 9-tuple (the 7 above + color_grad, invdepth_grad)"""
    L = lib()
    dev = means3D.device
    g, k1 = _gaussians(means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, sh, degree)
    gt, k3 = _tangent_gaussians(means3D_grad, colors_grad, opacity_grad, scales_grad, rotations_grad,
                                cov3D_precomp_grad, sh_grad)
    cam, k2 = _camera(background, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, campos,
                      prefiltered, antialiasing, debug)
    k4 = [_f32(background_grad), _f32(viewmatrix_grad), _f32(projmatrix_grad), _f32(campos_grad)]
    ct = GslmCameraTangent()
    ct.tan_fovx, ct.tan_fovy = _as_float(tan_fovx_grad), _as_float(tan_fovy_grad)
    ct.scale_modifier = _as_float(scale_modifier_grad)
    ct.bg, ct.viewmatrix, ct.projmatrix, ct.campos = [_ptr(k) for k in k4]
    P, H, W = g.P, int(image_height), int(image_width)
    out_color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
    out_invdepth = torch.empty((1, H, W), dtype=torch.float32, device=dev)
    out_color_t = torch.empty((3, H, W), dtype=torch.float32, device=dev)
    out_invdepth_t = torch.empty((1, H, W), dtype=torch.float32, device=dev)
    radii = torch.zeros((P,), dtype=torch.int32, device=dev)
    if P == 0:
        for t in (out_color, out_invdepth, out_color_t, out_invdepth_t):
            t.zero_()
        e = torch.empty(0, dtype=torch.uint8, device=dev)
        return NumRendered(0, 0), out_color, radii, e, e.clone(), e.clone(), out_invdepth, out_color_t, out_invdepth_t
    al = _Alloc(dev)
    n, nc = C.c_int32(0), C.c_int32(0)
    _check(L.gslm_forward_jvp(C.addressof(g), C.addressof(gt), C.addressof(cam), C.addressof(ct), al.cb, None,
                              out_color.data_ptr(), out_invdepth.data_ptr(), radii.data_ptr(),
                              out_color_t.data_ptr(), out_invdepth_t.data_ptr(), C.addressof(n), C.addressof(nc),
                              _stream()))
    _after_forward(al)
    return (NumRendered(n.value, nc.value), out_color, radii, al.get(0), al.get(1), al.get(2), out_invdepth, out_color_t,
            out_invdepth_t)


def rasterize_gaussians_backward(background, means3D, radii, colors, opacities, scales, rotations, scale_modifier,
                                 cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color,
                                 dL_dout_invdepth, sh, degree, campos, geomBuffer, R, binningBuffer, imageBuffer,
                                 antialiasing, debug, accumulate_into=None):
    """DGR/rasterize_points.cu:300-399 -> (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh,
    dL_dscales, dL_drotations).  `accumulate_into` (optional 8-tuple) adds into existing tensors instead."""
    L = lib()
    dev = means3D.device
    g, k1 = _gaussians(means3D, colors, opacities, scales, rotations, scale_modifier, cov3D_precomp, sh, degree)
    H, W = int(dL_dout_color.shape[-2]), int(dL_dout_color.shape[-1])
    cam, k2 = _camera(background, viewmatrix, projmatrix, tan_fovx, tan_fovy, H, W, campos, False, antialiasing, debug)
    P, M = g.P, g.M
    if accumulate_into is None:
        # overwrite mode: the kernel writes every element (zeros for culled Gaussians), so no memset
        # of the eight tensors is needed (the reference zero-fills ten, rasterize_points.cu:339-348)
        z = lambda *s: torch.empty(s, dtype=torch.float32, device=dev)
        outs = (z(P, 3), z(P, 3), z(P, 1), z(P, 3), z(P, 6), z(P, M, 3), z(P, 3), z(P, 4))
    else:
        outs = accumulate_into
    if P == 0:
        return outs
    dcol = _f32(dL_dout_color)
    dinv = _f32(dL_dout_invdepth) if (dL_dout_invdepth is not None and dL_dout_invdepth.numel() != 0) else None
    gr = GslmGrads()
    (gr.dL_dmeans2D, gr.dL_dcolors, gr.dL_dopacity, gr.dL_dmeans3D, gr.dL_dcov3D, gr.dL_dsh, gr.dL_dscales,
     gr.dL_drotations) = [_ptr(o) for o in outs]
    gr.accumulate = 0 if accumulate_into is None else 1
    Rc = _cached(R, imageBuffer, W, H)
    scratch = torch.empty(L.gslm_backward_scratch_bytes(Rc), dtype=torch.uint8, device=dev)
    _check(L.gslm_backward(C.addressof(g), C.addressof(cam), _ptr(radii), _ptr(geomBuffer), _ptr(binningBuffer),
                           _ptr(imageBuffer), Rc, _ptr(dcol), _ptr(dinv), scratch.data_ptr(), C.addressof(gr),
                           _stream()))
    return outs


def mark_visible(means3D, viewmatrix, projmatrix):
    """DGR/rasterize_points.cu:401-420"""
    m = _f32(means3D)
    P = m.shape[0]
    present = torch.zeros((P,), dtype=torch.bool, device=m.device)
    if P:
        v, p = _f32(viewmatrix), _f32(projmatrix)
        _check(lib().gslm_mark_visible(P, _ptr(m), _ptr(v), _ptr(p), present.data_ptr(), _stream()))
    return present


# ---------------------------------------------------------------------------------------
# additive: linearisation-cached entry points (used by solver.solver_functions)
# ---------------------------------------------------------------------------------------
def jvp_cached(gauss_args, tangent_args, cam_args, radii, geomBuffer, R, binningBuffer, imageBuffer,
               want_invdepth=False):
    """J v for a linearised view. gauss_args = (means3D, colors, opacity, scales, rotations, scale_modifier,
    cov3D_precomp, sh, degree); tangent_args = (means3D_t, colors_t, opacity_t, scales_t, rotations_t, cov3D_t,
    sh_t); cam_args = (bg, viewmatrix, projmatrix, tan_fovx, tan_fovy, H, W, campos, antialiasing, debug)."""
    L = lib()
    g, k1 = _gaussians(*gauss_args)
    gt, k3 = _tangent_gaussians(*tangent_args)
    bg, vm, pm, tfx, tfy, H, W, campos, aa, dbg = cam_args
    cam, k2 = _camera(bg, vm, pm, tfx, tfy, H, W, campos, False, aa, dbg)
    dev = k1[0].device
    out_t = torch.empty((3, H, W), dtype=torch.float32, device=dev)
    out_d = torch.empty((1, H, W), dtype=torch.float32, device=dev) if want_invdepth else None
    scratch = torch.empty(L.gslm_jvp_scratch_bytes(g.P), dtype=torch.uint8, device=dev)
    _check(L.gslm_jvp_cached(C.addressof(g), C.addressof(gt), C.addressof(cam), _ptr(radii), _ptr(geomBuffer),
                             _ptr(binningBuffer), _ptr(imageBuffer), _cached(R, imageBuffer, W, H), scratch.data_ptr(),
                             out_t.data_ptr(),
                             _ptr(out_d), _stream()))
    return out_t, out_d


NOALLOC_MAX_GAUSSIANS = 262144   # GSLM_NOALLOC_MAX_GAUSSIANS in include/gslm_b200.h


def jvp_cached_noalloc(gauss_args, tangent_args, cam_args, radii, geomBuffer, R, binningBuffer, imageBuffer,
                       out_color_tangent, out_invdepth_tangent=None):
    """gslm_jvp_cached_noalloc: J v for a linearised view without any device allocation inside the call.  Meant for
    hosts that cannot give the library scratch memory (no device allocator in their language, HBM already owned by
    another pool, or a frame loop that must not synchronise on cudaMalloc).  The caller brings the outputs
    (`out_color_tangent` [3,H,W] float32, optionally `out_invdepth_tangent` [1,H,W]); the tangent records go to a
    12 MiB array inside the library, shared by every call on the device: keep calls on one stream.  Raises for more
    than NOALLOC_MAX_GAUSSIANS Gaussians.  Arguments otherwise as jvp_cached."""
    L = lib()
    g, k1 = _gaussians(*gauss_args)
    gt, k3 = _tangent_gaussians(*tangent_args)
    bg, vm, pm, tfx, tfy, H, W, campos, aa, dbg = cam_args
    cam, k2 = _camera(bg, vm, pm, tfx, tfy, H, W, campos, False, aa, dbg)
    assert out_color_tangent.is_cuda and out_color_tangent.dtype == torch.float32 and \
        out_color_tangent.is_contiguous() and tuple(out_color_tangent.shape) == (3, H, W)
    if out_invdepth_tangent is not None:
        assert out_invdepth_tangent.dtype == torch.float32 and out_invdepth_tangent.is_contiguous() and \
            tuple(out_invdepth_tangent.shape) == (1, H, W)
    _check(L.gslm_jvp_cached_noalloc(C.addressof(g), C.addressof(gt), C.addressof(cam), _ptr(radii), _ptr(geomBuffer),
                                     _ptr(binningBuffer), _ptr(imageBuffer), _cached(R, imageBuffer, W, H),
                                     out_color_tangent.data_ptr(), _ptr(out_invdepth_tangent), _stream()))
    return out_color_tangent, out_invdepth_tangent


def tangent_records(gauss_args, tangent_args, cam_args, geomBuffer, cam_tangent=None):
    """Introspection for tests (gslm_tangent_records): [P,12] per-Gaussian tangent records of a linearised view; rows of
    Gaussians without a cached instance stay NaN.  cam_tangent: dict with any of bg, viewmatrix, projmatrix, campos
    (tensors), tan_fovx, tan_fovy, scale_modifier (floats)."""
    L = lib()
    g, k1 = _gaussians(*gauss_args)
    gt, k3 = _tangent_gaussians(*tangent_args)
    bg, vm, pm, tfx, tfy, H, W, campos, aa, dbg = cam_args
    cam, k2 = _camera(bg, vm, pm, tfx, tfy, H, W, campos, False, aa, dbg)
    ct, k4 = None, []
    if cam_tangent:
        ct = GslmCameraTangent()
        k4 = [_f32(cam_tangent.get(k)) if cam_tangent.get(k) is not None else None
              for k in ("bg", "viewmatrix", "projmatrix", "campos")]
        ct.bg, ct.viewmatrix, ct.projmatrix, ct.campos = [_ptr(k) if k is not None else None for k in k4]
        ct.tan_fovx, ct.tan_fovy = float(cam_tangent.get("tan_fovx", 0.0)), float(cam_tangent.get("tan_fovy", 0.0))
        ct.scale_modifier = float(cam_tangent.get("scale_modifier", 0.0))
    out = torch.full((g.P, 12), float("nan"), dtype=torch.float32, device=k1[0].device)
    _check(L.gslm_tangent_records(C.addressof(g), C.addressof(gt), C.addressof(cam),
                                  C.addressof(ct) if ct is not None else None, _ptr(geomBuffer), out.data_ptr(),
                                  _stream()))
    return out


def jtjv(gauss_args, tangent_args, cam_args, radii, geomBuffer, R, binningBuffer, imageBuffer, weight, outs,
         out_vJJv=None, scratch=None):
    """outs (8-tuple like rasterize_gaussians_backward, entries may be None) += J^T diag(weight) J v."""
    L = lib()
    g, k1 = _gaussians(*gauss_args)
    gt, k3 = _tangent_gaussians(*tangent_args)
    bg, vm, pm, tfx, tfy, H, W, campos, aa, dbg = cam_args
    cam, k2 = _camera(bg, vm, pm, tfx, tfy, H, W, campos, False, aa, dbg)
    dev = k1[0].device
    w = _f32(weight)
    Rc = _cached(R, imageBuffer, W, H)
    need = L.gslm_jtjv_scratch_bytes(g.P, Rc)
    if scratch is None or scratch.numel() < need:
        scratch = torch.empty(need, dtype=torch.uint8, device=dev)
    gr = GslmGrads()
    (gr.dL_dmeans2D, gr.dL_dcolors, gr.dL_dopacity, gr.dL_dmeans3D, gr.dL_dcov3D, gr.dL_dsh, gr.dL_dscales,
     gr.dL_drotations) = [_ptr(o) for o in outs]
    gr.accumulate = 1
    _check(L.gslm_jtjv(C.addressof(g), C.addressof(gt), C.addressof(cam), _ptr(radii), _ptr(geomBuffer),
                       _ptr(binningBuffer), _ptr(imageBuffer), Rc, _ptr(w), scratch.data_ptr(),
                       C.addressof(gr), _ptr(out_vJJv), _stream()))
    return scratch


class JtjvBatchPlan:
    """Host-side argument block of gslm_jtjv_batch for a fixed set of linearised views (built once per
    linearisation, reused by every CG iteration): ctypes arrays of cameras, buffer pointers, compact counts and
    weights, the scratch tensor, and references that keep every tensor alive."""

    def __init__(self, views, device):
        """views: list of (cam_args, radii, geomBuffer, R, binningBuffer, imageBuffer, weight-or-None)."""
        L = lib()
        n = len(views)
        self.n = n
        self.keep = []
        self.cams = (GslmCamera * max(n, 1))()
        VP = C.c_void_p * max(n, 1)
        self.radii, self.geom, self.binning, self.image, self.weights = VP(), VP(), VP(), VP(), VP()
        self.num_cached = (C.c_int32 * max(n, 1))()
        any_weight = False
        for i, (cam_args, radii, geomBuffer, R, binningBuffer, imageBuffer, weight) in enumerate(views):
            bg, vm, pm, tfx, tfy, H, W, campos, aa, dbg = cam_args
            cam, k = _camera(bg, vm, pm, tfx, tfy, H, W, campos, False, aa, dbg)
            self.cams[i] = cam
            w = _f32(weight)
            self.keep.append((k, radii, geomBuffer, binningBuffer, imageBuffer, w))
            self.radii[i], self.geom[i] = _ptr(radii), _ptr(geomBuffer)
            self.binning[i], self.image[i] = _ptr(binningBuffer), _ptr(imageBuffer)
            self.weights[i] = _ptr(w)
            any_weight = any_weight or w is not None
            self.num_cached[i] = _cached(R, imageBuffer, W, H)
        self.any_weight = any_weight
        self.device = device
        self.scratch = None
        self.P = None

    def scratch_for(self, P):
        if self.scratch is None or self.P != P:
            need = lib().gslm_jtjv_batch_scratch_bytes(int(P), self.n, C.cast(self.num_cached, C.c_void_p))
            self.scratch = torch.empty(max(int(need), 1), dtype=torch.uint8, device=self.device)
            self.P = P
        return self.scratch


class _SliceCallback:
    """Wraps a Python callable(slice, g_lo, g_hi) as a GslmSliceFn; exceptions raised inside the C call are kept and
    re-raised by check() afterwards (ctypes would only print them)."""

    def __init__(self, fn):
        self.err = None

        def tramp(_ctx, k, g_lo, g_hi):
            if self.err is None:
                try:
                    fn(int(k), int(g_lo), int(g_hi))
                except BaseException as e:      # noqa: BLE001 -- re-raised by check()
                    self.err = e
        self.c = SLICE_FN(tramp)

    def check(self):
        # cb -> c -> tramp -> cb is a reference cycle that would keep the callable (and the 2 x 59 P floats it owns)
        # alive until the cycle collector runs: every product would then cudaMalloc fresh buffers
        self.c = None
        err, self.err = self.err, None
        if err is not None:
            raise err


def jtjv_batch(gauss_args, tangent_args, plan, outs, out_vJJv=None, accumulate=True, n_slices=1, on_slice=None):
    """outs (8-tuple like rasterize_gaussians_backward, entries may be None) (+)= sum over the plan's views of
    J_v^T diag(weight_v) J_v v -- ONE call into the library per CG iteration.  on_slice(k, g_lo, g_hi): host callback
    after the launch that completes the Gaussian rows [g_lo, g_hi) (gslm_jtjv_batch_sliced)."""
    L = lib()
    g, k1 = _gaussians(*gauss_args)
    gt, k3 = _tangent_gaussians(*tangent_args)
    scratch = plan.scratch_for(g.P)
    gr = GslmGrads()
    (gr.dL_dmeans2D, gr.dL_dcolors, gr.dL_dopacity, gr.dL_dmeans3D, gr.dL_dcov3D, gr.dL_dsh, gr.dL_dscales,
     gr.dL_drotations) = [_ptr(o) for o in outs]
    gr.accumulate = 1 if accumulate else 0
    vp = lambda a: C.cast(a, C.c_void_p)
    cb = _SliceCallback(on_slice) if on_slice is not None else None
    _check(L.gslm_jtjv_batch_sliced(C.addressof(g), C.addressof(gt), plan.n, vp(plan.cams), vp(plan.radii),
                                    vp(plan.geom), vp(plan.binning), vp(plan.image), vp(plan.num_cached),
                                    vp(plan.weights) if plan.any_weight else None, scratch.data_ptr(),
                                    scratch.numel(), C.addressof(gr), _ptr(out_vJJv), int(n_slices),
                                    cb.c if cb else SLICE_FN(), None, _stream()))
    if cb:
        cb.check()


def jvp_batch(gauss_args, tangent_args, plan, out_images):
    """out_images[v] ([3,H_v,W_v] contiguous fp32 CUDA tensors, one per view of the plan) <- J_v v."""
    L = lib()
    g, k1 = _gaussians(*gauss_args)
    gt, k3 = _tangent_gaussians(*tangent_args)
    scratch = plan.scratch_for(g.P)            # the J^T J v scratch is a superset of what this call needs
    outs = (C.c_void_p * max(plan.n, 1))(*[_ptr(t) for t in out_images])
    vp = lambda a: C.cast(a, C.c_void_p)
    _check(L.gslm_jvp_batch(C.addressof(g), C.addressof(gt), plan.n, vp(plan.cams), vp(plan.geom), vp(plan.binning),
                            vp(plan.image), vp(plan.num_cached), scratch.data_ptr(), scratch.numel(), vp(outs),
                            _stream()))


def backward_batch(gauss_args, plan, cot_images, outs, accumulate=True, n_slices=1, on_slice=None):
    """outs (8-tuple like rasterize_gaussians_backward) (+)= sum over the plan's views of J_v^T cot_images[v];
    on_slice as in jtjv_batch."""
    L = lib()
    g, k1 = _gaussians(*gauss_args)
    scratch = plan.scratch_for(g.P)
    cots = (C.c_void_p * max(plan.n, 1))(*[_ptr(t) for t in cot_images])
    gr = GslmGrads()
    (gr.dL_dmeans2D, gr.dL_dcolors, gr.dL_dopacity, gr.dL_dmeans3D, gr.dL_dcov3D, gr.dL_dsh, gr.dL_dscales,
     gr.dL_drotations) = [_ptr(o) for o in outs]
    gr.accumulate = 1 if accumulate else 0
    vp = lambda a: C.cast(a, C.c_void_p)
    cb = _SliceCallback(on_slice) if on_slice is not None else None
    _check(L.gslm_backward_batch_sliced(C.addressof(g), plan.n, vp(plan.cams), vp(plan.radii), vp(plan.geom),
                                        vp(plan.binning), vp(plan.image), vp(plan.num_cached), vp(cots),
                                        scratch.data_ptr(), scratch.numel(), C.addressof(gr), int(n_slices),
                                        cb.c if cb else SLICE_FN(), None, _stream()))
    if cb:
        cb.check()


def buffer_view(P, W, H, R, geomBuffer, binningBuffer, imageBuffer, sortScratch=None):
    """Test helper: torch views of the state inside the opaque buffers (bit-exactness checks).
    sortScratch (the which = 3 allocation of the forward call, see KEEP_SORT_SCRATCH) adds `keys_sorted` and the
    full sorted `point_list`."""
    Rc = _cached(R, imageBuffer, W, H)
    R = int(R)
    v = GslmBufferView()
    _check(lib().gslm_buffer_view(int(P), int(W), int(H), R, Rc, _ptr(geomBuffer), _ptr(binningBuffer),
                                  _ptr(imageBuffer), _ptr(sortScratch), C.addressof(v)))

    def view(buf, ptr, nbytes, dtype, shape):
        off = ptr - buf.data_ptr()
        return buf[off:off + nbytes].view(dtype).view(shape)

    T = ((W + 15) // 16) * ((H + 15) // 16)
    N = W * H
    out = {"num_cached": Rc}
    if binningBuffer is not None and binningBuffer.numel() and Rc > 0:
        out["records"] = view(binningBuffer, v.records, 48 * Rc, torch.float32, (Rc, 12))
        out["masks"] = view(binningBuffer, v.masks, 32 * Rc, torch.int32, (Rc, 8))
        out["cid"] = view(binningBuffer, v.cid, 4 * Rc, torch.int32, (Rc,))
        out["g_list"] = view(binningBuffer, v.g_list, 4 * Rc, torch.int32, (Rc,))
    if sortScratch is not None and sortScratch.numel() and R > 0:
        out["tile_keys_sorted"] = view(sortScratch, v.tile_keys_sorted, 4 * R, torch.int32, (R,))
        out["point_list"] = view(sortScratch, v.point_list, 4 * R, torch.int32, (R,))
    if imageBuffer is not None and imageBuffer.numel():
        out["ranges"] = view(imageBuffer, v.ranges, 8 * T, torch.int32, (T, 2))
        out["n_contrib"] = view(imageBuffer, v.n_contrib, 4 * N, torch.int32, (H, W))
        out["n_contrib_c"] = view(imageBuffer, v.n_contrib_c, 4 * N, torch.int32, (H, W))
        out["final_T"] = view(imageBuffer, v.final_T, 4 * N, torch.float32, (H, W))
        out["cstart"] = view(imageBuffer, v.cstart, 4 * (T + 1), torch.int32, (T + 1,))
    if geomBuffer is not None and geomBuffer.numel():
        out["tiles_touched"] = view(geomBuffer, v.tiles_touched, 4 * P, torch.int32, (P,))
        out["point_offsets"] = view(geomBuffer, v.point_offsets, 4 * P, torch.int32, (P,))
        out["depths"] = view(geomBuffer, v.depths, 4 * P, torch.float32, (P,))
        out["means2D"] = view(geomBuffer, v.means2D, 8 * P, torch.float32, (P, 2))
        out["conic_opacity"] = view(geomBuffer, v.conic_opacity, 16 * P, torch.float32, (P, 4))
        if "tile_keys_sorted" in out:
            # the reference's 64-bit sort key (rasterizer_impl.cu:97-103): tile << 32 | depth bits of the instance
            dbits = out["depths"].view(torch.int32)[out["point_list"].long()].long() & 0xffffffff
            out["keys_sorted"] = (out["tile_keys_sorted"].long() << 32) | dbits
        out["rgb"] = view(geomBuffer, v.rgb, 12 * P, torch.float32, (P, 3))
        out["cov3D"] = view(geomBuffer, v.cov3D, 24 * P, torch.float32, (P, 6))
        out["clamped"] = view(geomBuffer, v.clamped, P, torch.uint8, (P,))
        out["g_off"] = view(geomBuffer, v.g_off, 4 * (P + 1), torch.int32, (P + 1,))
    return out


# ---------------------------------------------------------------------- residual side (csrc/loss_blocks.cu)
class LossBatch:
    """One batch of image planes [B,3,H,W] at a fixed linearisation point: holds the contiguous fp32 tensors the
    three loss-block calls share (images, alpha masks [B,1,H,W] or None, targets, per-image alphas / betas [B])."""

    def __init__(self, images, alpha_masks, gt_images, per_image_alphas, per_image_betas, tangent_masks=None):
        if images.dim() != 4 or images.shape[1] != 3:
            raise ValueError("images must be [B,3,H,W]")
        if not images.is_cuda:
            raise RuntimeError("gslm_b200: tensors must live on a CUDA device (no CPU path)")
        self.images = _f32(images.detach())
        self.B, _, self.H, self.W = self.images.shape
        self.gt = _f32(gt_images.detach().expand_as(self.images))
        self.masks = None
        if alpha_masks is not None:
            self.masks = _f32(alpha_masks.detach().expand(self.B, 1, self.H, self.W))
        dev = self.images.device
        self.alphas = _f32(torch.as_tensor(per_image_alphas, device=dev).detach().reshape(-1).expand(self.B))
        self.betas = _f32(torch.as_tensor(per_image_betas, device=dev).detach().reshape(-1).expand(self.B))
        self.lin = None
        # [B,3,H,W] factor between the caller's tangent / gradient and the images (clamp mask), fused into jvp / vjp
        self.tmask = None if tangent_masks is None else _f32(tangent_masks.detach().expand_as(self.images))
        self.c = GslmLossBatch(self.B, self.H, self.W, _ptr(self.images), _ptr(self.masks), _ptr(self.gt),
                               _ptr(self.alphas), _ptr(self.betas), _ptr(self.tmask))

    def blocks(self, want_lin=True):
        """(Ll1_per_pixel, ssim_loss_per_pixel) of compute_batch_loss_block (disable_ssim=False); keeps the
        linearisation record for jvp / vjp."""
        r1, r2 = torch.empty_like(self.images), torch.empty_like(self.images)
        if want_lin:
            n = lib().gslm_loss_lin_bytes(self.B, self.H, self.W)
            self.lin = torch.empty(max(n, 1), dtype=torch.uint8, device=self.images.device)
        _check(lib().gslm_loss_blocks(C.addressof(self.c), _ptr(r1), _ptr(r2), _ptr(self.lin) if want_lin else None,
                                      _stream()))
        return r1, r2

    def jvp(self, t_images):
        """Tangents of the two blocks for an image tangent [B,3,H,W]."""
        t = _f32(t_images)
        t1, t2 = torch.empty_like(self.images), torch.empty_like(self.images)
        _check(lib().gslm_loss_blocks_jvp(C.addressof(self.c), _ptr(self.lin), _ptr(t), _ptr(t1), _ptr(t2), _stream()))
        return t1, t2

    def vjp(self, u1, u2, out=None):
        """Pull-back of the block cotangents to the images; out: accumulate into this [B,3,H,W] tensor."""
        u1, u2 = _f32(u1), _f32(u2)
        g = torch.empty_like(self.images) if out is None else out
        _check(lib().gslm_loss_blocks_vjp(C.addressof(self.c), _ptr(self.lin), _ptr(u1), _ptr(u2), _ptr(g),
                                          0 if out is None else 1, _stream()))
        return g


# ---------------------------------------------------------------------- parameter side (csrc/param_space.cu)
def _raw_vector(state):
    """GslmRawVector over the views of a packed GaussianModelState (solver/gaussian_model_state.py)."""
    ts = (state.xyz_grad, state.features_dc_grad, state.features_rest_grad, state.scaling_grad, state.rotation_grad,
          state.opacity_grad)
    for t in ts:
        if t.numel() and (not t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous()):
            raise RuntimeError("gslm_b200: parameter-space vectors must be contiguous fp32 CUDA tensors")
    return GslmRawVector(*[_ptr(t) for t in ts])


class Activations:
    """Activated parameters at the linearisation point + the two linear maps between raw and activated space."""

    def __init__(self, scales, rotations, rot_norm, opacities, M):
        self.keep = tuple(_f32(t) for t in (scales, rotations, rot_norm, opacities))
        self.P, self.M = int(scales.shape[0]), int(M)
        self.c = GslmActivations(self.P, self.M, *[_ptr(t) for t in self.keep])

    def jvp(self, v_state, shs_t, scales_t, rotations_t, opacities_t):
        """Tangents of (shs, scales, rotations, opacities) for the raw tangent `v_state`, written into the given
        tensors (the tangent of means3D is v_state.xyz_grad itself)."""
        rv = _raw_vector(v_state)
        av = GslmActVector(None, _ptr(shs_t), _ptr(scales_t), _ptr(rotations_t), _ptr(opacities_t))
        _check(lib().gslm_activation_jvp(C.addressof(self.c), C.addressof(rv), C.addressof(av), _stream()))

    def vjp(self, means3D_g, shs_g, scales_g, rotations_g, opacities_g, out_state, v_state=None, damp6=None,
            g_lo=0, g_hi=None):
        """out_state <- activation VJP of the activated-space gradient (+ damp6[k] * v_state per group), for the
        Gaussians [g_lo, g_hi) (default: all)."""
        av = GslmActVector(*[_ptr(t) for t in (means3D_g, shs_g, scales_g, rotations_g, opacities_g)])
        ov = _raw_vector(out_state)
        g_hi = self.P if g_hi is None else int(g_hi)
        if v_state is None:
            _check(lib().gslm_activation_vjp_range(C.addressof(self.c), C.addressof(av), None, None, C.addressof(ov),
                                                   int(g_lo), g_hi, _stream()))
        else:
            vv = _raw_vector(v_state)
            d = (C.c_float * 6)(*[float(x) for x in damp6])
            _check(lib().gslm_activation_vjp_range(C.addressof(self.c), C.addressof(av), C.addressof(vv), d,
                                                   C.addressof(ov), int(g_lo), g_hi, _stream()))


def multimem_allreduce(multicast_ptr, n_floats, rank, world):
    """gslm_multimem_allreduce_f32 on the current stream (see include/gslm_b200.h)."""
    _check(lib().gslm_multimem_allreduce_f32(int(multicast_ptr), int(n_floats), int(rank), int(world), _stream()))


def densification_stats(radii, viewspace_grad, max_radii2D, xyz_gradient_accum, denom):
    """gslm_densification_stats: the bookkeeping of train_jvp.py:311-316 in one pass (in place on the three
    accumulators).  radii: int32 [P]; viewspace_grad: [P,3] float32 or None."""
    P = int(radii.shape[0])
    if radii.dtype != torch.int32:
        radii = radii.to(torch.int32)
    vg = _f32(viewspace_grad)
    for t in (max_radii2D, xyz_gradient_accum, denom):
        if t is not None and not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.numel() == P):
            raise RuntimeError("gslm_b200: densification accumulators must be contiguous float32 CUDA tensors of P elements")
    _check(lib().gslm_densification_stats(P, _ptr(radii.contiguous()), _ptr(vg), _ptr(max_radii2D),
                                          _ptr(xyz_gradient_accum), _ptr(denom), _stream()))


def timing_enable(on=True):
    lib().gslm_timing_enable(1 if on else 0)


def timing_collect():
    """{stage name: (total ms, launches)} since the previous collect; synchronises the device."""
    L = lib()
    n = L.gslm_timing_num_stages()
    ms = (C.c_float * n)()
    cnt = (C.c_int32 * n)()
    _check(L.gslm_timing_collect(C.addressof(ms), C.addressof(cnt)))
    return {L.gslm_timing_stage_name(i).decode(): (float(ms[i]), int(cnt[i])) for i in range(n)}
