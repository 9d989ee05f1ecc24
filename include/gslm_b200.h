/* =============================================================================
 * gslm_b200.h -- C ABI of the B200-native Gaussian-splatting LM hot path.
 *
 * Drop-in boundary for the reference's torch extension `diff_gaussian_rasterization._C`
 * (submodules/diff-gaussian-rasterization, "DGR" below).  Plain pointers and sizes
 * only; every pointer is a DEVICE pointer unless stated; nothing here owns device
 * memory: outputs are caller-allocated, scratch comes from the caller's allocator
 * callback (mirrors resizeFunctional, DGR/rasterize_points.cu:28-34).
 *
 * All functions return 0 on success, non-zero on failure; gslm_last_error() gives the
 * message (thread-local).  Kernels are launched on `stream` (a cudaStream_t passed as
 * void*); the only host synchronisations are two 4-byte read-backs inside gslm_forward /
 * gslm_forward_jvp -- num_rendered R (the reference has the same one,
 * DGR/cuda_rasterizer/rasterizer_impl.cu:325) and num_cached Rc, the number of instances the
 * forward pass actually blended, which sizes the compact linearisation cache -- plus a full
 * sync when camera.debug != 0 (DGR/cuda_rasterizer/auxiliary.h:262-269).  The derivative
 * entry points (backward, jvp_cached, jtjv) never synchronise.
 *
 * Two instance counts cross the ABI: num_rendered = R = sum of tiles_touched, the reference's
 * num_rendered (rasterize_points.cu:117); num_cached = Rc <= R, the instances kept for the
 * derivative passes (all others have exactly zero tangent and gradient).
 * ========================================================================== */
#ifndef GSLM_B200_H
#define GSLM_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Gaussian parameters as the reference binding receives them
 * (DGR/rasterize_points.cu:36-57): activations already applied by the caller. */
typedef struct GslmGaussians {
    int32_t P;                   /* number of Gaussians                                   */
    int32_t sh_degree;           /* active SH degree D (0..3)                              */
    int32_t M;                   /* SH coefficients stored per channel (sh.size(1)), 0 if none */
    float scale_modifier;
    const float* means3D;        /* [P,3]                                                 */
    const float* opacities;      /* [P]                                                   */
    const float* scales;         /* [P,3]   or NULL when cov3D_precomp is given           */
    const float* rotations;      /* [P,4]   (r,x,y,z), NOT normalised by the kernels      */
    const float* cov3D_precomp;  /* [P,6]   or NULL                                       */
    const float* shs;            /* [P,M,3] or NULL when colors_precomp is given          */
    const float* colors_precomp; /* [P,3]   or NULL                                       */
} GslmGaussians;

/* One camera = the per-view part of GaussianRasterizationSettings
 * (DGR/diff_gaussian_rasterization/__init__.py:287-300). Matrices are the 16 floats
 * of the transposed (row-vector) torch tensors, i.e. maths element (i,j) at [i + 4 j]. */
typedef struct GslmCamera {
    int32_t width, height;
    float tan_fovx, tan_fovy;
    const float* bg;             /* [3]  */
    const float* viewmatrix;     /* [16] */
    const float* projmatrix;     /* [16] */
    const float* campos;         /* [3]  */
    int32_t prefiltered, antialiasing, debug;
    /* 0: bin every tile of the 3-sigma bounding square like the reference (auxiliary.h:60-70): num_rendered, point_list
     *    and ranges match rasterizer_impl.cu:321-364 bit for bit.
     * 1: exact tile culling -- a (tile, Gaussian) instance that provably stays below alpha = 1/255 at every pixel of
     *    the tile is not emitted.  Image, inverse depth, radii, final_T, the compact linearisation cache and hence every
     *    derivative are bit-identical to mode 0; num_rendered, the sort and the blend's list walk shrink. */
    int32_t tile_culling;
} GslmCamera;

/* Tangents of the camera-side inputs for the forward-mode entry points
 * (the `_grad` partners of DGR/rasterize_points.cu:143-180).  NULL / 0 => zero tangent. */
typedef struct GslmCameraTangent {
    float tan_fovx, tan_fovy, scale_modifier;
    const float* bg;
    const float* viewmatrix;
    const float* projmatrix;
    const float* campos;
} GslmCameraTangent;

/* Gradient sinks = the 8-tuple returned by RasterizeGaussiansBackwardCUDA
 * (DGR/rasterize_points.cu:300-399).  Any pointer may be NULL (not wanted).
 * accumulate != 0: add into the buffers (batched multi-view sum, replaces the Python
 * `+=` loop of DGR/diff_gaussian_rasterization/batch_render.py:330-346);
 * accumulate == 0: every element of every non-NULL buffer is overwritten. */
typedef struct GslmGrads {
    float* dL_dmeans2D;   /* [P,3] */
    float* dL_dcolors;    /* [P,3]   (only meaningful with colors_precomp) */
    float* dL_dopacity;   /* [P]   */
    float* dL_dmeans3D;   /* [P,3] */
    float* dL_dcov3D;     /* [P,6] */
    float* dL_dsh;        /* [P,M,3] */
    float* dL_dscales;    /* [P,3] */
    float* dL_drotations; /* [P,4] */
    int32_t accumulate;
} GslmGrads;

/* Scratch allocator: must return a device pointer to at least `bytes` bytes, 128-byte
 * aligned, that stays valid until the caller frees it.  `which`: 0 geometry, 1 binning,
 * 2 image buffer -- the three opaque buffers of the reference
 * (geomBuffer / binningBuffer / imgBuffer, DGR/rasterize_points.cu:80-88) -- 3 = sort scratch
 * (keys, radix-sort temporaries, uncompacted records and hit masks) that may be freed or reused as soon as the
 * call returns (the reference trims the same data off its binning buffer after the forward pass,
 * rasterizer_impl.cu:383-388), -1 = anonymous scratch for the duration of the call. */
typedef void* (*GslmAllocFn)(void* ctx, int32_t which, size_t bytes);
enum { GSLM_BUF_GEOM = 0, GSLM_BUF_BINNING = 1, GSLM_BUF_IMAGE = 2, GSLM_BUF_BINNING_TEMP = 3 };

const char* gslm_last_error(void);
int32_t gslm_abi_version(void);
/* sizeof() of an ABI struct by name ("GslmCamera", ...; 0 for an unknown name) and the byte offset of a field of
 * GslmCamera: lets a binding written in another language check its mirror of the layouts before the first call
 * (tests/test_abi.py does it for the ctypes mirrors). */
size_t gslm_sizeof(const char* struct_name);
size_t gslm_camera_offsetof(const char* field_name);

/* Buffer sizes (bytes) -- replaces required<GeometryState/ImageState/BinningState>
 * (DGR/cuda_rasterizer/rasterizer_impl.h:29-43). */
size_t gslm_geom_bytes(int32_t P);
size_t gslm_image_bytes(int32_t width, int32_t height);
size_t gslm_binning_bytes(int64_t num_cached);        /* persistent compact cache: 92 B / cached instance */
size_t gslm_binning_temp_bytes(int32_t P, int64_t num_rendered); /* sort scratch: ~105 B / rendered instance, dead after forward */

/* Replaces _C.mark_visible (DGR/rasterize_points.cu:401-420). present: uint8 [P]. */
int32_t gslm_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                          uint8_t* present, void* stream);

/* Replaces _C.rasterize_gaussians (DGR/rasterize_points.cu:36-141;
 * CudaRasterizer::Rasterizer::forward, DGR/cuda_rasterizer/rasterizer_impl.cu:239-391).
 * out_color [3,H,W], out_invdepth [1,H,W], radii int32 [P], *num_rendered host int. */
int32_t gslm_forward(const GslmGaussians* g, const GslmCamera* cam, GslmAllocFn alloc, void* alloc_ctx,
                     float* out_color, float* out_invdepth, int32_t* radii, int32_t* num_rendered,
                     int32_t* num_cached, void* stream);

/* The same forward pass WITHOUT the linearisation cache: image, inverse depth, radii and final_T only (the
 * binning buffer is never requested from the allocator, n_contrib is not written, there is one host read-back
 * instead of two).  For evaluations that are never differentiated -- the back-tracking line search of
 * train_jvp.py:262-279 renders 7 x 50 views per LM step this way, and so does inference.  The reference has no
 * such split (its forward always keeps the state for backward).  Buffers left by this call must NOT be handed to
 * gslm_backward / gslm_jvp_cached / gslm_jtjv*. */
int32_t gslm_render(const GslmGaussians* g, const GslmCamera* cam, GslmAllocFn alloc, void* alloc_ctx,
                    float* out_color, float* out_invdepth, int32_t* radii, int32_t* num_rendered, void* stream);

/* n_views render-only passes of the same Gaussians in one call with ONE host read-back: every view is preprocessed
 * and scanned first, the n_views instance counts come back together, then binning and blending run view after view
 * with nothing but kernel launches in between (the reference's forward synchronises once per view,
 * rasterizer_impl.cu:325, and its batch wrapper loops over views in Python,
 * diff_gaussian_rasterization/batch_render.py:171-213).  This is the evaluation the back-tracking line search repeats
 * for every step size (train_jvp.py:262-279).  cams / out_color / num_rendered: HOST arrays of n_views entries;
 * out_invdepth and radii may be NULL (or hold NULL entries): the radii then live in the geometry allocation.
 * Independent views are issued on up to 4 internal streams ("lanes", forked from and joined to `stream` with events;
 * GSLM_RENDER_LANES overrides) so that the many short launches of small scenes overlap.  The allocator is asked once
 * for GSLM_BUF_GEOM (all views) and, per lane k, for GSLM_BUF_IMAGE | (k << 8) (largest view, reused) and
 * GSLM_BUF_BINNING_TEMP | (k << 8) (largest view).  Results are bit-identical to n_views gslm_render calls. */
int32_t gslm_render_batch(const GslmGaussians* g, int32_t n_views, const GslmCamera* cams, GslmAllocFn alloc,
                          void* alloc_ctx, float* const* out_color, float* const* out_invdepth, int32_t* const* radii,
                          int32_t* num_rendered, void* stream);

/* The instance counts (num_rendered) of n_views cameras without rendering them: preprocess + (culling) + scan per view
 * and ONE host synchronisation for all counts.  What the multi-GPU view assignment balances on ("greedy
 * balancing on per-view R").  The allocator is asked once for anonymous scratch (which = -1). */
int32_t gslm_count_instances(const GslmGaussians* g, int32_t n_views, const GslmCamera* cams, GslmAllocFn alloc,
                             void* alloc_ctx, int32_t* num_rendered, void* stream);

/* What the LM solver derives from each rendered image (gaussian_renderer/batch_render.py:118 clamp(0,1);
 * solver/batch_training_loss.py:10-17 alpha mask and residual), produced by the forward batch in one pass per view
 * instead of ~8 torch elementwise kernels.  With m = 1[0 <= colour <= 1] (torch's clamp derivative), am = alpha mask:
 *   out_image  = clamp(colour, 0, 1) [* am if apply_alpha_mask] [- gt if subtract_gt]   (plane / row strides in floats)
 *   out_mask   = m [* am if apply_alpha_mask]                                          (plane / row strides in floats)
 *   out_weight = weight_scale * out_mask^2                                             (contiguous [3,H,W])
 * Any output pointer may be NULL. */
typedef struct GslmResidualEpilogue {
    const float* gt;          /* [3,H,W] with strides gt_plane / gt_row, or NULL                              */
    const float* alpha_mask;  /* [H,W] with row stride alpha_row, or NULL                                     */
    float* out_image;
    float* out_mask;
    float* out_weight;
    int64_t gt_plane, gt_row, alpha_row, image_plane, image_row, mask_plane, mask_row;
    float weight_scale;
    int32_t apply_alpha_mask, subtract_gt;
} GslmResidualEpilogue;

/* n_views LINEARISING forward passes (gslm_forward) of the same Gaussians in one call: the twin of gslm_render_batch
 * for the passes that are differentiated afterwards (LinearSolverFunctions.linearize; the reference renders its views
 * one by one with a host synchronisation each, rasterizer_impl.cu:325, DGR/diff_gaussian_rasterization/batch_render.py:
 * 171-213).  Every view is preprocessed and scanned first and all instance counts come back with ONE synchronisation;
 * then the views are binned and blended in rounds on up to 4 internal streams ("lanes", as in gslm_render_batch), and
 * the host waits for a view's compact count only after the whole round is queued, so the device never idles on a
 * read-back.  The allocator is called with
 *   which = kind | (view << 8)   for GSLM_BUF_GEOM, GSLM_BUF_IMAGE, GSLM_BUF_BINNING (one each per view), and
 *   which = GSLM_BUF_BINNING_TEMP | (lane << 8)   for the sort scratches (dead when the call returns).
 * out_color / radii / num_rendered / num_cached: HOST arrays of n_views entries (radii entries must be non-NULL);
 * out_invdepth and epilogues may be NULL.  Buffers and results are bit-identical to n_views gslm_forward calls. */
int32_t gslm_forward_batch(const GslmGaussians* g, int32_t n_views, const GslmCamera* cams, GslmAllocFn alloc,
                           void* alloc_ctx, float* const* out_color, float* const* out_invdepth,
                           int32_t* const* radii, const GslmResidualEpilogue* epilogues, int32_t* num_rendered,
                           int32_t* num_cached, void* stream);

/* Replaces _C.rasterize_gaussians_jvp (DGR/rasterize_points.cu:143-298; forwardJvp,
 * DGR/cuda_rasterizer/rasterizer_jvp_impl.h:125-325).  `gt` holds the tangent pointers in
 * the same fields as `g` (NULL => zero tangent; P/sh_degree/M/scale_modifier ignored,
 * the scale_modifier tangent travels in camt).  The buffers it leaves behind are
 * accepted by gslm_backward, like the reference's. */
int32_t gslm_forward_jvp(const GslmGaussians* g, const GslmGaussians* gt, const GslmCamera* cam,
                         const GslmCameraTangent* camt, GslmAllocFn alloc, void* alloc_ctx, float* out_color,
                         float* out_invdepth, int32_t* radii, float* out_color_tangent, float* out_invdepth_tangent,
                         int32_t* num_rendered, int32_t* num_cached, void* stream);

/* Replaces _C.rasterize_gaussians_backward (DGR/rasterize_points.cu:300-399;
 * Rasterizer::backward, DGR/cuda_rasterizer/rasterizer_impl.cu:395-500).
 * dL_dinvdepth may be NULL.  `scratch` = gslm_backward_scratch_bytes(num_cached) bytes. */
size_t gslm_backward_scratch_bytes(int64_t num_cached);
int32_t gslm_backward(const GslmGaussians* g, const GslmCamera* cam, const int32_t* radii, const void* geom_buffer,
                      const void* binning_buffer, const void* image_buffer, int32_t num_cached,
                      const float* dL_dcolor, const float* dL_dinvdepth, void* scratch, GslmGrads* grads,
                      void* stream);

/* ---- additive entry points: the linearisation-cached LM / CG path -------------------
 * A "linearised view" is the triple of buffers left by gslm_forward for one camera at the
 * current parameters.  The functions below reuse it, so a CG iteration costs no
 * preprocess-primal / scan / sort / ranges work and no host synchronisation. */

/* J v for one linearised view: image tangent [3,H,W] (+ optional invdepth tangent [1,H,W])
 * for Gaussian-parameter tangents `gt` (camera tangents are zero on this path,
 * DGR/diff_gaussian_rasterization/utils.py:18-19). scratch: gslm_jvp_scratch_bytes(P). */
size_t gslm_jvp_scratch_bytes(int32_t P);
int32_t gslm_jvp_cached(const GslmGaussians* g, const GslmGaussians* gt, const GslmCamera* cam, const int32_t* radii,
                        const void* geom_buffer, const void* binning_buffer, const void* image_buffer,
                        int32_t num_cached, void* scratch, float* out_color_tangent, float* out_invdepth_tangent,
                        void* stream);

/* gslm_jvp_cached for callers that cannot supply device scratch.  (No counterpart in the reference: its binding
 * always allocates through torch.)  Who needs it: an embedder that links this library from C, Go or Java and has no
 * device allocator at hand; a process whose tensor framework has already taken all of HBM into its own pool, so a
 * plain cudaMalloc fails; a viewer whose frame loop must not hit the device-wide synchronisation cudaMalloc/cudaFree
 * bring.  Same arguments and same result as gslm_jvp_cached, bit for bit, minus `scratch`: the per-Gaussian tangent
 * records (48 B each) go to an array built into the library, which every GPU that loads the library holds until the
 * process exits (GSLM_NOALLOC_MAX_GAUSSIANS x 48 B = 12 MiB).
 *  - g->P > GSLM_NOALLOC_MAX_GAUSSIANS is refused with an error naming the limit; nothing is launched.
 *  - The function calls no cudaMalloc / cudaFree and no allocating library routine.
 *  - The first call on each device looks the array's device address up (cudaGetSymbolAddress, with that device
 *    current -- so the device of `stream` must be the current one).  Make that first call outside any stream capture.
 *  - All calls on a device share the one array, and it keeps whatever the last call wrote.  Nothing in it needs
 *    clearing between calls (each call writes every record it reads), but two calls that can overlap on the GPU --
 *    different streams, or two host threads without ordering -- race on it.  Issue them on one stream, or order
 *    them with events. */
#define GSLM_NOALLOC_MAX_GAUSSIANS 262144
int32_t gslm_jvp_cached_noalloc(const GslmGaussians* g, const GslmGaussians* gt, const GslmCamera* cam,
                                const int32_t* radii, const void* geom_buffer, const void* binning_buffer,
                                const void* image_buffer, int32_t num_cached, float* out_color_tangent,
                                float* out_invdepth_tangent, void* stream);

/* Fused Gauss-Newton product for one linearised view, J never materialised:
 *   grads (+)= J^T diag(w) J v      with per-pixel weight image `weight` [3,H,W]
 * (w = residual multiplicity x clamp/alpha mask squared; solver/batch_training_loss.py:10-31,
 * gaussian_renderer/batch_render.py:118), NULL => w = 1.
 * out_vJJv (device float[1], may be NULL) accumulates v^T J^T W J v = sum w * (Jv)^2, the
 * p^T A p that CG needs (solver/conjugate_gradient.py:28), for free.
 * scratch: gslm_jtjv_scratch_bytes(P, num_cached). */
size_t gslm_jtjv_scratch_bytes(int32_t P, int64_t num_cached);
int32_t gslm_jtjv(const GslmGaussians* g, const GslmGaussians* gt, const GslmCamera* cam, const int32_t* radii,
                  const void* geom_buffer, const void* binning_buffer, const void* image_buffer,
                  int32_t num_cached, const float* weight, void* scratch, GslmGrads* grads, float* out_vJJv,
                  void* stream);

/* The same product summed over n_views linearised views in one call:
 *   grads (+)= sum_v J_v^T diag(w_v) J_v v,     out_vJJv += sum_v v^T J_v^T W_v J_v v.
 * All array arguments are HOST arrays of n_views entries (device pointers inside).  Views are processed in chunks:
 * ONE tangent-preprocess launch and ONE fused tile launch per chunk of up to 16 views, then the per-Gaussian adjoint
 * view by view, in order (deterministic sum).  This is what solver.LinearSolverFunctions.normal_matvec calls once per
 * CG iteration (solver/conjugate_gradient.py:28); it replaces n_views round trips through matvec / matvec_T
 * (solver/solver_functions.py:74-123).  grads->accumulate == 0: the first view overwrites, the others add.
 * scratch: gslm_jtjv_batch_scratch_bytes(P, n_views, num_cached) bytes. */
size_t gslm_jtjv_batch_scratch_bytes(int32_t P, int32_t n_views, const int32_t* num_cached);
int32_t gslm_jtjv_batch(const GslmGaussians* g, const GslmGaussians* gt, int32_t n_views, const GslmCamera* cams,
                        const int32_t* const* radii, const void* const* geom_buffers,
                        const void* const* binning_buffers, const void* const* image_buffers,
                        const int32_t* num_cached, const float* const* weights, void* scratch, size_t scratch_bytes,
                        GslmGrads* grads, float* out_vJJv, void* stream);

/* The same call with the per-Gaussian adjoint of the LAST chunk of views cut into n_slices ranges of Gaussian ids
 * (multiples of 256).  After the launch that completes slice k -- rows [g_lo, g_hi) of every sink are final from then
 * on, in stream order -- the HOST callback on_slice(ctx, k, g_lo, g_hi) runs: the multi-GPU caller records an event and
 * starts the allreduce and the activation VJP of that slice on a communication stream while the next slice is still
 * being produced (SURVEY 8e "Overlap": the reference has no counterpart, its sum over views is the Python `+=` of
 * DGR/diff_gaussian_rasterization/batch_render.py:330-346).  n_slices <= 1: one callback for [0, P) at the end
 * (on_slice may be NULL).  Results are bit-identical to the unsliced call. */
typedef void (*GslmSliceFn)(void* ctx, int32_t slice, int32_t g_lo, int32_t g_hi);
int32_t gslm_jtjv_batch_sliced(const GslmGaussians* g, const GslmGaussians* gt, int32_t n_views, const GslmCamera* cams,
                               const int32_t* const* radii, const void* const* geom_buffers,
                               const void* const* binning_buffers, const void* const* image_buffers,
                               const int32_t* num_cached, const float* const* weights, void* scratch,
                               size_t scratch_bytes, GslmGrads* grads, float* out_vJJv, int32_t n_slices,
                               GslmSliceFn on_slice, void* slice_ctx, void* stream);

/* J v and J^T u for n_views linearised views per call (colour only), <= 16 views per kernel launch: what
 * LinearSolverFunctions.matvec / matvec_T issue (solver/solver_functions.py:75-113; the reference's CGLS,
 * solver/conjugate_gradient.py:51-127, calls both every iteration, and the L1 + SSIM normal product needs the
 * image-space vector between them).  out_color_tangent[v] / dL_dcolor[v]: [3,H_v,W_v] device images.
 * gslm_backward_batch sums the views' gradients into `grads` in view order (deterministic). */
size_t gslm_jvp_batch_scratch_bytes(int32_t P, int32_t n_views);
int32_t gslm_jvp_batch(const GslmGaussians* g, const GslmGaussians* gt, int32_t n_views, const GslmCamera* cams,
                       const void* const* geom_buffers, const void* const* binning_buffers,
                       const void* const* image_buffers, const int32_t* num_cached, void* scratch, size_t scratch_bytes,
                       float* const* out_color_tangent, void* stream);
size_t gslm_backward_batch_scratch_bytes(int32_t n_views, const int32_t* num_cached);
int32_t gslm_backward_batch(const GslmGaussians* g, int32_t n_views, const GslmCamera* cams,
                            const int32_t* const* radii, const void* const* geom_buffers,
                            const void* const* binning_buffers, const void* const* image_buffers,
                            const int32_t* num_cached, const float* const* dL_dcolor, void* scratch,
                            size_t scratch_bytes, GslmGrads* grads, void* stream);

/* gslm_backward_batch with the sliced adjoint of gslm_jtjv_batch_sliced (same callback contract). */
int32_t gslm_backward_batch_sliced(const GslmGaussians* g, int32_t n_views, const GslmCamera* cams,
                                   const int32_t* const* radii, const void* const* geom_buffers,
                                   const void* const* binning_buffers, const void* const* image_buffers,
                                   const int32_t* num_cached, const float* const* dL_dcolor, void* scratch,
                                   size_t scratch_bytes, GslmGrads* grads, int32_t n_slices, GslmSliceFn on_slice,
                                   void* slice_ctx, void* stream);

/* ---- residual side: the two per-pixel loss blocks and their derivatives --------------------------------------
 * Replaces compute_batch_loss_block (solver/batch_training_loss.py:10-31) with the SSIM map of ssim_per_pixel /
 * _ssim_per_pixel (utils/loss_utils.py:91-117; 11x11 Gaussian window, sigma 1.5, zero padding, C1 = 0.01^2,
 * C2 = 0.03^2) for disable_ssim=False:
 *     r1 = alphas[i] sqrt(|x - gt| + 1e-6),   r2 = betas[i] sqrt(|1 - SSIM(x, gt)| + 1e-6),   x = images * alpha_masks
 * and what PyTorch derives from it in the reference: the forward-AD tangent used by LinearSolverFunctions.matvec
 * (solver/solver_functions.py:75-90) and the autograd pull-back used by matvec_T (:92-113).  It also stands in for
 * the reference's fused-ssim extension pair fusedssim / fusedssim_backward (utils/loss_utils.py:24-38), which has
 * no forward-mode rule.  All arrays are [B,3,H,W] fp32 device tensors (zero padded to the batch's Hmax x Wmax as
 * batch_training_loss.py:54-66 lays them out); alpha_masks is [B,1,H,W] or NULL; alphas / betas are DEVICE arrays
 * of B per-image factors (sqrt((1-lambda)/n_i), sqrt(lambda/n_i), batch_training_loss.py:68-77).
 * `lin` is an opaque per-pixel linearisation record (gslm_loss_lin_bytes) written by gslm_loss_blocks (NULL: not
 * wanted) and read by the two derivative calls, which are linear maps at that linearisation point. */
typedef struct GslmLossBatch {
    int32_t B, H, W;
    const float* images;       /* [B,3,H,W] rendered images (already clamped by the renderer glue)   */
    const float* alpha_masks;  /* [B,1,H,W] or NULL                                                  */
    const float* gt;           /* [B,3,H,W]                                                          */
    const float* alphas;       /* [B] device                                                         */
    const float* betas;        /* [B] device                                                         */
    const float* tangent_masks;/* [B,3,H,W] or NULL: d images / d (what the caller differentiates), e.g. the 0/1 mask of
                                * the renderer's clamp(0,1) (gaussian_renderer/batch_render.py:118): multiplies
                                * t_images on the way into gslm_loss_blocks_jvp and g_images on the way out of
                                * gslm_loss_blocks_vjp; ignored by gslm_loss_blocks                       */
} GslmLossBatch;
size_t gslm_loss_lin_bytes(int32_t B, int32_t H, int32_t W);       /* 16 B per pixel and channel */
int32_t gslm_loss_blocks(const GslmLossBatch* batch, float* r1, float* r2, void* lin, void* stream);
/* (t1, t2) = d(r1, r2)/d images . t_images */
int32_t gslm_loss_blocks_jvp(const GslmLossBatch* batch, const void* lin, const float* t_images, float* t1, float* t2,
                             void* stream);
/* g_images (+)= (d(r1, r2)/d images)^T (u1, u2); accumulate != 0 adds into g_images */
int32_t gslm_loss_blocks_vjp(const GslmLossBatch* batch, const void* lin, const float* u1, const float* u2,
                             float* g_images, int32_t accumulate, void* stream);

/* ---- parameter side: the activation layer, differentiated ---------------------------------------------------
 * The solver's vectors live in RAW parameter space (GaussianModelState: xyz, features_dc, features_rest, scaling,
 * rotation, opacity; solver/gaussian_model_state.py:4-60) while the rasterizer consumes ACTIVATED inputs
 * (get_scaling = exp, get_rotation = normalize, get_opacity = sigmoid, get_features = cat(dc, rest);
 * scene/gaussian_model.py:192-220).  The reference lets PyTorch differentiate those activations (forward AD inside
 * make_dual, autograd in loss.backward: solver/solver_functions.py:75-113); here each direction is one pass over the
 * 59 P floats.  `GslmActivations` holds the activated values at the linearisation point. */
typedef struct GslmActivations {
    int32_t P, M;             /* Gaussians; SH coefficients per channel (features_dc + features_rest rows)  */
    const float* scales;      /* [P,3] exp(raw scaling)                                                      */
    const float* rotations;   /* [P,4] raw quaternion / its norm                                             */
    const float* rot_norm;    /* [P]   norm of the raw quaternion                                            */
    const float* opacities;   /* [P]   sigmoid(raw opacity)                                                  */
} GslmActivations;
typedef struct GslmRawVector {   /* a GaussianModelState without the exposure block */
    float* xyz;            /* [P,3]     */
    float* features_dc;    /* [P,1,3]   */
    float* features_rest;  /* [P,M-1,3] (may be NULL when M == 1) */
    float* scaling;        /* [P,3]     */
    float* rotation;       /* [P,4]     */
    float* opacity;        /* [P,1]     */
} GslmRawVector;
typedef struct GslmActVector {   /* tangents or gradients of the rasterizer inputs */
    float* means3D;        /* [P,3]   (JVP output: may be NULL, the tangent of means3D IS v.xyz) */
    float* shs;            /* [P,M,3] */
    float* scales;         /* [P,3]   */
    float* rotations;      /* [P,4]   */
    float* opacities;      /* [P]     */
} GslmActVector;
/* t = d(activated inputs)/d(raw parameters) . v */
int32_t gslm_activation_jvp(const GslmActivations* act, const GslmRawVector* v, const GslmActVector* t, void* stream);
/* out = (d(activated)/d(raw))^T g  [+ damp6[k] * v per parameter group k = xyz, dc, rest, scaling, rotation, opacity
 * when v != NULL; damp6 is a HOST array]: the tail of J^T J v + D v (solver/conjugate_gradient.py:3-49 with the
 * damping of train_jvp.py:237-260) */
int32_t gslm_activation_vjp(const GslmActivations* act, const GslmActVector* g, const GslmRawVector* v,
                            const float* damp6, const GslmRawVector* out, void* stream);
/* the same restricted to the Gaussians [g_lo, g_hi): the per-slice tail of the sliced products above */
int32_t gslm_activation_vjp_range(const GslmActivations* act, const GslmActVector* g, const GslmRawVector* v,
                                  const float* damp6, const GslmRawVector* out, int32_t g_lo, int32_t g_hi,
                                  void* stream);

/* ---- multi-GPU: the sum of the per-rank parameter-space partials over NVSwitch peer memory ----------------------
 * (no reference counterpart: the reference is single-GPU)  `multicast_ptr` is the NVSwitch multicast
 * address of a symmetric allocation of n_floats floats that exists on every rank of the node (the caller creates it,
 * e.g. torch.distributed._symmetric_memory: empty + rendezvous -> multicast_ptr).  Rank `rank` of `world` sums ITS
 * 1/world slice over all ranks with multimem.ld_reduce (the switch adds the copies in the fabric) and writes the sums
 * back into every rank's copy with multimem.st: after all ranks have run it, every copy holds the all-reduced
 * vector.  The call does not synchronise the ranks: bracket it with a cross-rank barrier on the same stream (all
 * partials written before; all slices stored after).  n_floats must be a multiple of 4, the buffer 16-byte aligned. */
int32_t gslm_multimem_allreduce_f32(void* multicast_ptr, size_t n_floats, int32_t rank, int32_t world, void* stream);

/* Densification bookkeeping of the training driver (train_jvp.py:311-316: max_radii2D[vis] = max(max_radii2D[vis],
 * radii[vis]); GaussianModel.add_densification_stats, scene/gaussian_model.py:561-563: xyz_gradient_accum[vis] +=
 * ||viewspace_grad[vis, :2]||, denom[vis] += 1) with vis = radii > 0, in one pass.  radii: int32 [P] (max over the
 * views of the batch), viewspace_grad: [P,3] = dL_dmeans2D summed over the batch (NULL: only the radii are tracked),
 * max_radii2D / xyz_gradient_accum / denom: float [P] accumulators (max_radii2D may be NULL). */
int32_t gslm_densification_stats(int32_t P, const int32_t* radii, const float* viewspace_grad, float* max_radii2D,
                                 float* xyz_gradient_accum, float* denom, void* stream);

/* ---- optional per-stage device timing (no reference counterpart; bench.py's live roofline numbers) ----
 * When enabled, each stage brackets its kernels with a cudaEvent pair on the launching stream.
 * gslm_timing_collect synchronises the device and returns, per stage, the summed elapsed milliseconds and the
 * number of launches since the previous collect (arrays of gslm_timing_num_stages() entries). */
int32_t gslm_timing_num_stages(void);
const char* gslm_timing_stage_name(int32_t stage);
void gslm_timing_enable(int32_t on);
int32_t gslm_timing_collect(float* ms, int32_t* counts);

/* FP32 peak probe (no reference counterpart): the denominator of the tile kernels' roofline, measured in the run that
 * quotes it.  mode 0: scalar FFMA on independent chains (the FP32 FLOP peak); 1: packed FFMA2 with three distinct 64-bit
 * register operands (what the tile sweeps issue); 2: mode 1 + one MUFU.EX2 per 12 FFMA2 (the sweeps' mix).
 * scratch: gslm_fp32_probe_scratch_bytes() device bytes.  *out_tflops: FMA = 2 flops; best of 3 launches of ~10 ms each,
 * timed with CUDA events on `stream`; synchronises the stream. */
size_t gslm_fp32_probe_scratch_bytes(void);
int32_t gslm_fp32_peak_probe(int32_t mode, void* scratch, float* out_tflops, float* out_ms, void* stream);

/* Introspection for tests (bit-exact binning checks): device pointers INTO the buffers. */
typedef struct GslmBufferView {
    const uint32_t* point_list;   /* [R] sorted Gaussian ids (BinningState::point_list; lives in the sort scratch) */
    const uint32_t* tile_keys_sorted; /* [R] tile id of every sorted instance (sort scratch); the reference's 64-bit key is
                                       * tile << 32 | depth bits of point_list[i]                        */
    const uint32_t* ranges;       /* [T,2] per-tile [start,end) in the full list (ImageState::ranges)   */
    const uint32_t* n_contrib;    /* [H*W] reference numbering                                          */
    const float* final_T;         /* [H*W]                                                              */
    const uint32_t* tiles_touched;/* [P]                                                                */
    const uint32_t* point_offsets;/* [P] inclusive scan                                                 */
    const float* depths;          /* [P]                                                                */
    const float* means2D;         /* [P,2]                                                              */
    const float* conic_opacity;   /* [P,4]                                                              */
    const float* rgb;             /* [P,3]                                                              */
    const float* cov3D;           /* [P,6]                                                              */
    const uint8_t* clamped;       /* [P] bit c set <=> channel c clamped                                */
    const uint32_t* n_contrib_c;  /* [H*W] last blended position in the compact tile list               */
    const uint32_t* cstart;       /* [T+1] compact tile offsets, cstart[T] = Rc                         */
    const float* records;         /* [Rc,12] compact instance records (derivative form: log2-scaled conic, log2 o) */
    const uint32_t* masks;        /* [Rc,8] hit masks: bit l of word k <=> pixel (row 2k + l/16, column l%16) blended it */
    const uint32_t* cid;          /* [Rc] Gaussian id per compact instance                              */
    const uint32_t* g_off;        /* [P+1] CSR offsets into g_list                                      */
    const uint32_t* g_list;       /* [Rc] compact positions grouped by Gaussian                         */
} GslmBufferView;
int32_t gslm_buffer_view(int32_t P, int32_t width, int32_t height, int64_t num_rendered, int64_t num_cached,
                         const void* geom_buffer, const void* binning_buffer, const void* image_buffer,
                         const void* binning_temp_buffer, GslmBufferView* out);

/* Introspection for tests: the per-Gaussian TANGENT records of one linearised view, i.e. the output of the tangent
 * preprocess (the restatement of preprocessCUDAJvp, DGR/cuda_rasterizer/forward_impl.h:208-367) exactly as the tile
 * sweeps consume it.  out_records: [P,12] floats, row of Gaussian i =
 *   (a1, a2, a3, a4 | a5, o_t / o, r_t, g_t | b_t, (1/depth)_t, 0, 0)   with, for the reference's per-Gaussian
 * tangents (means2D_t = (x_t, y_t), conic_t = (A_t, B_t, C_t), opacity o_t, rgb_t, depth_t) of
 * DGR/tests/render_data/{means2D,conic_opacity,rgb,depth}_grad.csv:
 *   a1 = -(A x_t + B y_t), a2 = -(C y_t + B x_t), a3 = -A_t / 2, a4 = -C_t / 2, a5 = -B_t, (1/depth)_t = -depth_t / depth^2.
 * Only rows of Gaussians that own a cached instance in this view are written (the others are never read).
 * camt may be NULL.  tests/test_gpu_parity.py compares the rows with the reference's golden tangents. */
int32_t gslm_tangent_records(const GslmGaussians* g, const GslmGaussians* gt, const GslmCamera* cam,
                             const GslmCameraTangent* camt, const void* geom_buffer, float* out_records, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GSLM_B200_H */
