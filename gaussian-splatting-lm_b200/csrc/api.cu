// =============================================================================
// api.cu -- the C ABI (include/gslm_b200.h): orchestration of the stages, one function per
// entry point of the reference binding (DGR/rasterize_points.cu) plus the cached CG path.
// No device memory is allocated here; no torch types anywhere.
// =============================================================================
#include "gslm_common.cuh"
#include <cstddef>
#include <cstdlib>
#include <cstring>
#include <pthread.h>
#include <vector>

using namespace gslm;

static TangentPtrs make_tangents(const GslmGaussians* gt, const GslmCameraTangent* ct) {
    TangentPtrs t;
    memset(&t, 0, sizeof(t));
    if (gt) {
        t.means3D = gt->means3D; t.opacities = gt->opacities; t.scales = gt->scales; t.rotations = gt->rotations;
        t.cov3D_precomp = gt->cov3D_precomp; t.shs = gt->shs; t.colors_precomp = gt->colors_precomp;
    }
    if (ct) {
        t.viewmatrix = ct->viewmatrix; t.projmatrix = ct->projmatrix; t.campos = ct->campos;
        t.tan_fovx = ct->tan_fovx; t.tan_fovy = ct->tan_fovy; t.scale_modifier = ct->scale_modifier;
    }
    return t;
}

static int check_inputs(const GslmGaussians* g, const GslmCamera* cam) {
    if (!g || !cam) { set_error("null argument"); return 1; }
    if (g->P < 0 || cam->width <= 0 || cam->height <= 0) { set_error("bad sizes"); return 1; }
    if (g->P > 0) {
        if (!g->means3D || !g->opacities) { set_error("means3D / opacities missing"); return 1; }
        // rasterizer_impl.cu:285-288 / __init__.py:322-326
        if (!g->shs && !g->colors_precomp) { set_error("For non-RGB, provide precomputed Gaussian colors!"); return 1; }
        if (!g->cov3D_precomp && (!g->scales || !g->rotations)) {
            set_error("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!");
            return 1;
        }
        if (g->shs && !g->colors_precomp) {
            const int need = (g->sh_degree + 1) * (g->sh_degree + 1);
            if (g->sh_degree < 0 || g->sh_degree > 3 || g->M < need) { set_error("sh_degree / M mismatch"); return 1; }
        }
    }
    if (!cam->bg || !cam->viewmatrix || !cam->projmatrix) { set_error("camera pointers missing"); return 1; }
    return 0;
}

static int debug_sync(const GslmCamera* cam, cudaStream_t s, const char* what) {
    if (cam->debug) {   // auxiliary.h:262-269 CHECK_CUDA
        cudaError_t e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) { set_error("%s: %s", what, cudaGetErrorString(e)); return 1; }
    }
    return 0;
}

// point_offsets[P-1] (the instance count) of each view of a render batch, gathered for one read-back
static __global__ void k_gather_totals(const char* geom_base, size_t view_stride, size_t offset_of_last, int n_views,
                                       uint32_t* totals) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n_views) totals[v] = *reinterpret_cast<const uint32_t*>(geom_base + v * view_stride + offset_of_last);
}

// What the solver derives from each rendered image, in one pass (the reference does it with ~8 torch elementwise
// kernels per view: clamp, compares, mask products, the residual subtraction; gaussian_renderer/batch_render.py:118,
// solver/batch_training_loss.py:10-17).
static __global__ void k_residual_epilogue(const float* __restrict__ color, int W, int H, GslmResidualEpilogue e) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, c = blockIdx.z;
    if (x >= W) return;
    const float v = color[((size_t)c * H + y) * W + x];
    const float am = e.alpha_mask ? e.alpha_mask[(size_t)y * e.alpha_row + x] : 1.f;
    // d clamp(x, 0, 1) / dx as torch defines it: 1 on the closed interval
    float m = (v >= 0.f && v <= 1.f) ? 1.f : 0.f;
    float img = fminf(fmaxf(v, 0.f), 1.f);
    if (v != v) img = v;      // torch.clamp propagates NaN
    // individually rounded like the torch composition it replaces (a contracted fma would differ in the last bit)
    if (e.apply_alpha_mask) { img = __fmul_rn(img, am); m = __fmul_rn(m, am); }
    if (e.gt && e.subtract_gt) img = __fsub_rn(img, e.gt[(size_t)c * e.gt_plane + (size_t)y * e.gt_row + x]);
    if (e.out_image) e.out_image[(size_t)c * e.image_plane + (size_t)y * e.image_row + x] = img;
    if (e.out_mask) e.out_mask[(size_t)c * e.mask_plane + (size_t)y * e.mask_row + x] = m;
    if (e.out_weight) e.out_weight[((size_t)c * H + y) * W + x] = __fmul_rn(__fmul_rn(e.weight_scale, m), m);
}

static int launch_residual_epilogue(const float* color, int W, int H, const GslmResidualEpilogue& e, cudaStream_t s) {
    if (!e.out_image && !e.out_mask && !e.out_weight) return 0;
    k_residual_epilogue<<<dim3((W + 255) / 256, H, 3), 256, 0, s>>>(color, W, H, e);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

static int clear_sinks(const GslmGrads& acc, int P, int M, cudaStream_t s) {
    const size_t n = (size_t)P;
    struct { float* p; size_t count; } sinks[8] = {
        {acc.dL_dmeans2D, 3 * n}, {acc.dL_dcolors, 3 * n}, {acc.dL_dopacity, n}, {acc.dL_dmeans3D, 3 * n},
        {acc.dL_dcov3D, 6 * n}, {acc.dL_dsh, 3 * n * (size_t)M}, {acc.dL_dscales, 3 * n}, {acc.dL_drotations, 4 * n}};
    for (auto& k : sinks)
        if (k.p && k.count) GSLM_CUDA_OK(cudaMemsetAsync(k.p, 0, k.count * sizeof(float), s));
    return 0;
}

// The per-Gaussian adjoint of one chunk of views (always accumulating: overwrite-mode products clear the sinks
// first).  On the LAST chunk the launch may be cut into n_slices Gaussian ranges with a host callback after
// each (the caller starts the collective / the activation VJP of that slice while the next one is produced).
// the callbacks a product with NO local views owes its caller (same slice boundaries as adjoint_chunk: a rank without
// views must still enter the same sequence of collectives as the others)
static void empty_product_callbacks(int P, int n_slices, GslmSliceFn on_slice, void* slice_ctx) {
    if (!on_slice) return;
    const int nranges = (P + ACT_RANGE - 1) / ACT_RANGE;
    if (n_slices <= 1) { on_slice(slice_ctx, 0, 0, P); return; }
    if (n_slices > nranges) n_slices = nranges;
    for (int k = 0; k < n_slices; k++) {
        const int lo = (int)((int64_t)nranges * k / n_slices), hi = (int)((int64_t)nranges * (k + 1) / n_slices);
        const int64_t g_hi = (int64_t)hi * ACT_RANGE;
        on_slice(slice_ctx, k, lo * ACT_RANGE, (int32_t)(g_hi < P ? g_hi : P));
    }
}

static int adjoint_chunk(const GslmGaussians& g, int n, const GslmCamera* cams, const int32_t* const* radii,
                         const GeomView* gv, const float4* const* inst, const GslmGrads& acc, bool last,
                         int n_slices, GslmSliceFn on_slice, void* slice_ctx, cudaStream_t s) {
    const int nranges = (g.P + ACT_RANGE - 1) / ACT_RANGE;
    if (!last || n_slices <= 1) {
        if (launch_preprocess_bwd_views(g, n, cams, radii, gv, inst, acc, s, 0, nranges)) return 1;
        if (last && on_slice) on_slice(slice_ctx, 0, 0, g.P);
        return 0;
    }
    if (n_slices > nranges) n_slices = nranges;
    for (int k = 0; k < n_slices; k++) {
        const int lo = (int)((int64_t)nranges * k / n_slices), hi = (int)((int64_t)nranges * (k + 1) / n_slices);
        if (launch_preprocess_bwd_views(g, n, cams, radii, gv, inst, acc, s, lo, hi)) return 1;
        if (on_slice) {
            const int64_t g_hi = (int64_t)hi * ACT_RANGE;
            on_slice(slice_ctx, k, lo * ACT_RANGE, (int32_t)(g_hi < g.P ? g_hi : g.P));
        }
    }
    return 0;
}

// Internal streams ("lanes") of the batched render: small scenes leave most of the GPU idle behind ~25 short launches
// per view (CUB scans and sorts), so independent views are issued on several streams and overlap.  Created once per
// host thread and device; forked from / joined to the caller's stream with events, so the call stays stream-ordered for the caller.
struct LanePool {
    static constexpr int MAX = 4;
    cudaStream_t s[MAX] = {};
    cudaEvent_t done[MAX] = {};
    cudaEvent_t fork = nullptr;
    int n = 0;
    int get(int want) {
        if (want > MAX) want = MAX;
        if (!fork && cudaEventCreateWithFlags(&fork, cudaEventDisableTiming) != cudaSuccess) return 0;
        while (n < want) {
            if (cudaStreamCreateWithFlags(&s[n], cudaStreamNonBlocking) != cudaSuccess) break;
            if (cudaEventCreateWithFlags(&done[n], cudaEventDisableTiming) != cudaSuccess) break;
            n++;
        }
        return n < want ? n : want;
    }
};
// One pool per host thread AND device: streams and events belong to the device that was current when they were
// created, so a thread that moves to another GPU must not be handed the first one's lanes.
static LanePool* lane_pool() {
    static constexpr int MAX_DEV = 64;
    static thread_local LanePool p[MAX_DEV];
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEV) return nullptr;
    return &p[dev];
}
static int render_lanes(int n_views) {
    static int cfg = -1;
    if (cfg < 0) {
        const char* e = getenv("GSLM_RENDER_LANES");
        cfg = e ? atoi(e) : 4;
        if (cfg < 1) cfg = 1;
    }
    int k = cfg < n_views ? cfg : n_views;
    if (stage_timing_enabled()) k = 1;      // the per-stage timers bracket one stream
    return k;
}

// small grow-only pinned host buffer for the instance counts of a forward batch (so that the read-backs are truly
// asynchronous and the host can run ahead of the device)
static uint32_t* pinned_counts(size_t n) {
    static thread_local uint32_t* buf = nullptr;
    static thread_local size_t cap = 0;
    if (n > cap) {
        if (buf) cudaFreeHost(buf);
        buf = nullptr;
        cap = 0;
        const size_t want = n < 256 ? 256 : n;
        if (cudaMallocHost(reinterpret_cast<void**>(&buf), want * sizeof(uint32_t)) != cudaSuccess) return nullptr;
        cap = want;
    }
    return buf;
}

extern "C" {

const char* gslm_last_error(void) { return last_error(); }
int32_t gslm_abi_version(void) { return 3; }    // 2: GslmCamera::tile_culling, gslm_render_batch, gslm_sizeof; 3: gslm_forward_batch, gslm_count_instances

size_t gslm_sizeof(const char* n) {
    if (!n) return 0;
#define GSLM_SZ(T) if (!strcmp(n, #T)) return sizeof(T);
    GSLM_SZ(GslmGaussians) GSLM_SZ(GslmCamera) GSLM_SZ(GslmCameraTangent) GSLM_SZ(GslmGrads) GSLM_SZ(GslmLossBatch)
    GSLM_SZ(GslmActivations) GSLM_SZ(GslmRawVector) GSLM_SZ(GslmActVector) GSLM_SZ(GslmBufferView)
    GSLM_SZ(GslmResidualEpilogue)
#undef GSLM_SZ
    return 0;
}
size_t gslm_camera_offsetof(const char* n) {
    if (!n) return (size_t)-1;
#define GSLM_OFF(F) if (!strcmp(n, #F)) return offsetof(GslmCamera, F);
    GSLM_OFF(width) GSLM_OFF(height) GSLM_OFF(tan_fovx) GSLM_OFF(tan_fovy) GSLM_OFF(bg) GSLM_OFF(viewmatrix)
    GSLM_OFF(projmatrix) GSLM_OFF(campos) GSLM_OFF(prefiltered) GSLM_OFF(antialiasing) GSLM_OFF(debug)
    GSLM_OFF(tile_culling)
#undef GSLM_OFF
    return (size_t)-1;
}

size_t gslm_geom_bytes(int32_t P) { return carve_geom(nullptr, P).total; }
size_t gslm_image_bytes(int32_t width, int32_t height) { return carve_image(nullptr, width, height).total; }
size_t gslm_binning_bytes(int64_t Rc) { return carve_binning(nullptr, Rc).total; }
size_t gslm_binning_temp_bytes(int32_t P, int64_t R) { return carve_binning_temp(nullptr, R, P).total; }
size_t gslm_backward_scratch_bytes(int64_t Rc) { return align_up((size_t)Rc * REC_F4 * sizeof(float4)) + 128; }
size_t gslm_jvp_scratch_bytes(int32_t P) { return align_up((size_t)P * REC_F4 * sizeof(float4)) + 128; }
size_t gslm_jtjv_scratch_bytes(int32_t P, int64_t Rc) {
    return gslm_jvp_scratch_bytes(P) + gslm_backward_scratch_bytes(Rc);
}

int32_t gslm_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                          uint8_t* present, void* stream) {
    (void)projmatrix;
    if (P > 0 && (!means3D || !viewmatrix || !present)) { set_error("null argument"); return 1; }
    return launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream);
}

static int32_t forward_impl(const GslmGaussians* g, const GslmCamera* cam, GslmAllocFn alloc, void* alloc_ctx,
                            float* out_color, float* out_invdepth, int32_t* radii, int32_t* num_rendered,
                            int32_t* num_cached, bool record, void* stream) {
    if (check_inputs(g, cam)) return 1;
    if (!alloc || !out_color || !num_rendered || (record && !num_cached) || (g->P > 0 && !radii)) {
        set_error("null argument");
        return 1;
    }
    cudaStream_t s = (cudaStream_t)stream;
    const int P = g->P, W = cam->width, H = cam->height;
    const int T = ((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
    void* gbuf = alloc(alloc_ctx, GSLM_BUF_GEOM, gslm_geom_bytes(P));
    void* ibuf = alloc(alloc_ctx, GSLM_BUF_IMAGE, gslm_image_bytes(W, H));
    if (!gbuf || !ibuf) { set_error("allocator returned null"); return 1; }
    GeomView gv = carve_geom(gbuf, P);
    ImageView iv = carve_image(ibuf, W, H);
    uint32_t R = 0, Rc = 0;
    if (P > 0) {
        if (launch_preprocess(*g, *cam, gv, radii, s)) return 1;
        if (debug_sync(cam, s, "preprocess")) return 1;
        if (cam->tile_culling && run_tile_cull(gv, radii, P, W, H, s)) return 1;
        if (run_scan(gv, P, s)) return 1;
        // rasterizer_impl.cu:325 -- host synchronisation #1 of the uncached path (sizes the sort scratch)
        GSLM_CUDA_OK(cudaMemcpyAsync(&R, gv.point_offsets + (P - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        GSLM_CUDA_OK(cudaStreamSynchronize(s));
    }
    *num_rendered = (int32_t)R;
    void* tbuf = alloc(alloc_ctx, GSLM_BUF_BINNING_TEMP, gslm_binning_temp_bytes(P, R));
    if (!tbuf) { set_error("allocator returned null"); return 1; }
    BinningTemp bt = carve_binning_temp(tbuf, R, P);
    if (run_binning(gv, radii, bt, iv, P, W, H, R, cam->tile_culling != 0, s)) return 1;
    if (debug_sync(cam, s, "binning")) return 1;
    const float* features = g->colors_precomp ? g->colors_precomp : gv.rgb;
    if (launch_blend_fwd(gv, bt, iv, W, H, features, cam->bg, out_color, out_invdepth, record, s)) return 1;
    if (debug_sync(cam, s, "render")) return 1;
    if (!record) return 0;      // render-only: no linearisation cache, no second read-back
    if (run_tile_scan(iv, W, H, s)) return 1;
    // host synchronisation #2: the compact cache is sized by what the forward pass actually blended
    GSLM_CUDA_OK(cudaMemcpyAsync(&Rc, iv.cstart + T, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    GSLM_CUDA_OK(cudaStreamSynchronize(s));
    *num_cached = (int32_t)Rc;
    void* bbuf = alloc(alloc_ctx, GSLM_BUF_BINNING, gslm_binning_bytes(Rc));
    if (!bbuf) { set_error("allocator returned null"); return 1; }
    BinningView bv = carve_binning(bbuf, Rc);
    if (P > 0 && run_compaction(gv, bv, bt, iv, P, W, H, Rc, s)) return 1;
    return debug_sync(cam, s, "compaction");
}

int32_t gslm_forward(const GslmGaussians* g, const GslmCamera* cam, GslmAllocFn alloc, void* alloc_ctx,
                     float* out_color, float* out_invdepth, int32_t* radii, int32_t* num_rendered,
                     int32_t* num_cached, void* stream) {
    return forward_impl(g, cam, alloc, alloc_ctx, out_color, out_invdepth, radii, num_rendered, num_cached, true,
                        stream);
}

int32_t gslm_render(const GslmGaussians* g, const GslmCamera* cam, GslmAllocFn alloc, void* alloc_ctx,
                    float* out_color, float* out_invdepth, int32_t* radii, int32_t* num_rendered, void* stream) {
    return forward_impl(g, cam, alloc, alloc_ctx, out_color, out_invdepth, radii, num_rendered, nullptr, false,
                        stream);
}

int32_t gslm_render_batch(const GslmGaussians* g, int32_t n_views, const GslmCamera* cams, GslmAllocFn alloc,
                          void* alloc_ctx, float* const* out_color, float* const* out_invdepth, int32_t* const* radii,
                          int32_t* num_rendered, void* stream) {
    if (n_views < 0 || (n_views > 0 && (!cams || !out_color || !num_rendered)) || !alloc) {
        set_error("null argument");
        return 1;
    }
    if (n_views == 0) return 0;
    for (int v = 0; v < n_views; v++) {
        if (check_inputs(g, &cams[v])) return 1;
        if (!out_color[v]) { set_error("null argument"); return 1; }
        num_rendered[v] = 0;
    }
    cudaStream_t s = (cudaStream_t)stream;
    const int P = g->P;
    // lanes: views are dealt round robin to K internal streams, each with its own image state and sort scratch
    int K = render_lanes(n_views);
    LanePool* pool_p = lane_pool();
    if (!pool_p) { set_error("no current CUDA device for the render lanes"); return 1; }
    LanePool& pool = *pool_p;
    if (K > 1) K = pool.get(K);
    if (K < 1) K = 1;
    const bool lanes = K > 1;
    auto lane_stream = [&](int v) { return lanes ? pool.s[v % K] : s; };
    auto fork_lanes = [&]() -> int {
        if (!lanes) return 0;
        GSLM_CUDA_OK(cudaEventRecord(pool.fork, s));
        for (int k = 0; k < K; k++) GSLM_CUDA_OK(cudaStreamWaitEvent(pool.s[k], pool.fork, 0));
        return 0;
    };
    auto join_lanes = [&]() -> int {
        if (!lanes) return 0;
        for (int k = 0; k < K; k++) {
            GSLM_CUDA_OK(cudaEventRecord(pool.done[k], pool.s[k]));
            GSLM_CUDA_OK(cudaStreamWaitEvent(s, pool.done[k], 0));
        }
        return 0;
    };
    // error exit after a fork: the caller is about to release the buffers the lanes may still be working on, and
    // nothing orders its stream behind them any more -- drain the lanes on the host before reporting the failure
    auto fail_lanes = [&]() -> int {
        if (lanes)
            for (int k = 0; k < K; k++) cudaStreamSynchronize(pool.s[k]);
        return 1;
    };
    // one geometry allocation: n_views per-view states, the radii of views the caller does not want back, the totals
    const size_t gstride = align_up(gslm_geom_bytes(P)), rstride = align_up((size_t)P * sizeof(int32_t));
    size_t ibytes = 0;
    for (int v = 0; v < n_views; v++) {
        const size_t b = gslm_image_bytes(cams[v].width, cams[v].height);
        ibytes = b > ibytes ? b : ibytes;
    }
    char* gbase = (char*)alloc(alloc_ctx, GSLM_BUF_GEOM,
                               128 + (size_t)n_views * (gstride + rstride) + align_up(n_views * sizeof(uint32_t)));
    if (!gbase) { set_error("allocator returned null"); return 1; }
    void* ibuf[LanePool::MAX] = {};
    for (int k = 0; k < K; k++) {      // per lane, reused view after view in stream order
        ibuf[k] = alloc(alloc_ctx, GSLM_BUF_IMAGE | (k << 8), ibytes);
        if (!ibuf[k]) { set_error("allocator returned null"); return 1; }
    }
    gbase = (char*)align_up((size_t)gbase);
    char* rbase = gbase + (size_t)n_views * gstride;
    uint32_t* totals = reinterpret_cast<uint32_t*>(rbase + (size_t)n_views * rstride);
    auto radii_of = [&](int v) { return (radii && radii[v]) ? radii[v] : reinterpret_cast<int32_t*>(rbase + v * rstride); };
    std::vector<uint32_t> R(n_views, 0u);
    if (P > 0) {
        if (fork_lanes()) return fail_lanes();
        for (int v = 0; v < n_views; v++) {
            cudaStream_t ls = lane_stream(v);
            GeomView gv = carve_geom(gbase + v * gstride, P);
            if (launch_preprocess(*g, cams[v], gv, radii_of(v), ls)) return fail_lanes();
            if (cams[v].tile_culling && run_tile_cull(gv, radii_of(v), P, cams[v].width, cams[v].height, ls)) return fail_lanes();
            if (run_scan(gv, P, ls)) return fail_lanes();
        }
        if (join_lanes()) return fail_lanes();
        GeomView g0 = carve_geom(gbase, P);
        const size_t off = (size_t)((const char*)(g0.point_offsets + (P - 1)) - gbase);
        k_gather_totals<<<(n_views + 127) / 128, 128, 0, s>>>(gbase, gstride, off, n_views, totals);
        GSLM_CUDA_OK(cudaGetLastError());
        // the ONE host synchronisation of the batch (rasterizer_impl.cu:325 has one per view)
        GSLM_CUDA_OK(cudaMemcpyAsync(R.data(), totals, n_views * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        GSLM_CUDA_OK(cudaStreamSynchronize(s));
    }
    uint32_t Rmax = 0;
    for (int v = 0; v < n_views; v++) { num_rendered[v] = (int32_t)R[v]; Rmax = R[v] > Rmax ? R[v] : Rmax; }
    void* tbuf[LanePool::MAX] = {};
    for (int k = 0; k < K; k++) {
        tbuf[k] = alloc(alloc_ctx, GSLM_BUF_BINNING_TEMP | (k << 8), gslm_binning_temp_bytes(P, Rmax));
        if (!tbuf[k]) { set_error("allocator returned null"); return fail_lanes(); }
    }
    if (fork_lanes()) return fail_lanes();
    for (int v = 0; v < n_views; v++) {
        const GslmCamera& cam = cams[v];
        cudaStream_t ls = lane_stream(v);
        GeomView gv = carve_geom(gbase + v * gstride, P);
        ImageView iv = carve_image(ibuf[v % K], cam.width, cam.height);
        BinningTemp bt = carve_binning_temp(tbuf[v % K], R[v], P);
        if (run_binning(gv, radii_of(v), bt, iv, P, cam.width, cam.height, R[v], cam.tile_culling != 0, ls)) return fail_lanes();
        const float* features = g->colors_precomp ? g->colors_precomp : gv.rgb;
        if (launch_blend_fwd(gv, bt, iv, cam.width, cam.height, features, cam.bg, out_color[v],
                             out_invdepth ? out_invdepth[v] : nullptr, false, ls)) return fail_lanes();
        if (debug_sync(&cam, ls, "render_batch")) return fail_lanes();
    }
    return join_lanes() ? fail_lanes() : 0;
}

int32_t gslm_count_instances(const GslmGaussians* g, int32_t n_views, const GslmCamera* cams, GslmAllocFn alloc,
                             void* alloc_ctx, int32_t* num_rendered, void* stream) {
    if (n_views < 0 || (n_views > 0 && (!cams || !num_rendered)) || !alloc || !g) { set_error("null argument"); return 1; }
    for (int v = 0; v < n_views; v++) {
        if (check_inputs(g, &cams[v])) return 1;
        num_rendered[v] = 0;
    }
    const int P = g->P;
    if (n_views == 0 || P == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    // one geometry state + radii, reused view after view (stream order); the counts collect in a device array
    const size_t gbytes = align_up(gslm_geom_bytes(P)), rbytes = align_up((size_t)P * sizeof(int32_t));
    char* base = (char*)alloc(alloc_ctx, -1, 128 + gbytes + rbytes + align_up((size_t)n_views * sizeof(uint32_t)));
    if (!base) { set_error("allocator returned null"); return 1; }
    base = (char*)align_up((size_t)base);
    int32_t* radii = reinterpret_cast<int32_t*>(base + gbytes);
    uint32_t* totals = reinterpret_cast<uint32_t*>(base + gbytes + rbytes);
    GeomView gv = carve_geom(base, P);
    for (int v = 0; v < n_views; v++) {
        if (launch_preprocess(*g, cams[v], gv, radii, s)) return 1;
        if (cams[v].tile_culling && run_tile_cull(gv, radii, P, cams[v].width, cams[v].height, s)) return 1;
        if (run_scan(gv, P, s)) return 1;
        GSLM_CUDA_OK(cudaMemcpyAsync(totals + v, gv.point_offsets + (P - 1), sizeof(uint32_t), cudaMemcpyDeviceToDevice, s));
    }
    std::vector<uint32_t> R(n_views, 0u);
    GSLM_CUDA_OK(cudaMemcpyAsync(R.data(), totals, n_views * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    GSLM_CUDA_OK(cudaStreamSynchronize(s));
    for (int v = 0; v < n_views; v++) num_rendered[v] = (int32_t)R[v];
    return 0;
}

int32_t gslm_forward_batch(const GslmGaussians* g, int32_t n_views, const GslmCamera* cams, GslmAllocFn alloc,
                           void* alloc_ctx, float* const* out_color, float* const* out_invdepth,
                           int32_t* const* radii, const GslmResidualEpilogue* epilogues, int32_t* num_rendered,
                           int32_t* num_cached, void* stream) {
    if (n_views < 0 || (n_views > 0 && (!cams || !out_color || !radii || !num_rendered || !num_cached)) || !alloc || !g) {
        set_error("null argument");
        return 1;
    }
    if (n_views == 0) return 0;
    for (int v = 0; v < n_views; v++) {
        if (check_inputs(g, &cams[v])) return 1;
        if (!out_color[v] || (g->P > 0 && !radii[v])) { set_error("null argument"); return 1; }
        num_rendered[v] = num_cached[v] = 0;
    }
    cudaStream_t s = (cudaStream_t)stream;
    const int P = g->P;
    std::vector<GeomView> gv(n_views);
    std::vector<ImageView> iv(n_views);
    for (int v = 0; v < n_views; v++) {
        void* gb = alloc(alloc_ctx, GSLM_BUF_GEOM | (v << 8), gslm_geom_bytes(P));
        void* ib = alloc(alloc_ctx, GSLM_BUF_IMAGE | (v << 8), gslm_image_bytes(cams[v].width, cams[v].height));
        if (!gb || !ib) { set_error("allocator returned null"); return 1; }
        gv[v] = carve_geom(gb, P);
        iv[v] = carve_image(ib, cams[v].width, cams[v].height);
    }
    uint32_t* host = pinned_counts(2 * (size_t)n_views);
    if (!host) { set_error("pinned host allocation failed"); return 1; }
    uint32_t* R = host;
    uint32_t* Rc = host + n_views;
    for (int v = 0; v < 2 * n_views; v++) host[v] = 0u;
    // lanes: views are dealt round robin to K >= 2 internal streams (one sort scratch each); see gslm_render_batch
    int K = render_lanes(n_views);
    if (K < 2 && !stage_timing_enabled()) K = n_views > 1 ? 2 : 1;
    LanePool* pool_p = lane_pool();
    if (!pool_p) { set_error("no current CUDA device for the render lanes"); return 1; }
    LanePool& pool = *pool_p;
    if (K > 1) K = pool.get(K);
    if (K < 1) K = 1;
    const bool lanes = K > 1;
    auto lane_stream = [&](int v) { return lanes ? pool.s[v % K] : s; };
    auto fork_lanes = [&]() -> int {
        if (!lanes) return 0;
        GSLM_CUDA_OK(cudaEventRecord(pool.fork, s));
        for (int k = 0; k < K; k++) GSLM_CUDA_OK(cudaStreamWaitEvent(pool.s[k], pool.fork, 0));
        return 0;
    };
    auto join_lanes = [&]() -> int {
        if (!lanes) return 0;
        for (int k = 0; k < K; k++) {
            GSLM_CUDA_OK(cudaEventRecord(pool.done[k], pool.s[k]));
            GSLM_CUDA_OK(cudaStreamWaitEvent(s, pool.done[k], 0));
        }
        return 0;
    };
    // error exit after a fork: the caller is about to release the buffers the lanes may still be working on, and
    // nothing orders its stream behind them any more -- drain the lanes on the host before reporting the failure
    auto fail_lanes = [&]() -> int {
        if (lanes)
            for (int k = 0; k < K; k++) cudaStreamSynchronize(pool.s[k]);
        return 1;
    };
    // phase 1: every view preprocessed, culled and scanned; ONE synchronisation for all instance counts
    if (P > 0) {
        if (fork_lanes()) return fail_lanes();
        for (int v = 0; v < n_views; v++) {
            cudaStream_t ls = lane_stream(v);
            if (launch_preprocess(*g, cams[v], gv[v], radii[v], ls)) return fail_lanes();
            if (cams[v].tile_culling && run_tile_cull(gv[v], radii[v], P, cams[v].width, cams[v].height, ls)) return fail_lanes();
            if (run_scan(gv[v], P, ls)) return fail_lanes();
            if (cudaMemcpyAsync(R + v, gv[v].point_offsets + (P - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, ls) != cudaSuccess) {
                set_error("forward_batch: count read-back failed");
                return fail_lanes();
            }
        }
        if (join_lanes()) return fail_lanes();
        if (cudaStreamSynchronize(s) != cudaSuccess) { set_error("forward_batch: synchronisation failed"); return fail_lanes(); }
    }
    uint32_t Rmax = 0;
    for (int v = 0; v < n_views; v++) { num_rendered[v] = (int32_t)R[v]; Rmax = R[v] > Rmax ? R[v] : Rmax; }
    // phase 2, in rounds of K views: bin + blend + tile scan of the round's views are queued on their lanes, THEN the
    // host waits for each view's compact count (sizes the persistent cache) and queues its compaction -- while it
    // waits the device has the other lanes' work, and by the time it returns to a lane that lane's scratch is free
    // again in stream order (compaction(v) precedes binning(v + K)).  With one lane (stage timers on) the wait is exposed.
    void* tbuf[LanePool::MAX] = {};
    for (int k = 0; k < K; k++) {
        tbuf[k] = alloc(alloc_ctx, GSLM_BUF_BINNING_TEMP | (k << 8), gslm_binning_temp_bytes(P, Rmax));
        if (!tbuf[k]) { set_error("allocator returned null"); return 1; }
    }
    cudaEvent_t ev[LanePool::MAX];
    for (int k = 0; k < K; k++) GSLM_CUDA_OK(cudaEventCreateWithFlags(&ev[k], cudaEventDisableTiming));
    int rc = fork_lanes();
    for (int v0 = 0; v0 < n_views && !rc; v0 += K) {
        const int n = (n_views - v0) < K ? (n_views - v0) : K;
        for (int i = 0; i < n && !rc; i++) {
            const int v = v0 + i;
            const GslmCamera& cam = cams[v];
            cudaStream_t ls = lane_stream(v);
            const int W = cam.width, H = cam.height;
            const int T = ((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
            BinningTemp bt = carve_binning_temp(tbuf[v % K], R[v], P);
            const float* features = g->colors_precomp ? g->colors_precomp : gv[v].rgb;
            rc = run_binning(gv[v], radii[v], bt, iv[v], P, W, H, R[v], cam.tile_culling != 0, ls)
                 || launch_blend_fwd(gv[v], bt, iv[v], W, H, features, cam.bg, out_color[v],
                                     out_invdepth ? out_invdepth[v] : nullptr, true, ls)
                 || run_tile_scan(iv[v], W, H, ls);
            if (rc) break;
            if (cudaMemcpyAsync(Rc + v, iv[v].cstart + T, sizeof(uint32_t), cudaMemcpyDeviceToHost, ls) != cudaSuccess ||
                cudaEventRecord(ev[v % K], ls) != cudaSuccess) { set_error("forward_batch: read-back failed"); rc = 1; break; }
            if (epilogues && launch_residual_epilogue(out_color[v], W, H, epilogues[v], ls)) { rc = 1; break; }
        }
        for (int i = 0; i < n && !rc; i++) {
            const int u = v0 + i;
            const GslmCamera& cam = cams[u];
            cudaStream_t ls = lane_stream(u);
            if (cudaEventSynchronize(ev[u % K]) != cudaSuccess) { set_error("forward_batch: event wait failed"); rc = 1; break; }
            num_cached[u] = (int32_t)Rc[u];
            void* bbuf = alloc(alloc_ctx, GSLM_BUF_BINNING | (u << 8), gslm_binning_bytes(Rc[u]));
            if (!bbuf) { set_error("allocator returned null"); rc = 1; break; }
            BinningView bv = carve_binning(bbuf, Rc[u]);
            BinningTemp bt = carve_binning_temp(tbuf[u % K], R[u], P);
            if (P > 0 && run_compaction(gv[u], bv, bt, iv[u], P, cam.width, cam.height, Rc[u], ls)) { rc = 1; break; }
            if (cam.debug) rc = debug_sync(&cam, ls, "forward_batch");
        }
    }
    if (join_lanes()) rc = 1;
    if (rc) fail_lanes();
    for (int k = 0; k < K; k++) cudaEventDestroy(ev[k]);
    return rc;
}

int32_t gslm_jvp_cached(const GslmGaussians* g, const GslmGaussians* gt, const GslmCamera* cam, const int32_t* radii,
                        const void* geom_buffer, const void* binning_buffer, const void* image_buffer,
                        int32_t num_cached, void* scratch, float* out_color_tangent, float* out_invdepth_tangent,
                        void* stream) {
    if (check_inputs(g, cam)) return 1;
    if (!out_color_tangent || !image_buffer || !binning_buffer || !geom_buffer || !scratch) {
        set_error("null argument");
        return 1;
    }
    cudaStream_t s = (cudaStream_t)stream;
    GeomView gv = carve_geom(const_cast<void*>(geom_buffer), g->P);
    ImageView iv = carve_image(const_cast<void*>(image_buffer), cam->width, cam->height);
    BinningView bv = carve_binning(const_cast<void*>(binning_buffer), num_cached);
    float4* tgeom = reinterpret_cast<float4*>(align_up((size_t)scratch));
    TangentPtrs t = make_tangents(gt, nullptr);
    if (launch_preprocess_tangent(*g, t, *cam, gv, tgeom, s)) return 1;
    if (launch_blend_jvp(bv, iv, cam->width, cam->height, tgeom, cam->bg, nullptr, out_color_tangent,
                         out_invdepth_tangent, s)) return 1;
    return debug_sync(cam, s, "jvp_cached");
}

int32_t gslm_forward_jvp(const GslmGaussians* g, const GslmGaussians* gt, const GslmCamera* cam,
                         const GslmCameraTangent* camt, GslmAllocFn alloc, void* alloc_ctx, float* out_color,
                         float* out_invdepth, int32_t* radii, float* out_color_tangent, float* out_invdepth_tangent,
                         int32_t* num_rendered, int32_t* num_cached, void* stream) {
    // rasterizer_jvp_impl.h:125-325 runs primal and tangent in one pass and then shrinks its
    // buffers to the primal layout; here the primal pass IS the plain forward, followed by the
    // tangent stages on the compact state it left behind.
    struct Capture { GslmAllocFn fn; void* ctx; void* p[4]; } cap = {alloc, alloc_ctx, {nullptr, nullptr, nullptr, nullptr}};
    auto hook = [](void* c, int32_t which, size_t bytes) -> void* {
        Capture* k = (Capture*)c;
        void* r = k->fn(k->ctx, which, bytes);
        if (which >= 0 && which < 4) k->p[which] = r;
        return r;
    };
    if (!out_color_tangent) { set_error("null argument"); return 1; }
    if (gslm_forward(g, cam, hook, &cap, out_color, out_invdepth, radii, num_rendered, num_cached, stream)) return 1;
    cudaStream_t s = (cudaStream_t)stream;
    const int P = g->P, W = cam->width, H = cam->height;
    GeomView gv = carve_geom(cap.p[GSLM_BUF_GEOM], P);
    ImageView iv = carve_image(cap.p[GSLM_BUF_IMAGE], W, H);
    BinningView bv = carve_binning(cap.p[GSLM_BUF_BINNING], *num_cached);
    // the tangent geometry (48 B per Gaussian) borrows the sort scratch, which is dead once the
    // forward pass has returned, when it fits; otherwise one more scratch allocation (which = -1)
    float4* tgeom = nullptr;
    const size_t need = gslm_jvp_scratch_bytes(P);
    if (cap.p[GSLM_BUF_BINNING_TEMP] && gslm_binning_temp_bytes(P, *num_rendered) >= need)
        tgeom = reinterpret_cast<float4*>(cap.p[GSLM_BUF_BINNING_TEMP]);
    else {
        void* extra = alloc(alloc_ctx, -1, need);
        if (!extra) { set_error("allocator returned null"); return 1; }
        tgeom = reinterpret_cast<float4*>(align_up((size_t)extra));
    }
    TangentPtrs t = make_tangents(gt, camt);
    if (launch_preprocess_tangent(*g, t, *cam, gv, tgeom, s)) return 1;
    if (launch_blend_jvp(bv, iv, W, H, tgeom, cam->bg, camt ? camt->bg : nullptr, out_color_tangent,
                         out_invdepth_tangent, s)) return 1;
    return debug_sync(cam, s, "forward_jvp");
}

int32_t gslm_backward(const GslmGaussians* g, const GslmCamera* cam, const int32_t* radii, const void* geom_buffer,
                      const void* binning_buffer, const void* image_buffer, int32_t num_cached,
                      const float* dL_dcolor, const float* dL_dinvdepth, void* scratch, GslmGrads* grads,
                      void* stream) {
    if (check_inputs(g, cam)) return 1;
    if (!grads || !dL_dcolor || !geom_buffer || !binning_buffer || !image_buffer || !scratch) {
        set_error("null argument");
        return 1;
    }
    cudaStream_t s = (cudaStream_t)stream;
    const int P = g->P, W = cam->width, H = cam->height;
    if (P == 0) return 0;
    GeomView gv = carve_geom(const_cast<void*>(geom_buffer), P);
    ImageView iv = carve_image(const_cast<void*>(image_buffer), W, H);
    BinningView bv = carve_binning(const_cast<void*>(binning_buffer), num_cached);
    float4* inst = reinterpret_cast<float4*>(align_up((size_t)scratch));
    if (num_cached > 0)
        if (launch_blend_bwd(bv, iv, W, H, cam->bg, dL_dcolor, dL_dinvdepth, inst, s)) return 1;
    if (debug_sync(cam, s, "render bwd")) return 1;
    if (launch_preprocess_bwd(*g, *cam, radii, gv, bv, inst, dL_dinvdepth != nullptr, *grads, s)) return 1;
    return debug_sync(cam, s, "preprocess bwd");
}

int32_t gslm_jtjv(const GslmGaussians* g, const GslmGaussians* gt, const GslmCamera* cam, const int32_t* radii,
                  const void* geom_buffer, const void* binning_buffer, const void* image_buffer,
                  int32_t num_cached, const float* weight, void* scratch, GslmGrads* grads, float* out_vJJv,
                  void* stream) {
    if (check_inputs(g, cam)) return 1;
    if (!grads || !geom_buffer || !binning_buffer || !image_buffer || !scratch) { set_error("null argument"); return 1; }
    cudaStream_t s = (cudaStream_t)stream;
    const int P = g->P, W = cam->width, H = cam->height;
    if (P == 0) return 0;
    GeomView gv = carve_geom(const_cast<void*>(geom_buffer), P);
    ImageView iv = carve_image(const_cast<void*>(image_buffer), W, H);
    BinningView bv = carve_binning(const_cast<void*>(binning_buffer), num_cached);
    float4* tgeom = reinterpret_cast<float4*>(align_up((size_t)scratch));
    float4* inst = reinterpret_cast<float4*>((char*)tgeom + gslm_jvp_scratch_bytes(P) - 128);
    inst = reinterpret_cast<float4*>(align_up((size_t)inst));
    TangentPtrs t = make_tangents(gt, nullptr);
    if (launch_preprocess_tangent(*g, t, *cam, gv, tgeom, s)) return 1;
    if (num_cached > 0)
        if (launch_blend_jtjv(bv, iv, W, H, cam->bg, tgeom, weight, inst, out_vJJv, s)) return 1;
    if (launch_preprocess_bwd(*g, *cam, radii, gv, bv, inst, false, *grads, s)) return 1;
    return debug_sync(cam, s, "jtjv");
}

size_t gslm_jtjv_batch_scratch_bytes(int32_t P, int32_t n_views, const int32_t* num_cached) {
    size_t worst = 0;
    for (int v0 = 0; v0 < n_views; v0 += MAX_BATCH_VIEWS) {
        size_t need = 0;
        for (int v = v0; v < n_views && v < v0 + MAX_BATCH_VIEWS; v++)
            need += gslm_jtjv_scratch_bytes(P, num_cached ? num_cached[v] : 0);
        worst = need > worst ? need : worst;
    }
    return worst;
}

int32_t gslm_jtjv_batch(const GslmGaussians* g, const GslmGaussians* gt, int32_t n_views, const GslmCamera* cams,
                        const int32_t* const* radii, const void* const* geom_buffers,
                        const void* const* binning_buffers, const void* const* image_buffers,
                        const int32_t* num_cached, const float* const* weights, void* scratch, size_t scratch_bytes,
                        GslmGrads* grads, float* out_vJJv, void* stream) {
    return gslm_jtjv_batch_sliced(g, gt, n_views, cams, radii, geom_buffers, binning_buffers, image_buffers, num_cached,
                                  weights, scratch, scratch_bytes, grads, out_vJJv, 1, nullptr, nullptr, stream);
}

int32_t gslm_jtjv_batch_sliced(const GslmGaussians* g, const GslmGaussians* gt, int32_t n_views, const GslmCamera* cams,
                               const int32_t* const* radii, const void* const* geom_buffers,
                               const void* const* binning_buffers, const void* const* image_buffers,
                               const int32_t* num_cached, const float* const* weights, void* scratch,
                               size_t scratch_bytes, GslmGrads* grads, float* out_vJJv, int32_t n_slices,
                               GslmSliceFn on_slice, void* slice_ctx, void* stream) {
    if (n_views < 0 || (n_views > 0 && (!cams || !radii || !geom_buffers || !binning_buffers || !image_buffers ||
                                        !num_cached || !scratch)) || !grads || !g) {
        set_error("null argument");
        return 1;
    }
    if (scratch_bytes < gslm_jtjv_batch_scratch_bytes(g->P, n_views, num_cached)) {
        set_error("scratch too small: need %zu bytes", gslm_jtjv_batch_scratch_bytes(g->P, n_views, num_cached));
        return 1;
    }
    for (int v = 0; v < n_views; v++)
        if (check_inputs(g, &cams[v])) return 1;
    cudaStream_t s = (cudaStream_t)stream;
    const int P = g->P;
    if (P == 0) return 0;
    TangentPtrs t = make_tangents(gt, nullptr);
    GslmGrads acc = *grads;
    // overwrite semantics = clear, then accumulate every view.  (cudaMemsetAsync beat zero-filling inside the first
    // adjoint launch by 0.06 ms per product at P = 1M -- the stores delay the start of every warp's dependent
    // chain -- so that variant was removed.)
    if (!acc.accumulate && clear_sinks(acc, P, g->M, s)) return 1;
    acc.accumulate = 1;
    if (n_views == 0) {
        empty_product_callbacks(P, n_slices, on_slice, slice_ctx);
        return 0;
    }
    for (int v0 = 0; v0 < n_views; v0 += MAX_BATCH_VIEWS) {
        const int n = (n_views - v0) < MAX_BATCH_VIEWS ? (n_views - v0) : MAX_BATCH_VIEWS;
        GeomView gv[MAX_BATCH_VIEWS];
        ImageView iv[MAX_BATCH_VIEWS];
        BinningView bv[MAX_BATCH_VIEWS];
        float4* tgeom[MAX_BATCH_VIEWS];
        float4* inst[MAX_BATCH_VIEWS];
        char* p = reinterpret_cast<char*>(align_up((size_t)scratch));
        for (int i = 0; i < n; i++) {
            const int v = v0 + i;
            gv[i] = carve_geom(const_cast<void*>(geom_buffers[v]), P);
            iv[i] = carve_image(const_cast<void*>(image_buffers[v]), cams[v].width, cams[v].height);
            bv[i] = carve_binning(const_cast<void*>(binning_buffers[v]), num_cached[v]);
            tgeom[i] = reinterpret_cast<float4*>(p);
            p += align_up((size_t)P * REC_F4 * sizeof(float4));
            inst[i] = reinterpret_cast<float4*>(p);
            p += align_up((size_t)num_cached[v] * REC_F4 * sizeof(float4));
        }
        if (launch_preprocess_tangent_batch(*g, t, n, cams + v0, gv, tgeom, s)) return 1;
        if (launch_blend_jtjv_batch(n, bv, iv, cams + v0, tgeom, weights ? weights + v0 : nullptr, inst, out_vJJv, s))
            return 1;
        if (adjoint_chunk(*g, n, cams + v0, radii + v0, gv, inst, acc,
                          v0 + MAX_BATCH_VIEWS >= n_views, n_slices, on_slice, slice_ctx, s)) return 1;
    }
    return 0;
}

// ---- batched linear maps on cached linearisations (matvec / matvec_T of the solver) ----
size_t gslm_jvp_batch_scratch_bytes(int32_t P, int32_t n_views) {
    if (n_views <= 0) return 0;     // a rank that owns no view (multi-GPU) brings no scratch
    const int n = n_views < MAX_BATCH_VIEWS ? n_views : MAX_BATCH_VIEWS;
    return (size_t)n * align_up((size_t)P * REC_F4 * sizeof(float4)) + 128;
}

int32_t gslm_jvp_batch(const GslmGaussians* g, const GslmGaussians* gt, int32_t n_views, const GslmCamera* cams,
                       const void* const* geom_buffers, const void* const* binning_buffers,
                       const void* const* image_buffers, const int32_t* num_cached, void* scratch, size_t scratch_bytes,
                       float* const* out_color_tangent, void* stream) {
    if (n_views < 0 || (n_views > 0 && (!cams || !geom_buffers || !binning_buffers || !image_buffers || !num_cached ||
                                        !scratch || !out_color_tangent)) || !g) {
        set_error("null argument");
        return 1;
    }
    if (scratch_bytes < gslm_jvp_batch_scratch_bytes(g->P, n_views)) {
        set_error("scratch too small: need %zu bytes", gslm_jvp_batch_scratch_bytes(g->P, n_views));
        return 1;
    }
    for (int v = 0; v < n_views; v++)
        if (check_inputs(g, &cams[v])) return 1;
    cudaStream_t s = (cudaStream_t)stream;
    const int P = g->P;
    TangentPtrs t = make_tangents(gt, nullptr);
    for (int v0 = 0; v0 < n_views; v0 += MAX_BATCH_VIEWS) {
        const int n = (n_views - v0) < MAX_BATCH_VIEWS ? (n_views - v0) : MAX_BATCH_VIEWS;
        GeomView gv[MAX_BATCH_VIEWS];
        ImageView iv[MAX_BATCH_VIEWS];
        BinningView bv[MAX_BATCH_VIEWS];
        float4* tgeom[MAX_BATCH_VIEWS];
        char* p = reinterpret_cast<char*>(align_up((size_t)scratch));
        for (int i = 0; i < n; i++) {
            const int v = v0 + i;
            gv[i] = carve_geom(const_cast<void*>(geom_buffers[v]), P);
            iv[i] = carve_image(const_cast<void*>(image_buffers[v]), cams[v].width, cams[v].height);
            bv[i] = carve_binning(const_cast<void*>(binning_buffers[v]), num_cached[v]);
            tgeom[i] = reinterpret_cast<float4*>(p);
            p += align_up((size_t)P * REC_F4 * sizeof(float4));
        }
        if (P > 0 && launch_preprocess_tangent_batch(*g, t, n, cams + v0, gv, tgeom, s)) return 1;
        if (launch_blend_jvp_batch(n, bv, iv, cams + v0, tgeom, out_color_tangent + v0, s)) return 1;
    }
    return 0;
}

size_t gslm_backward_batch_scratch_bytes(int32_t n_views, const int32_t* num_cached) {
    size_t worst = 0;
    for (int v0 = 0; v0 < n_views; v0 += MAX_BATCH_VIEWS) {
        size_t need = 0;
        for (int v = v0; v < n_views && v < v0 + MAX_BATCH_VIEWS; v++)
            need += gslm_backward_scratch_bytes(num_cached ? num_cached[v] : 0);
        worst = need > worst ? need : worst;
    }
    return worst;
}

int32_t gslm_backward_batch(const GslmGaussians* g, int32_t n_views, const GslmCamera* cams,
                            const int32_t* const* radii, const void* const* geom_buffers,
                            const void* const* binning_buffers, const void* const* image_buffers,
                            const int32_t* num_cached, const float* const* dL_dcolor, void* scratch,
                            size_t scratch_bytes, GslmGrads* grads, void* stream) {
    return gslm_backward_batch_sliced(g, n_views, cams, radii, geom_buffers, binning_buffers, image_buffers, num_cached,
                                      dL_dcolor, scratch, scratch_bytes, grads, 1, nullptr, nullptr, stream);
}

int32_t gslm_backward_batch_sliced(const GslmGaussians* g, int32_t n_views, const GslmCamera* cams,
                                   const int32_t* const* radii, const void* const* geom_buffers,
                                   const void* const* binning_buffers, const void* const* image_buffers,
                                   const int32_t* num_cached, const float* const* dL_dcolor, void* scratch,
                                   size_t scratch_bytes, GslmGrads* grads, int32_t n_slices, GslmSliceFn on_slice,
                                   void* slice_ctx, void* stream) {
    if (n_views < 0 || (n_views > 0 && (!cams || !radii || !geom_buffers || !binning_buffers || !image_buffers ||
                                        !num_cached || !scratch || !dL_dcolor)) || !grads || !g) {
        set_error("null argument");
        return 1;
    }
    if (scratch_bytes < gslm_backward_batch_scratch_bytes(n_views, num_cached)) {
        set_error("scratch too small: need %zu bytes", gslm_backward_batch_scratch_bytes(n_views, num_cached));
        return 1;
    }
    for (int v = 0; v < n_views; v++)
        if (check_inputs(g, &cams[v])) return 1;
    cudaStream_t s = (cudaStream_t)stream;
    const int P = g->P;
    if (P == 0) return 0;
    GslmGrads acc = *grads;
    if (!acc.accumulate && clear_sinks(acc, P, g->M, s)) return 1;
    acc.accumulate = 1;
    if (n_views == 0) {
        empty_product_callbacks(P, n_slices, on_slice, slice_ctx);
        return 0;
    }
    for (int v0 = 0; v0 < n_views; v0 += MAX_BATCH_VIEWS) {
        const int n = (n_views - v0) < MAX_BATCH_VIEWS ? (n_views - v0) : MAX_BATCH_VIEWS;
        GeomView gv[MAX_BATCH_VIEWS];
        ImageView iv[MAX_BATCH_VIEWS];
        BinningView bv[MAX_BATCH_VIEWS];
        float4* inst[MAX_BATCH_VIEWS];
        char* p = reinterpret_cast<char*>(align_up((size_t)scratch));
        for (int i = 0; i < n; i++) {
            const int v = v0 + i;
            gv[i] = carve_geom(const_cast<void*>(geom_buffers[v]), P);
            iv[i] = carve_image(const_cast<void*>(image_buffers[v]), cams[v].width, cams[v].height);
            bv[i] = carve_binning(const_cast<void*>(binning_buffers[v]), num_cached[v]);
            inst[i] = reinterpret_cast<float4*>(p);
            p += align_up((size_t)num_cached[v] * REC_F4 * sizeof(float4));
        }
        if (launch_blend_bwd_batch(n, bv, iv, cams + v0, dL_dcolor + v0, inst, s)) return 1;
        if (adjoint_chunk(*g, n, cams + v0, radii + v0, gv, inst, acc,
                          v0 + MAX_BATCH_VIEWS >= n_views, n_slices, on_slice, slice_ctx, s)) return 1;
    }
    return 0;
}

int32_t gslm_tangent_records(const GslmGaussians* g, const GslmGaussians* gt, const GslmCamera* cam,
                             const GslmCameraTangent* camt, const void* geom_buffer, float* out_records, void* stream) {
    if (check_inputs(g, cam)) return 1;
    if (!geom_buffer || !out_records) { set_error("null argument"); return 1; }
    if (g->P == 0) return 0;
    GeomView gv = carve_geom(const_cast<void*>(geom_buffer), g->P);
    TangentPtrs t = make_tangents(gt, camt);
    if (launch_preprocess_tangent(*g, t, *cam, gv, reinterpret_cast<float4*>(out_records), (cudaStream_t)stream)) return 1;
    return debug_sync(cam, (cudaStream_t)stream, "tangent_records");
}

int32_t gslm_buffer_view(int32_t P, int32_t width, int32_t height, int64_t num_rendered, int64_t num_cached,
                         const void* geom_buffer, const void* binning_buffer, const void* image_buffer,
                         const void* binning_temp_buffer, GslmBufferView* out) {
    if (!out) { set_error("null argument"); return 1; }
    memset(out, 0, sizeof(*out));
    if (geom_buffer) {
        GeomView gv = carve_geom(const_cast<void*>(geom_buffer), P);
        out->tiles_touched = gv.tiles_touched; out->point_offsets = gv.point_offsets; out->depths = gv.depths;
        out->means2D = reinterpret_cast<const float*>(gv.means2D);
        out->conic_opacity = reinterpret_cast<const float*>(gv.conic_opacity);
        out->rgb = gv.rgb; out->cov3D = gv.cov3D; out->clamped = gv.clamped; out->g_off = gv.g_off;
    }
    if (binning_buffer) {
        BinningView bv = carve_binning(const_cast<void*>(binning_buffer), num_cached);
        out->records = reinterpret_cast<const float*>(bv.records); out->cid = bv.cid; out->g_list = bv.g_list;
        out->masks = reinterpret_cast<const uint32_t*>(bv.masks);
    }
    if (binning_temp_buffer) {
        BinningTemp bt = carve_binning_temp(const_cast<void*>(binning_temp_buffer), num_rendered, P);
        out->tile_keys_sorted = bt.keys_sorted; out->point_list = bt.point_list;
    }
    if (image_buffer) {
        ImageView iv = carve_image(const_cast<void*>(image_buffer), width, height);
        out->ranges = reinterpret_cast<const uint32_t*>(iv.ranges); out->n_contrib = iv.n_contrib; out->final_T = iv.final_T;
        out->n_contrib_c = iv.n_contrib_c; out->cstart = iv.cstart;
    }
    return 0;
}

// ---- J v with built-in scratch: gslm_jvp_cached_noalloc ---------------------------------------
// Some hosts of this library have no way to hand us device scratch: a C / Go / Java embedder that links the
// rasterizer without torch has no device allocator of its own; a process whose framework pool has already claimed
// all of HBM cannot get more from cudaMalloc; an interactive viewer cannot afford the device-wide synchronisation
// cudaMalloc / cudaFree imply inside its frame loop.  For them the per-Gaussian tangent records (the only scratch
// J v needs: 48 B per Gaussian) live in the library image itself.  Every GPU that loads the library carries these
// GSLM_NOALLOC_MAX_GAUSSIANS * 48 B = 12 MiB until the process exits, used or not.
static __device__ __align__(256) float4 g_builtin_tangent_records[(size_t)GSLM_NOALLOC_MAX_GAUSSIANS * REC_F4];

// `g_builtin_tangent_records` in HOST code is only the handle the runtime knows the symbol by, not a device
// address: a kernel given it would fault (or scribble over host memory on systems where the GPU can reach it).
// The real address differs per device and comes from cudaGetSymbolAddress with that device current.
static const int MAX_DEVICES = 64;
static float4* g_builtin_records_addr[MAX_DEVICES];   // [device ordinal] -> device address, 0 = not looked up yet
static pthread_mutex_t g_builtin_records_lock = PTHREAD_MUTEX_INITIALIZER;

static int builtin_tangent_records(float4** out) {
    int dev = -1;
    GSLM_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEVICES) { set_error("device ordinal %d outside 0..%d", dev, MAX_DEVICES - 1); return 1; }
    pthread_mutex_lock(&g_builtin_records_lock);
    float4* p = g_builtin_records_addr[dev];
    if (!p) {
        void* sym = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&sym, g_builtin_tangent_records);
        if (e != cudaSuccess || !sym) {
            pthread_mutex_unlock(&g_builtin_records_lock);
            set_error("cudaGetSymbolAddress(built-in tangent records) on device %d: %s", dev, cudaGetErrorString(e));
            return 1;
        }
        p = g_builtin_records_addr[dev] = reinterpret_cast<float4*>(sym);
    }
    pthread_mutex_unlock(&g_builtin_records_lock);
    *out = p;
    return 0;
}

int32_t gslm_jvp_cached_noalloc(const GslmGaussians* g, const GslmGaussians* gt, const GslmCamera* cam,
                                const int32_t* radii, const void* geom_buffer, const void* binning_buffer,
                                const void* image_buffer, int32_t num_cached, float* out_color_tangent,
                                float* out_invdepth_tangent, void* stream) {
    (void)radii;
    if (check_inputs(g, cam)) return 1;
    if (!out_color_tangent || !image_buffer || !binning_buffer || !geom_buffer) { set_error("null argument"); return 1; }
    if (g->P > GSLM_NOALLOC_MAX_GAUSSIANS) {
        set_error("gslm_jvp_cached_noalloc: P = %d exceeds the built-in scratch (GSLM_NOALLOC_MAX_GAUSSIANS = %d); "
                  "use gslm_jvp_cached with caller-provided scratch", g->P, GSLM_NOALLOC_MAX_GAUSSIANS);
        return 1;
    }
    // first call on this device: looks the array's address up (nothing is allocated; the array is part of the
    // loaded module).  Later calls only read the table.
    float4* tgeom = nullptr;
    if (builtin_tangent_records(&tgeom)) return 1;
    cudaStream_t s = (cudaStream_t)stream;
    GeomView gv = carve_geom(const_cast<void*>(geom_buffer), g->P);
    ImageView iv = carve_image(const_cast<void*>(image_buffer), cam->width, cam->height);
    BinningView bv = carve_binning(const_cast<void*>(binning_buffer), num_cached);
    TangentPtrs t = make_tangents(gt, nullptr);
    // Nothing here has to start from zero: the tangent stage writes the record of every Gaussian that is active in
    // this view, and the tile sweep reads records of cached instances only, all of which are active.  Records left
    // by an earlier call (another scene, another view) stay in the other rows and are never read.
    if (launch_preprocess_tangent(*g, t, *cam, gv, tgeom, s)) return 1;
    if (launch_blend_jvp(bv, iv, cam->width, cam->height, tgeom, cam->bg, nullptr, out_color_tangent,
                         out_invdepth_tangent, s)) return 1;
    return debug_sync(cam, s, "jvp_cached_noalloc");
}

}  // extern "C"
