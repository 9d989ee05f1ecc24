// =============================================================================
// blend_kernels.cu -- tile kernels of the derivative paths:
//   k_blend_jvp   J v     image tangent of one linearised view          (a11)
//   k_blend_bwd   J^T u   per-instance gradient partials from dL/dimage (a13)
//   k_blend_jtjv  J^T W J v fused: tangent sweep, per-pixel weighting in registers, adjoint
//                 sweep -- the image-space vector J v never reaches HBM   
// The sweeps themselves live in blend_sweeps.cuh.  One WARP per 16x16 tile, one thread per
// eight pixels of a column; DWARPS independent warps (tiles) per CTA, each with its own slice
// of shared memory -- warps never synchronise with each other.
// =============================================================================
#include "blend_sweeps.cuh"
#include <cstring>

namespace gslm {

#ifndef GSLM_DWARPS
#define GSLM_DWARPS 1
#endif
constexpr int DWARPS = GSLM_DWARPS;                     // tiles (independent warps) per CTA
constexpr int DTHREADS = DWARPS * 32;
#ifndef GSLM_DMINB
#define GSLM_DMINB 18
#endif
constexpr int RING_F4 = 2 * WBATCH * REC_F4;            // float4 per warp: record ring (3 KB)
constexpr int RING_U4 = 2 * WBATCH * MSK_U4;            // uint4 per warp: mask ring (2 KB)

struct TileArgs {          // one linearised view
    const uint32_t* cstart;
    const float4* records;
    const uint4* masks;
    const uint32_t* cid;
    const uint32_t* g_inv;
    const float* final_T;
    const uint32_t* order;     // launch slot -> tile id (longest lists first)
    int W, H, tiles;
    const float* bg;
};

__device__ __forceinline__ bool make_tile(const TileArgs& a, int warp, WTile& c) {
    c.lane = threadIdx.x & 31;
    c.lbit = 1u << c.lane;
    const int slot = blockIdx.x * DWARPS + warp;
    if (slot >= a.tiles) return false;
    const int tile = (int)__ldg(a.order + slot);
    const int gx = (a.W + TILE_X - 1) / TILE_X;
    c.W = a.W; c.H = a.H;
    c.px = (tile % gx) * TILE_X + (c.lane & 15);
    c.pyb = (tile / gx) * TILE_Y + (c.lane >> 4);
    c.pxf = (float)c.px; c.pyf = (float)c.pyb;
    c.range = make_uint2(a.cstart[tile], a.cstart[tile + 1]);
    c.total = (int)(c.range.y - c.range.x);
    return true;
}

__device__ __forceinline__ void init_ring(WarpRing& ring, const WTile& c, const TileArgs& a, float4* s_rec, uint4* s_msk,
                                          uint64_t* s_bar) {
    ring.rec = s_rec; ring.msk = s_msk; ring.bar = s_bar;
    ring.src_rec = a.records + (size_t)c.range.x * REC_F4;
    ring.src_msk = a.masks + (size_t)c.range.x * MSK_U4;
    ring.total = c.total; ring.phase = 0;
    if (c.lane == 0) {
        mbar_init(&s_bar[0], 1);
        mbar_init(&s_bar[1], 1);
        mbar_fence_init();
    }
    __syncwarp();
}

template <bool WITH_D>
__device__ __forceinline__ void
blend_jvp_body(const TileArgs& a, const float4* __restrict__ tgeom, const float* __restrict__ bg_t,
               float* __restrict__ out_color_t, float* __restrict__ out_invdepth_t) {
    __shared__ __align__(128) float4 s_rec[DWARPS][RING_F4];
    __shared__ __align__(128) uint4 s_msk[DWARPS][RING_U4];
    __shared__ __align__(16) float4 s_tan[DWARPS][RING_F4];
    __shared__ __align__(8) uint64_t s_bar[DWARPS][2];
    const int warp = threadIdx.x >> 5;
    WTile c;
    if (!make_tile(a, warp, c)) return;
    WarpRing ring;
    init_ring(ring, c, a, s_rec[warp], s_msk[warp], s_bar[warp]);
    PairT st[4];
    tangent_sweep<WITH_D>(c, ring, s_tan[warp], a.cid, tgeom, st);
    const size_t HW = (size_t)a.H * a.W;
    const float b0 = a.bg[0], b1 = a.bg[1], b2 = a.bg[2];
    const float bt0 = bg_t ? bg_t[0] : 0.f, bt1 = bg_t ? bg_t[1] : 0.f, bt2 = bg_t ? bg_t[2] : 0.f;
#pragma unroll
    for (int k = 0; k < 8; k++) {   // forward_impl.h:531: C + T * bg on duals
        if (!c.inside(k)) continue;
        const PairT& s = st[k >> 1];
        const bool hi = k & 1;
        const float T = hi ? s.T.y : s.T.x, Tt = hi ? s.Tt.y : s.Tt.x;
        const size_t pix = c.pix(k);
        out_color_t[pix] = (hi ? s.C0.y : s.C0.x) + Tt * b0 + T * bt0;
        out_color_t[HW + pix] = (hi ? s.C1.y : s.C1.x) + Tt * b1 + T * bt1;
        out_color_t[2 * HW + pix] = (hi ? s.C2.y : s.C2.x) + Tt * b2 + T * bt2;
        if (WITH_D) out_invdepth_t[pix] = hi ? s.D.y : s.D.x;
    }
}

template <bool WITH_D>
__global__ void __launch_bounds__(DTHREADS, WITH_D ? 16 : GSLM_DMINB)
k_blend_jvp(const __grid_constant__ TileArgs a, const float4* __restrict__ tgeom, const float* __restrict__ bg_t,
            float* __restrict__ out_color_t, float* __restrict__ out_invdepth_t) {
    blend_jvp_body<WITH_D>(a, tgeom, bg_t, out_color_t, out_invdepth_t);
}

// Up to MAX_BATCH_VIEWS linearised views per launch (blockIdx.y), colour only: what LinearSolverFunctions.matvec /
// matvec_T issue per chunk of views (the reference's CGLS path, solver/conjugate_gradient.py:51-127, calls both
// every iteration), and the two tile passes of the L1 + SSIM normal product.
struct ViewBatch {
    TileArgs t[MAX_BATCH_VIEWS];
    const float4* tgeom[MAX_BATCH_VIEWS];      // JVP: per-Gaussian tangent records
    float* out_color_t[MAX_BATCH_VIEWS];       // JVP: image tangent [3,H,W]
    const float* dL_dcolor[MAX_BATCH_VIEWS];   // VJP: image cotangent [3,H,W]
    float4* inst_grads[MAX_BATCH_VIEWS];       // VJP: per-instance gradient records
};

__global__ void __launch_bounds__(DTHREADS, GSLM_DMINB) k_blend_jvp_batch(const __grid_constant__ ViewBatch vb) {
    blend_jvp_body<false>(vb.t[blockIdx.y], vb.tgeom[blockIdx.y], nullptr, vb.out_color_t[blockIdx.y], nullptr);
}

__device__ __forceinline__ void set_lane(f2& v, bool hi, float x) { if (hi) v.y = x; else v.x = x; }

template <bool WITH_D>
__device__ __forceinline__ void
blend_bwd_body(const TileArgs& a, const float* __restrict__ dL_dcolor, const float* __restrict__ dL_dinvdepth,
               float4* __restrict__ inst_grads) {
    __shared__ __align__(128) float4 s_rec[DWARPS][RING_F4];
    __shared__ __align__(128) uint4 s_msk[DWARPS][RING_U4];
    __shared__ float s_acc[DWARPS][(WBATCH + 1) * ACC_STRIDE];
    __shared__ __align__(8) uint64_t s_bar[DWARPS][2];
    const int warp = threadIdx.x >> 5;
    WTile c;
    if (!make_tile(a, warp, c)) return;
    WarpRing ring;
    init_ring(ring, c, a, s_rec[warp], s_msk[warp], s_bar[warp]);
    const size_t HW = (size_t)a.H * a.W;
    PairA sa[4];
#pragma unroll
    for (int k = 0; k < 8; k++) {
        PairA& s = sa[k >> 1];
        const bool hi = k & 1, in = c.inside(k);
        const size_t pix = in ? c.pix(k) : 0;
        set_lane(s.T, hi, in ? a.final_T[pix] : 0.f);
        set_lane(s.g0, hi, in ? dL_dcolor[pix] : 0.f);
        set_lane(s.g1, hi, in ? dL_dcolor[HW + pix] : 0.f);
        set_lane(s.g2, hi, in ? dL_dcolor[2 * HW + pix] : 0.f);
        set_lane(s.gd, hi, (in && WITH_D) ? dL_dinvdepth[pix] : 0.f);
    }
    const float bgv[3] = {a.bg[0], a.bg[1], a.bg[2]};
    adjoint_sweep<WITH_D>(c, ring, s_acc[warp], bgv, sa, a.g_inv, inst_grads);
}

template <bool WITH_D>
__global__ void __launch_bounds__(DTHREADS, WITH_D ? 16 : GSLM_DMINB)
k_blend_bwd(const __grid_constant__ TileArgs a, const float* __restrict__ dL_dcolor,
            const float* __restrict__ dL_dinvdepth, float4* __restrict__ inst_grads) {
    blend_bwd_body<WITH_D>(a, dL_dcolor, dL_dinvdepth, inst_grads);
}

__global__ void __launch_bounds__(DTHREADS, GSLM_DMINB) k_blend_bwd_batch(const __grid_constant__ ViewBatch vb) {
    blend_bwd_body<false>(vb.t[blockIdx.y], vb.dL_dcolor[blockIdx.y], nullptr, vb.inst_grads[blockIdx.y]);
}

// Up to MAX_BATCH_VIEWS linearised views per launch (blockIdx.y): the tail of one view's tiles overlaps the next
// view's, and a CG iteration needs n_views / MAX_BATCH_VIEWS launches of this kernel instead of n_views.
struct JtjvBatch {
    TileArgs t[MAX_BATCH_VIEWS];
    const float4* tgeom[MAX_BATCH_VIEWS];
    const float* weight[MAX_BATCH_VIEWS];
    float4* inst_grads[MAX_BATCH_VIEWS];
};

__global__ void __launch_bounds__(DTHREADS, GSLM_DMINB)
k_blend_jtjv(const __grid_constant__ JtjvBatch vb, float* __restrict__ out_vJJv) {
    __shared__ __align__(128) float4 s_rec[DWARPS][RING_F4];
    __shared__ __align__(128) uint4 s_msk[DWARPS][RING_U4];
    __shared__ __align__(16) float4 s_tan[DWARPS][RING_F4];
    __shared__ float s_acc[DWARPS][(WBATCH + 1) * ACC_STRIDE];
    __shared__ __align__(8) uint64_t s_bar[DWARPS][2];
    const TileArgs& a = vb.t[blockIdx.y];
    const float4* __restrict__ tgeom = vb.tgeom[blockIdx.y];
    const float* __restrict__ weight = vb.weight[blockIdx.y];
    float4* __restrict__ inst_grads = vb.inst_grads[blockIdx.y];
    const int warp = threadIdx.x >> 5;
    WTile c;
    if (!make_tile(a, warp, c)) return;
    if (c.total == 0) return;   // nothing blended in this tile: J v = 0 there and no instance to write
    WarpRing ring;
    init_ring(ring, c, a, s_rec[warp], s_msk[warp], s_bar[warp]);
    const size_t HW = (size_t)a.H * a.W;
    const float bgv[3] = {a.bg[0], a.bg[1], a.bg[2]};

    // pass 1: J v for the eight pixels, kept in registers
    PairT st[4];
    tangent_sweep<false>(c, ring, s_tan[warp], a.cid, tgeom, st);
    // u = W J v (cotangent) and the quadratic form v^T J^T W J v
    PairA sa[4];
    float quad = 0.f;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const PairT& s = st[k >> 1];
        PairA& d = sa[k >> 1];
        const bool hi = k & 1, in = c.inside(k);
        const size_t pix = in ? c.pix(k) : 0;
        const float Tt = hi ? s.Tt.y : s.Tt.x;
        const float jv0 = (hi ? s.C0.y : s.C0.x) + Tt * bgv[0];
        const float jv1 = (hi ? s.C1.y : s.C1.x) + Tt * bgv[1];
        const float jv2 = (hi ? s.C2.y : s.C2.x) + Tt * bgv[2];
        const float w0 = in ? (weight ? weight[pix] : 1.f) : 0.f;
        const float w1 = in ? (weight ? weight[HW + pix] : 1.f) : 0.f;
        const float w2 = in ? (weight ? weight[2 * HW + pix] : 1.f) : 0.f;
        const float g0 = w0 * jv0, g1 = w1 * jv1, g2 = w2 * jv2;
        quad += g0 * jv0 + g1 * jv1 + g2 * jv2;
        set_lane(d.T, hi, in ? a.final_T[pix] : 0.f);
        set_lane(d.g0, hi, g0);
        set_lane(d.g1, hi, g1);
        set_lane(d.g2, hi, g2);
        set_lane(d.gd, hi, 0.f);
    }
    // pass 2: J^T u
    adjoint_sweep<false>(c, ring, s_acc[warp], bgv, sa, a.g_inv, inst_grads);

    if (out_vJJv) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) quad += __shfl_xor_sync(0xffffffffu, quad, o);
        if (c.lane == 0) atomicAdd(out_vJJv, quad);
    }
}

static int tiles_of(int W, int H) { return ((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y); }
static TileArgs tile_args(const BinningView& bv, const ImageView& iv, int W, int H, const float* bg) {
    TileArgs a;
    a.cstart = iv.cstart; a.records = bv.records; a.masks = bv.masks; a.cid = bv.cid; a.g_inv = bv.g_inv; a.final_T = iv.final_T; a.order = iv.order;
    a.W = W; a.H = H; a.tiles = tiles_of(W, H); a.bg = bg;
    return a;
}
static int grid_of(int tiles) { return (tiles + DWARPS - 1) / DWARPS; }

int launch_blend_jvp(const BinningView& bv, const ImageView& iv, int W, int H, const float4* tgeom, const float* bg,
                     const float* bg_tangent, float* out_color_t, float* out_invdepth_t, cudaStream_t s) {
    const TileArgs a = tile_args(bv, iv, W, H, bg);
    if (a.tiles == 0) return 0;
    StageScope ts(ST_BLEND_JVP, s);
    if (out_invdepth_t)
        k_blend_jvp<true><<<grid_of(a.tiles), DTHREADS, 0, s>>>(a, tgeom, bg_tangent, out_color_t, out_invdepth_t);
    else
        k_blend_jvp<false><<<grid_of(a.tiles), DTHREADS, 0, s>>>(a, tgeom, bg_tangent, out_color_t, out_invdepth_t);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

int launch_blend_bwd(const BinningView& bv, const ImageView& iv, int W, int H, const float* bg,
                     const float* dL_dcolor, const float* dL_dinvdepth, float4* inst_grads, cudaStream_t s) {
    const TileArgs a = tile_args(bv, iv, W, H, bg);
    if (a.tiles == 0) return 0;
    StageScope ts(ST_BLEND_BWD, s);
    if (dL_dinvdepth)
        k_blend_bwd<true><<<grid_of(a.tiles), DTHREADS, 0, s>>>(a, dL_dcolor, dL_dinvdepth, inst_grads);
    else
        k_blend_bwd<false><<<grid_of(a.tiles), DTHREADS, 0, s>>>(a, dL_dcolor, dL_dinvdepth, inst_grads);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

int launch_blend_jtjv_batch(int n_views, const BinningView* bvs, const ImageView* ivs, const GslmCamera* cams,
                            const float4* const* tgeoms, const float* const* weights, float4* const* inst_grads,
                            float* out_vJJv, cudaStream_t s) {
    if (n_views > MAX_BATCH_VIEWS) { set_error("too many views in one batch"); return 1; }
    JtjvBatch vb;
    memset(&vb, 0, sizeof(vb));
    int max_tiles = 0;
    for (int i = 0; i < n_views; i++) {
        vb.t[i] = tile_args(bvs[i], ivs[i], cams[i].width, cams[i].height, cams[i].bg);
        vb.tgeom[i] = tgeoms[i];
        vb.weight[i] = weights ? weights[i] : nullptr;
        vb.inst_grads[i] = inst_grads[i];
        max_tiles = max_tiles > vb.t[i].tiles ? max_tiles : vb.t[i].tiles;
    }
    if (max_tiles == 0 || n_views == 0) return 0;
    StageScope ts(ST_BLEND_JTJV, s);
    k_blend_jtjv<<<dim3(grid_of(max_tiles), n_views), DTHREADS, 0, s>>>(vb, out_vJJv);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

int launch_blend_jvp_batch(int n_views, const BinningView* bvs, const ImageView* ivs, const GslmCamera* cams,
                           const float4* const* tgeoms, float* const* out_color_t, cudaStream_t s) {
    if (n_views > MAX_BATCH_VIEWS) { set_error("too many views in one batch"); return 1; }
    ViewBatch vb;
    memset(&vb, 0, sizeof(vb));
    int max_tiles = 0;
    for (int i = 0; i < n_views; i++) {
        vb.t[i] = tile_args(bvs[i], ivs[i], cams[i].width, cams[i].height, cams[i].bg);
        vb.tgeom[i] = tgeoms[i];
        vb.out_color_t[i] = out_color_t[i];
        max_tiles = max_tiles > vb.t[i].tiles ? max_tiles : vb.t[i].tiles;
    }
    if (max_tiles == 0 || n_views == 0) return 0;
    StageScope ts(ST_BLEND_JVP, s);
    k_blend_jvp_batch<<<dim3(grid_of(max_tiles), n_views), DTHREADS, 0, s>>>(vb);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

int launch_blend_bwd_batch(int n_views, const BinningView* bvs, const ImageView* ivs, const GslmCamera* cams,
                           const float* const* dL_dcolor, float4* const* inst_grads, cudaStream_t s) {
    if (n_views > MAX_BATCH_VIEWS) { set_error("too many views in one batch"); return 1; }
    ViewBatch vb;
    memset(&vb, 0, sizeof(vb));
    int max_tiles = 0;
    for (int i = 0; i < n_views; i++) {
        vb.t[i] = tile_args(bvs[i], ivs[i], cams[i].width, cams[i].height, cams[i].bg);
        vb.dL_dcolor[i] = dL_dcolor[i];
        vb.inst_grads[i] = inst_grads[i];
        max_tiles = max_tiles > vb.t[i].tiles ? max_tiles : vb.t[i].tiles;
    }
    if (max_tiles == 0 || n_views == 0) return 0;
    StageScope ts(ST_BLEND_BWD, s);
    k_blend_bwd_batch<<<dim3(grid_of(max_tiles), n_views), DTHREADS, 0, s>>>(vb);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

int launch_blend_jtjv(const BinningView& bv, const ImageView& iv, int W, int H, const float* bg,
                      const float4* tgeom, const float* weight, float4* inst_grads, float* out_vJJv, cudaStream_t s) {
    GslmCamera cam;
    memset(&cam, 0, sizeof(cam));
    cam.width = W; cam.height = H; cam.bg = bg;
    return launch_blend_jtjv_batch(1, &bv, &iv, &cam, &tgeom, weight ? &weight : nullptr, &inst_grads, out_vJJv, s);
}

}  // namespace gslm
