// =============================================================================
// blend_kernels.cu -- tile kernels of the derivative paths:
//   k_blend_jvp   J v     image tangent of one linearised view          (a11)
//   k_blend_bwd   J^T u   per-instance gradient partials from dL/dimage (a13)
//   k_blend_jtjv  J^T W J v fused: tangent sweep, per-pixel weighting in registers, adjoint
//                 sweep -- the image-space vector J v never reaches HBM   
// The sweeps themselves live in blend_sweeps.cuh.  One CTA of 128 threads per 16x16 tile,
// one thread per vertical pixel pair.
// =============================================================================
#include "blend_sweeps.cuh"

namespace gslm {

constexpr int REC_STAGE_F4 = 2 * DBATCH * REC_F4;                 // float4 per record ring
constexpr size_t SMEM_RING = REC_STAGE_F4 * sizeof(float4);       // 12 KB
constexpr size_t SMEM_ACC = DBATCH * ACC_STRIDE * sizeof(float);  // 5.5 KB

__device__ __forceinline__ TileCtx make_ctx(const uint32_t* __restrict__ cstart, int W, int H) {
    TileCtx c;
    c.tid = threadIdx.x; c.lane = c.tid & 31; c.warp = c.tid >> 5;
    const int gx = (W + TILE_X - 1) / TILE_X;
    const int tile = blockIdx.x;
    c.px = (tile % gx) * TILE_X + (c.tid & (TILE_X - 1));
    c.py0 = (tile / gx) * TILE_Y + 2 * (c.tid >> 4);
    c.in0 = c.px < W && c.py0 < H;
    c.in1 = c.px < W && c.py0 + 1 < H;
    c.pxf = (float)c.px; c.pyf0 = (float)c.py0; c.pyf1 = (float)(c.py0 + 1);
    c.range = make_uint2(cstart[tile], cstart[tile + 1]);
    c.total = (int)(c.range.y - c.range.x);
    return c;
}

__device__ __forceinline__ void init_ring(RecordRing& ring, const TileCtx& c, float4* s_rec, uint64_t* s_bar,
                                          const float4* __restrict__ records) {
    ring.buf = s_rec; ring.bar = s_bar; ring.src = records + (size_t)c.range.x * REC_F4;
    ring.total = c.total; ring.phase = 0;
    if (c.tid == 0) {
        mbar_init(&s_bar[0], 1);
        mbar_init(&s_bar[1], 1);
        mbar_fence_init();
    }
    __syncthreads();
}

template <bool WITH_D>
__global__ void __launch_bounds__(PAIR_THREADS)
k_blend_jvp(const uint32_t* __restrict__ cstart, const float4* __restrict__ records, const uint32_t* __restrict__ cid,
            const uint32_t* __restrict__ n_contrib_c, const float4* __restrict__ tgeom, int W, int H,
            const float* __restrict__ bg, const float* __restrict__ bg_t, float* __restrict__ out_color_t,
            float* __restrict__ out_invdepth_t) {
    __shared__ __align__(128) float4 s_rec[REC_STAGE_F4];
    __shared__ __align__(16) float4 s_tan[REC_STAGE_F4];
    __shared__ __align__(8) uint64_t s_bar[2];
    const TileCtx c = make_ctx(cstart, W, H);
    RecordRing ring;
    init_ring(ring, c, s_rec, s_bar, records);
    const size_t pix0 = (size_t)c.py0 * W + c.px, pix1 = pix0 + W, HW = (size_t)H * W;
    const uint32_t contrib0 = c.in0 ? n_contrib_c[pix0] : 0u, contrib1 = c.in1 ? n_contrib_c[pix1] : 0u;
    PixTangent t0, t1;
    tangent_sweep<WITH_D>(c, ring, s_tan, cid, tgeom, contrib0, contrib1, t0, t1);
#pragma unroll
    for (int ch = 0; ch < 3; ch++) {   // forward_impl.h:531: C + T * bg on duals
        const float b = bg[ch], bt = bg_t ? bg_t[ch] : 0.f;
        if (c.in0) out_color_t[ch * HW + pix0] = t0.Ct[ch] + t0.Tt * b + t0.T * bt;
        if (c.in1) out_color_t[ch * HW + pix1] = t1.Ct[ch] + t1.Tt * b + t1.T * bt;
    }
    if (WITH_D) {
        if (c.in0) out_invdepth_t[pix0] = t0.Dt;
        if (c.in1) out_invdepth_t[pix1] = t1.Dt;
    }
}

__device__ __forceinline__ void load_cotangent(PixAdjoint& p, bool inside, size_t pix, size_t HW,
                                               const float* __restrict__ final_T, const float* __restrict__ dL_dcolor,
                                               const float* __restrict__ dL_dinvdepth) {
    p.T = inside ? final_T[pix] : 0.f;
    p.S = 0.f;
#pragma unroll
    for (int ch = 0; ch < 3; ch++) p.g[ch] = inside ? dL_dcolor[ch * HW + pix] : 0.f;
    p.gd = (inside && dL_dinvdepth) ? dL_dinvdepth[pix] : 0.f;
}

template <bool WITH_D>
__global__ void __launch_bounds__(PAIR_THREADS)
k_blend_bwd(const uint32_t* __restrict__ cstart, const float4* __restrict__ records,
            const uint32_t* __restrict__ n_contrib_c, const float* __restrict__ final_T, int W, int H,
            const float* __restrict__ bg, const float* __restrict__ dL_dcolor,
            const float* __restrict__ dL_dinvdepth, float4* __restrict__ inst_grads) {
    __shared__ __align__(128) float4 s_rec[REC_STAGE_F4];
    __shared__ float s_acc[DBATCH * ACC_STRIDE];
    __shared__ __align__(8) uint64_t s_bar[2];
    const TileCtx c = make_ctx(cstart, W, H);
    for (int i = c.tid; i < DBATCH * ACC_STRIDE; i += PAIR_THREADS) s_acc[i] = 0.f;
    RecordRing ring;
    init_ring(ring, c, s_rec, s_bar, records);
    const size_t pix0 = (size_t)c.py0 * W + c.px, pix1 = pix0 + W, HW = (size_t)H * W;
    const uint32_t contrib0 = c.in0 ? n_contrib_c[pix0] : 0u, contrib1 = c.in1 ? n_contrib_c[pix1] : 0u;
    PixAdjoint p0, p1;
    load_cotangent(p0, c.in0, pix0, HW, final_T, dL_dcolor, dL_dinvdepth);
    load_cotangent(p1, c.in1, pix1, HW, final_T, dL_dcolor, dL_dinvdepth);
    const float bgv[3] = {bg[0], bg[1], bg[2]};
    adjoint_sweep<WITH_D>(c, ring, s_acc, W, H, bgv, p0, p1, contrib0, contrib1, inst_grads);
}

__device__ __forceinline__ float weigh(PixAdjoint& p, const PixTangent& t, bool inside, size_t pix, size_t HW,
                                       const float* __restrict__ final_T, const float* __restrict__ weight,
                                       const float bg[3]) {
    p.T = inside ? final_T[pix] : 0.f;
    p.S = 0.f;
    p.gd = 0.f;
    float quad = 0.f;
#pragma unroll
    for (int ch = 0; ch < 3; ch++) {
        const float jv = t.Ct[ch] + t.Tt * bg[ch];
        const float w = inside ? (weight ? weight[ch * HW + pix] : 1.f) : 0.f;
        p.g[ch] = w * jv;          // cotangent u = W J v
        quad += p.g[ch] * jv;      // v^T J^T W J v
    }
    return quad;
}

__global__ void __launch_bounds__(PAIR_THREADS)
k_blend_jtjv(const uint32_t* __restrict__ cstart, const float4* __restrict__ records, const uint32_t* __restrict__ cid,
             const uint32_t* __restrict__ n_contrib_c, const float* __restrict__ final_T,
             const float4* __restrict__ tgeom, int W, int H, const float* __restrict__ bg,
             const float* __restrict__ weight, float4* __restrict__ inst_grads, float* __restrict__ out_vJJv) {
    __shared__ __align__(128) float4 s_rec[REC_STAGE_F4];
    __shared__ __align__(16) float4 s_tan[REC_STAGE_F4];
    __shared__ float s_acc[DBATCH * ACC_STRIDE];
    __shared__ __align__(8) uint64_t s_bar[2];
    __shared__ float s_sum[PAIR_THREADS / 32];
    const TileCtx c = make_ctx(cstart, W, H);
    for (int i = c.tid; i < DBATCH * ACC_STRIDE; i += PAIR_THREADS) s_acc[i] = 0.f;
    RecordRing ring;
    init_ring(ring, c, s_rec, s_bar, records);
    const size_t pix0 = (size_t)c.py0 * W + c.px, pix1 = pix0 + W, HW = (size_t)H * W;
    const uint32_t contrib0 = c.in0 ? n_contrib_c[pix0] : 0u, contrib1 = c.in1 ? n_contrib_c[pix1] : 0u;
    const float bgv[3] = {bg[0], bg[1], bg[2]};

    // pass 1: J v for the two pixels, kept in registers
    PixTangent t0, t1;
    tangent_sweep<false>(c, ring, s_tan, cid, tgeom, contrib0, contrib1, t0, t1);
    PixAdjoint p0, p1;
    float quad = weigh(p0, t0, c.in0, pix0, HW, final_T, weight, bgv) + weigh(p1, t1, c.in1, pix1, HW, final_T, weight, bgv);
    // pass 2: J^T u
    adjoint_sweep<false>(c, ring, s_acc, W, H, bgv, p0, p1, contrib0, contrib1, inst_grads);

    if (out_vJJv) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) quad += __shfl_xor_sync(0xffffffffu, quad, o);
        if (c.lane == 0) s_sum[c.warp] = quad;
        __syncthreads();
        if (c.tid == 0) {
            float s = 0.f;
#pragma unroll
            for (int i = 0; i < PAIR_THREADS / 32; i++) s += s_sum[i];
            atomicAdd(out_vJJv, s);
        }
    }
}

static int tiles_of(int W, int H) { return ((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y); }

int launch_blend_jvp(const BinningView& bv, const ImageView& iv, int W, int H, const float4* tgeom, const float* bg,
                     const float* bg_tangent, float* out_color_t, float* out_invdepth_t, cudaStream_t s) {
    const int tiles = tiles_of(W, H);
    if (tiles == 0) return 0;
    StageScope ts(ST_BLEND_JVP, s);
    if (out_invdepth_t)
        k_blend_jvp<true><<<tiles, PAIR_THREADS, 0, s>>>(iv.cstart, bv.records, bv.cid, iv.n_contrib_c, tgeom, W, H, bg,
                                                        bg_tangent, out_color_t, out_invdepth_t);
    else
        k_blend_jvp<false><<<tiles, PAIR_THREADS, 0, s>>>(iv.cstart, bv.records, bv.cid, iv.n_contrib_c, tgeom, W, H, bg,
                                                         bg_tangent, out_color_t, out_invdepth_t);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

int launch_blend_bwd(const BinningView& bv, const ImageView& iv, int W, int H, const float* bg,
                     const float* dL_dcolor, const float* dL_dinvdepth, float4* inst_grads, cudaStream_t s) {
    const int tiles = tiles_of(W, H);
    if (tiles == 0) return 0;
    StageScope ts(ST_BLEND_BWD, s);
    if (dL_dinvdepth)
        k_blend_bwd<true><<<tiles, PAIR_THREADS, 0, s>>>(iv.cstart, bv.records, iv.n_contrib_c, iv.final_T, W, H, bg,
                                                        dL_dcolor, dL_dinvdepth, inst_grads);
    else
        k_blend_bwd<false><<<tiles, PAIR_THREADS, 0, s>>>(iv.cstart, bv.records, iv.n_contrib_c, iv.final_T, W, H, bg,
                                                         dL_dcolor, dL_dinvdepth, inst_grads);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

int launch_blend_jtjv(const BinningView& bv, const ImageView& iv, int W, int H, const float* bg,
                      const float4* tgeom, const float* weight, float4* inst_grads, float* out_vJJv, cudaStream_t s) {
    const int tiles = tiles_of(W, H);
    if (tiles == 0) return 0;
    StageScope ts(ST_BLEND_JTJV, s);
    k_blend_jtjv<<<tiles, PAIR_THREADS, 0, s>>>(iv.cstart, bv.records, bv.cid, iv.n_contrib_c, iv.final_T, tgeom, W, H,
                                               bg, weight, inst_grads, out_vJJv);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

}  // namespace gslm
