// =============================================================================
// blend_kernels.cu -- tile kernels of the derivative paths:
//   k_blend_jvp   J v     image tangent of one linearised view          (a11)
//   k_blend_bwd   J^T u   per-instance gradient partials from dL/dimage (a13)
//   k_blend_jtjv  J^T W J v fused: tangent sweep, per-pixel weighting in registers, adjoint
//                 sweep -- the image-space vector J v never reaches HBM   
// The sweeps themselves live in blend_sweeps.cuh.
// =============================================================================
#include "blend_sweeps.cuh"

namespace gslm {

constexpr int REC_STAGE_F4 = 2 * BATCH * REC_F4;                 // float4 per record ring
constexpr size_t SMEM_RING = REC_STAGE_F4 * sizeof(float4);      // 24 KB
constexpr size_t SMEM_ACC = BATCH * ACC_STRIDE * sizeof(float);  // 11 KB

__device__ __forceinline__ TileCtx make_ctx(const uint2* __restrict__ ranges,
                                            const uint32_t* __restrict__ tile_max_contrib, int W, int H) {
    TileCtx c;
    c.tid = threadIdx.x; c.lane = c.tid & 31; c.warp = c.tid >> 5;
    const int gx = (W + TILE_X - 1) / TILE_X;
    const int tile = blockIdx.x;
    c.px = (tile % gx) * TILE_X + (c.tid & (TILE_X - 1));
    c.py = (tile / gx) * TILE_Y + (c.tid >> 4);
    c.inside = c.px < W && c.py < H;
    c.pxf = (float)c.px; c.pyf = (float)c.py;
    c.range = ranges[tile];
    c.total = (int)(c.range.y - c.range.x);
    c.sweep_len = min(c.total, (int)tile_max_contrib[tile]);
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

__global__ void __launch_bounds__(TILE_PIX)
k_blend_jvp(const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_max_contrib,
            const float4* __restrict__ records, const uint32_t* __restrict__ point_list,
            const uint32_t* __restrict__ n_contrib, const float4* __restrict__ tgeom, int W, int H,
            const float* __restrict__ bg, const float* __restrict__ bg_t, float* __restrict__ out_color_t,
            float* __restrict__ out_invdepth_t) {
    extern __shared__ __align__(128) unsigned char smem[];
    float4* s_rec = reinterpret_cast<float4*>(smem);
    float4* s_tan = s_rec + REC_STAGE_F4;
    __shared__ __align__(8) uint64_t s_bar[2];
    const TileCtx c = make_ctx(ranges, tile_max_contrib, W, H);
    RecordRing ring;
    init_ring(ring, c, s_rec, s_bar, records);
    const size_t pix = (size_t)c.py * W + c.px, HW = (size_t)H * W;
    const uint32_t my_contrib = c.inside ? n_contrib[pix] : 0u;
    const PixTangent t = tangent_sweep(c, ring, s_tan, point_list, tgeom, my_contrib);
    if (c.inside) {
#pragma unroll
        for (int ch = 0; ch < 3; ch++)   // forward_impl.h:531: C + T * bg on duals
            out_color_t[ch * HW + pix] = t.Ct[ch] + t.Tt * bg[ch] + (bg_t ? t.T * bg_t[ch] : 0.f);
        if (out_invdepth_t) out_invdepth_t[pix] = t.Dt;
    }
}

__global__ void __launch_bounds__(TILE_PIX)
k_blend_bwd(const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_max_contrib,
            const float4* __restrict__ records, const uint32_t* __restrict__ n_contrib,
            const float* __restrict__ final_T, int W, int H, const float* __restrict__ bg,
            const float* __restrict__ dL_dcolor, const float* __restrict__ dL_dinvdepth,
            float4* __restrict__ inst_grads) {
    extern __shared__ __align__(128) unsigned char smem[];
    float4* s_rec = reinterpret_cast<float4*>(smem);
    float* s_acc = reinterpret_cast<float*>(s_rec + REC_STAGE_F4);
    __shared__ __align__(8) uint64_t s_bar[2];
    const TileCtx c = make_ctx(ranges, tile_max_contrib, W, H);
    for (int i = c.tid; i < BATCH * ACC_STRIDE; i += TILE_PIX) s_acc[i] = 0.f;
    RecordRing ring;
    init_ring(ring, c, s_rec, s_bar, records);
    const size_t pix = (size_t)c.py * W + c.px, HW = (size_t)H * W;
    const uint32_t my_contrib = c.inside ? n_contrib[pix] : 0u;
    const float Tf = c.inside ? final_T[pix] : 0.f;
    float g[3] = {0.f, 0.f, 0.f}, gd = 0.f;
    if (c.inside) {
#pragma unroll
        for (int ch = 0; ch < 3; ch++) g[ch] = dL_dcolor[ch * HW + pix];
        if (dL_dinvdepth) gd = dL_dinvdepth[pix];
    }
    const float bgv[3] = {bg[0], bg[1], bg[2]};
    adjoint_sweep(c, ring, s_acc, W, H, bgv, g, gd, Tf, my_contrib, inst_grads);
}

__global__ void __launch_bounds__(TILE_PIX)
k_blend_jtjv(const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_max_contrib,
             const float4* __restrict__ records, const uint32_t* __restrict__ point_list,
             const uint32_t* __restrict__ n_contrib, const float* __restrict__ final_T,
             const float4* __restrict__ tgeom, int W, int H, const float* __restrict__ bg,
             const float* __restrict__ weight, float4* __restrict__ inst_grads, float* __restrict__ out_vJJv) {
    extern __shared__ __align__(128) unsigned char smem[];
    float4* s_rec = reinterpret_cast<float4*>(smem);
    float4* s_tan = s_rec + REC_STAGE_F4;
    float* s_acc = reinterpret_cast<float*>(s_tan + REC_STAGE_F4);
    __shared__ __align__(8) uint64_t s_bar[2];
    __shared__ float s_sum[TILE_PIX / 32];
    const TileCtx c = make_ctx(ranges, tile_max_contrib, W, H);
    for (int i = c.tid; i < BATCH * ACC_STRIDE; i += TILE_PIX) s_acc[i] = 0.f;
    RecordRing ring;
    init_ring(ring, c, s_rec, s_bar, records);
    const size_t pix = (size_t)c.py * W + c.px, HW = (size_t)H * W;
    const uint32_t my_contrib = c.inside ? n_contrib[pix] : 0u;
    const float Tf = c.inside ? final_T[pix] : 0.f;
    const float bgv[3] = {bg[0], bg[1], bg[2]};

    // pass 1: J v for this pixel, kept in registers
    const PixTangent t = tangent_sweep(c, ring, s_tan, point_list, tgeom, my_contrib);
    float g[3] = {0.f, 0.f, 0.f};
    float quad = 0.f;
    if (c.inside) {
#pragma unroll
        for (int ch = 0; ch < 3; ch++) {
            const float jv = t.Ct[ch] + t.Tt * bgv[ch];
            const float w = weight ? weight[ch * HW + pix] : 1.f;
            g[ch] = w * jv;          // cotangent u = W J v
            quad += g[ch] * jv;      // v^T J^T W J v
        }
    }
    // pass 2: J^T u
    adjoint_sweep(c, ring, s_acc, W, H, bgv, g, 0.f, Tf, my_contrib, inst_grads);

    if (out_vJJv) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) quad += __shfl_xor_sync(0xffffffffu, quad, o);
        if (c.lane == 0) s_sum[c.warp] = quad;
        __syncthreads();
        if (c.tid == 0) {
            float s = 0.f;
#pragma unroll
            for (int i = 0; i < TILE_PIX / 32; i++) s += s_sum[i];
            atomicAdd(out_vJJv, s);
        }
    }
}

static int tiles_of(int W, int H) { return ((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y); }

int launch_blend_jvp(const BinningView& bv, const ImageView& iv, int W, int H, const float4* tgeom, const float* bg,
                     const float* bg_tangent, float* out_color_t, float* out_invdepth_t, cudaStream_t s) {
    const int tiles = tiles_of(W, H);
    if (tiles == 0) return 0;
    const size_t smem = 2 * SMEM_RING;
    static bool attr = false;
    if (!attr) {
        GSLM_CUDA_OK(cudaFuncSetAttribute(k_blend_jvp, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr = true;
    }
    StageScope ts(ST_BLEND_JVP, s);
    k_blend_jvp<<<tiles, TILE_PIX, smem, s>>>(iv.ranges, iv.tile_max_contrib, bv.records, bv.point_list, iv.n_contrib,
                                              tgeom, W, H, bg, bg_tangent, out_color_t, out_invdepth_t);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

int launch_blend_bwd(const BinningView& bv, const ImageView& iv, int W, int H, const float* bg,
                     const float* dL_dcolor, const float* dL_dinvdepth, float4* inst_grads, cudaStream_t s) {
    const int tiles = tiles_of(W, H);
    if (tiles == 0) return 0;
    const size_t smem = SMEM_RING + SMEM_ACC;
    StageScope ts(ST_BLEND_BWD, s);
    k_blend_bwd<<<tiles, TILE_PIX, smem, s>>>(iv.ranges, iv.tile_max_contrib, bv.records, iv.n_contrib, iv.final_T, W,
                                              H, bg, dL_dcolor, dL_dinvdepth, inst_grads);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

int launch_blend_jtjv(const BinningView& bv, const ImageView& iv, int W, int H, const float* bg,
                      const float4* tgeom, const float* weight, float4* inst_grads, float* out_vJJv, cudaStream_t s) {
    const int tiles = tiles_of(W, H);
    if (tiles == 0) return 0;
    const size_t smem = 2 * SMEM_RING + SMEM_ACC;
    static bool attr = false;
    if (!attr) {
        GSLM_CUDA_OK(cudaFuncSetAttribute(k_blend_jtjv, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr = true;
    }
    StageScope ts(ST_BLEND_JTJV, s);
    k_blend_jtjv<<<tiles, TILE_PIX, smem, s>>>(iv.ranges, iv.tile_max_contrib, bv.records, bv.point_list, iv.n_contrib,
                                               iv.final_T, tgeom, W, H, bg, weight, inst_grads, out_vJJv);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

}  // namespace gslm
