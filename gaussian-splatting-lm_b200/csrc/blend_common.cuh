// =============================================================================
// blend_common.cuh -- building blocks of the tile kernels (a9, a11, a13, fused J^T J v).
//
// One 16x16 tile per CTA in the primal kernel (one thread per pixel, as DGR/cuda_rasterizer/forward.cu:278-402);
// one tile per WARP in the derivative kernels (one thread = 8 pixels of one image column, blend_sweeps.cuh).
// What is different from the reference:
//  * a tile's compact instances are contiguous 48-byte records + 32-byte hit masks (binning.cu), staged 32 at a
//    time by cp.async.bulk (TMA bulk copy, SASS UBLKCP) into a 2-stage shared-memory ring guarded by
//    mbarriers -- no per-thread gathers, colour / depth never fetched from global per hit (forward.cu:377,380);
//  * the primal kernel evaluates alpha with the reference's own rounding sequence (alpha_column / alpha_pixel:
//    the image is a discontinuous function of alpha at the thresholds) and records, per compact instance, the
//    256-bit mask of the pixels that blended it; the derivative sweeps take every hit / skip / termination
//    decision from that mask, so they agree with the primal pass by construction and are free to evaluate
//    alpha with contracted, packed (FFMA2) arithmetic;
//  * the adjoint sweep pre-reduces its ten partials with a 12-shuffle recursive-halving warp reduction and
//    writes ONE record per (tile, Gaussian) instance -- no atomics at all (the reference does ten atomicAdd
//    per (pixel, Gaussian) hit, backward.cu:593-635).
// =============================================================================
#pragma once
#include "gslm_common.cuh"

namespace gslm {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {}
}
// TMA bulk copy global -> shared, completion signalled on an mbarrier (bytes % 16 == 0).
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

constexpr int WBATCH = 32;    // compact instances per staged batch in the derivative kernels (one per lane)
constexpr int MSK_U4 = 2;     // hit mask of one compact instance = 8 x u32 = 2 x uint4 (bit l of word k: see blend_sweeps.cuh)

// Two-stage ring of (records, hit masks) batches for ONE WARP = one tile.  All lanes keep an identical copy
// of the phase bits; lane 0 issues the two bulk copies of a batch on one mbarrier.
struct WarpRing {
    float4* rec;        // [2][WBATCH * REC_F4]
    uint4* msk;         // [2][WBATCH * MSK_U4]
    uint64_t* bar;      // [2]
    const float4* src_rec;   // records of this tile (global)
    const uint4* src_msk;    // masks of this tile (global)
    int total;          // compact instances of the tile
    uint32_t phase;     // bit s = parity of the next completion of stage s

    __device__ __forceinline__ float4* rec_stage(int s) const { return rec + s * (WBATCH * REC_F4); }
    __device__ __forceinline__ uint4* msk_stage(int s) const { return msk + s * (WBATCH * MSK_U4); }
    __device__ __forceinline__ int batch_size(int b) const { return min(WBATCH, total - b * WBATCH); }
    __device__ __forceinline__ void issue(int b, int s) const {   // one lane
        const uint32_t n = (uint32_t)batch_size(b);
        mbar_arrive_expect_tx(&bar[s], n * (REC_F4 * 16 + MSK_U4 * 16));
        bulk_g2s(rec_stage(s), src_rec + (size_t)b * WBATCH * REC_F4, n * (REC_F4 * 16), &bar[s]);
        bulk_g2s(msk_stage(s), src_msk + (size_t)b * WBATCH * MSK_U4, n * (MSK_U4 * 16), &bar[s]);
    }
    __device__ __forceinline__ void wait(int s) {
        mbar_wait(&bar[s], (phase >> s) & 1u);
        phase ^= (1u << s);
    }
};

// ---- packed fp32 pairs (sm_100a FFMA2 / FMUL2 / FADD2: one issue slot for two lanes' worth of FP32 work;
//      measured on B200 with tools/ffma2_bench.cu: same FLOP peak as scalar FFMA, half the issue slots) ----
typedef float2 f2;
__device__ __forceinline__ f2 splat(float a) { return make_float2(a, a); }
__device__ __forceinline__ f2 neg2(f2 a) { return make_float2(-a.x, -a.y); }
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ f2 mul2(f2 a, f2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ f2 add2(f2 a, f2 b) { return __fadd2_rn(a, b); }

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// forward.cu:354-369 / backward.cu:561-569: power, G = exp(power), alpha = min(0.99, o G) and the two
// skip tests, q0 = (x, y, conic.x, conic.y), q1 = (conic.z, opacity, ., .).
// The image is a DISCONTINUOUS function of alpha at the 1/255 and T < 1e-4 thresholds, so `power` follows the
// reference's own instruction sequence bit for bit (nvcc's contraction of forward.cu:358:
// power = fma(-0.5, fma(C dy, dy, (A dx) dx), -((B dx) dy))); a cheaper reassociation flips enough threshold decisions
// to move isolated pixels by 1e-3.  The exponential itself is ex2.approx(power log2e) where the reference (built
// without -use_fast_math) calls libdevice expf: a few ulp apart, so alpha, the image and final_T agree with the
// reference's kernels to rel 1e-4 but are NOT bit-identical, and a pixel whose alpha lies within that distance of
// 1/255 can take one blend the reference skips (or vice versa) -- tests/test_gpu_vs_reference.py bounds both; only
// the binning (radii, tiles_touched, point_list, ranges) is bit-exact.  The evaluation is split in a
// part shared by the pixels of one image column (alpha_column: dx, e = (A dx) dx, f = B dx) and a per-pixel part
// (alpha_pixel), and every operation is explicitly rounded so that every kernel of THIS library -- whatever its
// pixel-to-thread mapping, with or without tile culling -- reproduces the same bits and therefore the same hit /
// skip decisions.
__device__ __forceinline__ void alpha_column(const float4& q0, float pxf, float& dx, float& e, float& f) {
    dx = __fsub_rn(q0.x, pxf);
    e = __fmul_rn(__fmul_rn(q0.z, dx), dx);
    f = __fmul_rn(q0.w, dx);
}
__device__ __forceinline__ bool alpha_pixel(const float4& q0, const float4& q1, float e, float f, float pyf, float& dy,
                                            float& G, float& alpha) {
    dy = __fsub_rn(q0.y, pyf);
    const float t = __fmaf_rn(__fmul_rn(q1.x, dy), dy, e);
    const float power = __fmaf_rn(-0.5f, t, -__fmul_rn(f, dy));
    if (power > 0.0f) return false;
    G = ex2_approx(__fmul_rn(power, 1.4426950408889634f));
    alpha = fminf(0.99f, __fmul_rn(q1.y, G));
    return alpha >= (1.0f / 255.0f);
}
__device__ __forceinline__ float rcp_approx(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// Recursive-halving reduction of 10 values over a warp: 12 SHFL instead of 50.
// On return lane l holds in `out` the warp total of value slot10(l) (or garbage-free 0 for
// pad lanes); lanes with an odd index duplicate their even neighbour.
__device__ __forceinline__ int slot10(int lane) {
    const int b4 = (lane >> 4) & 1, b3 = (lane >> 3) & 1, b2 = (lane >> 2) & 1, b1 = (lane >> 1) & 1;
    int sub;                       // index within a group of five: {0,1,2 | 3,4,pad}
    if (!b3) sub = b2 ? (b1 ? -1 : 2) : b1;
    else sub = b2 ? -1 : 3 + b1;
    if (sub < 0 || (lane & 1)) return -1;
    return b4 * 5 + sub;
}
__device__ __forceinline__ float xhalf(float lo, float hi, bool up, int mask) {
    const float send = up ? lo : hi;
    const float keep = up ? hi : lo;
    return keep + __shfl_xor_sync(0xffffffffu, send, mask);
}
__device__ __forceinline__ float warp_reduce10(const float (&v)[10], int lane) {
    const bool u4 = lane & 16, u3 = lane & 8, u2 = lane & 4, u1 = lane & 2;
    float a0 = xhalf(v[0], v[5], u4, 16), a1 = xhalf(v[1], v[6], u4, 16), a2 = xhalf(v[2], v[7], u4, 16),
          a3 = xhalf(v[3], v[8], u4, 16), a4 = xhalf(v[4], v[9], u4, 16);
    float b0 = xhalf(a0, a3, u3, 8), b1 = xhalf(a1, a4, u3, 8), b2 = xhalf(a2, 0.f, u3, 8);
    float c0 = xhalf(b0, b2, u2, 4), c1 = xhalf(b1, 0.f, u2, 4);
    float d = xhalf(c0, c1, u1, 2);
    d += __shfl_xor_sync(0xffffffffu, d, 1);
    return d;
}

constexpr int ACC_STRIDE = 11;   // 10 partials + 1 pad word: conflict-free column reads

}  // namespace gslm
