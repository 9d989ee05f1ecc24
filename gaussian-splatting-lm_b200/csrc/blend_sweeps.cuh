// =============================================================================
// blend_sweeps.cuh -- the tangent (forward-mode) and adjoint (reverse-mode) tile sweeps as
// device functions, shared by the stand-alone J v / J^T u kernels and the fused J^T J v kernel.
//
// Tangent sweep: DGR/cuda_rasterizer/forward_impl.h:373-539 (renderCUDAJvpSplit) restated
// without dual-number objects: all branches on the primal (forward_impl.h:477-492), a clamped
// alpha has zero tangent (helper_math.h:70-80).
// Adjoint sweep: DGR/cuda_rasterizer/backward.cu:452-638 (bwd renderCUDA), including its
// quirk that the 0.99 clamp is NOT masked in the gradient (backward.cu:569).
// Both sweeps run on the COMPACT per-tile lists (only instances some pixel blended) and use the
// cached n_contrib_c / final_T of the primal pass, so they never test for termination: a hit at
// a compact position < n_contrib_c was blended by construction.
//
// Mapping: 128 threads per 16x16 tile, one thread = the vertical pixel pair (x, 2r), (x, 2r+1);
// record loads, the column part of the alpha evaluation, loop control and -- in the adjoint --
// the warp reduction are paid once per pair instead of once per pixel.
//
// Algebra (same mathematics as the reference, fewer operations per (pixel, Gaussian) pair):
//  * tangent of the exponent as a polynomial in (dx, dy) with per-Gaussian coefficients
//      power_t = dx (a3 dx + a5 dy + a1) + dy (a4 dy + a2)           (coefficients: preprocess.cu)
//  * adjoint: with q = G dL/dalpha the six geometric partials are moments of q,
//      sum q, q dx, q dy, q dx^2, q dx dy, q dy^2,
//    turned into (dmean2D, dconic, dopacity) ONCE per instance after the reduction; and
//      dL/dalpha_j = T_j (c_j . g) - (S_j + T_final bg . g) / (1 - alpha_j),  S_j = sum_{k>j} (c_k . g) alpha_k T_k
//    which is backward.cu:573-603's accum_rec recurrence contracted with the pixel's cotangent g.
// =============================================================================
#pragma once
#include "blend_common.cuh"

namespace gslm {

constexpr int PAIR_THREADS = TILE_PIX / 2;   // 128 threads per tile in the derivative kernels

struct TileCtx {
    int tid, lane, warp;
    int px, py0;            // pixel column and the first of the two rows
    bool in0, in1;
    float pxf, pyf0, pyf1;
    uint2 range;            // compact range of the tile
    int total;              // compact instances of the tile (every one was blended by some pixel)
};

// Per-thread gather of the tangent records of batch b into shared memory (cp.async, 3 x 16 B per record,
// two records per thread).
__device__ __forceinline__ void stage_tangents(const TileCtx& c, int b, const uint32_t* __restrict__ cid,
                                               const float4* __restrict__ tgeom, float4* dst) {
#pragma unroll
    for (int h = 0; h < DBATCH / PAIR_THREADS; h++) {
        const int slot = c.tid + h * PAIR_THREADS;
        const int k = b * DBATCH + slot;
        if (k < c.total) {
            const uint32_t id = __ldg(cid + c.range.x + k);
            const float4* src = tgeom + (size_t)REC_F4 * id;
            cp_async16(dst + REC_F4 * slot, src);
            cp_async16(dst + REC_F4 * slot + 1, src + 1);
            cp_async16(dst + REC_F4 * slot + 2, src + 2);
        }
    }
    cp_async_commit();
}

struct PixTangent {
    float T, Tt;            // transmittance and its tangent at the end of the sweep
    float Ct[3], Dt;        // tangents of colour (without background) and inverse depth
};
__device__ __forceinline__ void pix_tangent_init(PixTangent& o) {
    o.T = 1.f; o.Tt = 0.f; o.Ct[0] = o.Ct[1] = o.Ct[2] = 0.f; o.Dt = 0.f;
}

template <bool WITH_D>
__device__ __forceinline__ void tangent_update(PixTangent& o, float dx, float dy, float h, float G, float alpha,
                                               const float4& q1, const float4& q2, const float4& t0,
                                               const float4& t1, const float4& t2) {
    // power_t = dx (a3 dx + a1 + a5 dy) + dy (a4 dy + a2);  h = a3 dx + a1 is shared by the pixel pair
    const float pw_t = fmaf(dy, fmaf(t0.w, dy, t0.y), dx * fmaf(t1.x, dy, h));
    // alpha = min(0.99, o G): the constant branch carries no tangent
    const float alpha_t = alpha < 0.99f ? G * fmaf(q1.y, pw_t, t1.y) : 0.f;
    const float aT = alpha * o.T;
    const float aT_t = fmaf(alpha_t, o.T, alpha * o.Tt);
    o.Ct[0] += fmaf(t1.z, aT, q1.z * aT_t);
    o.Ct[1] += fmaf(t1.w, aT, q1.w * aT_t);
    o.Ct[2] += fmaf(t2.x, aT, q2.x * aT_t);
    if (WITH_D) o.Dt += fmaf(t2.y, aT, q2.y * aT_t);
    const float om = 1.f - alpha;
    o.Tt = fmaf(o.Tt, om, -o.T * alpha_t);
    o.T *= om;
}

// Front-to-back tangent sweep over the tile's compact list. s_tan: [2][DBATCH*REC_F4] shared memory.
template <bool WITH_D>
__device__ __forceinline__ void tangent_sweep(const TileCtx& c, RecordRing& ring, float4* s_tan,
                                              const uint32_t* __restrict__ cid, const float4* __restrict__ tgeom,
                                              uint32_t contrib0, uint32_t contrib1, PixTangent& o0, PixTangent& o1) {
    pix_tangent_init(o0);
    pix_tangent_init(o1);
    const int rounds = (c.total + DBATCH - 1) / DBATCH;
    if (rounds == 0) return;
    if (c.tid == 0) ring.issue(0, 0);
    stage_tangents(c, 0, cid, tgeom, s_tan);
    const int cmax = (int)max(contrib0, contrib1);
    for (int b = 0; b < rounds; b++) {
        __syncthreads();   // batch b-1 fully consumed: stage (b+1)&1 is free
        if (b + 1 < rounds) {
            if (c.tid == 0) ring.issue(b + 1, (b + 1) & 1);
            stage_tangents(c, b + 1, cid, tgeom, s_tan + ((b + 1) & 1) * (DBATCH * REC_F4));
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        ring.wait(b & 1);
        __syncthreads();   // everybody's cp.async for batch b is visible
        const float4* rec = ring.stage(b & 1);
        const float4* tan = s_tan + (b & 1) * (DBATCH * REC_F4);
        const int n = min(DBATCH, c.total - b * DBATCH);
        const int mine = min(n, cmax - b * DBATCH);   // a pixel blends compact positions < its n_contrib_c only
        for (int j = 0; j < mine; j++) {
            const float4 q0 = rec[REC_F4 * j], q1 = rec[REC_F4 * j + 1];
            float dx, e, f, dy0, dy1, G0, G1, al0, al1;
            alpha_column(q0, c.pxf, dx, e, f);
            const uint32_t pos = (uint32_t)(b * DBATCH + j);
            const bool hit0 = pos < contrib0 && alpha_pixel(q0, q1, e, f, c.pyf0, dy0, G0, al0);
            const bool hit1 = pos < contrib1 && alpha_pixel(q0, q1, e, f, c.pyf1, dy1, G1, al1);
            if (!(hit0 || hit1)) continue;
            const float4 q2 = rec[REC_F4 * j + 2];
            const float4 t0 = tan[REC_F4 * j], t1 = tan[REC_F4 * j + 1], t2 = tan[REC_F4 * j + 2];
            const float h = fmaf(t0.z, dx, t0.x);
            if (hit0) tangent_update<WITH_D>(o0, dx, dy0, h, G0, al0, q1, q2, t0, t1, t2);
            if (hit1) tangent_update<WITH_D>(o1, dx, dy1, h, G1, al1, q1, q2, t0, t1, t2);
        }
    }
}

// Per-pixel state of the back-to-front adjoint sweep.
struct PixAdjoint {
    float T;          // transmittance BEHIND the current instance, starts at T_final
    float S;          // sum over the instances behind of (c_k . g + invd_k gd) alpha_k T_k, plus T_final (bg . g)
    float g[3], gd;   // the pixel's cotangent
};

template <bool WITH_D>
__device__ __forceinline__ void adjoint_update(PixAdjoint& p, float dx, float dy, float G, float alpha,
                                               const float4& q1, const float4& q2, float (&v)[10]) {
    const float r = rcp_approx(1.f - alpha);
    p.T *= r;                                            // T_j = T_{j+1} / (1 - alpha_j)   (backward.cu:573)
    float cg = fmaf(q1.z, p.g[0], fmaf(q1.w, p.g[1], q2.x * p.g[2]));
    if (WITH_D) cg = fmaf(q2.y, p.gd, cg);
    const float dL_dalpha = fmaf(p.T, cg, -r * p.S);
    const float w = alpha * p.T;
    p.S = fmaf(w, cg, p.S);
    const float q = G * dL_dalpha;
    const float qx = q * dx, qy = q * dy;
    v[0] += q;
    v[1] += qx;
    v[2] += qy;
    v[3] = fmaf(qx, dx, v[3]);
    v[4] = fmaf(qx, dy, v[4]);
    v[5] = fmaf(qy, dy, v[5]);
    v[6] = fmaf(w, p.g[0], v[6]);
    v[7] = fmaf(w, p.g[1], v[7]);
    v[8] = fmaf(w, p.g[2], v[8]);
    if (WITH_D) v[9] = fmaf(w, p.gd, v[9]);
}

// Back-to-front adjoint sweep.  p0 / p1 arrive with T = T_final and the pixel's cotangent (g, gd).
// Writes one 48-byte partial per compact instance of the tile:
// (dmean2D.xy | dconic xx,xy,yy | dopacity | dcolour rgb | dinvdepth | 0 0).
// s_acc: [DBATCH * ACC_STRIDE] floats, zero on entry and on exit.
template <bool WITH_D>
__device__ __forceinline__ void adjoint_sweep(const TileCtx& c, RecordRing& ring, float* s_acc, int W, int H,
                                              const float bg[3], PixAdjoint& p0, PixAdjoint& p1, uint32_t contrib0,
                                              uint32_t contrib1, float4* __restrict__ inst_grads) {
    const int rounds = (c.total + DBATCH - 1) / DBATCH;
    if (rounds == 0) return;
    float4* out = inst_grads + (size_t)c.range.x * REC_F4;
    const int my_slot = slot10(c.lane);
    const float half_W = 0.5f * W, half_H = 0.5f * H;   // backward.cu:527-528
    // the background term of dL/dalpha, -T_final (bg . g) / (1 - alpha), rides in S
    p0.S = p0.T * (bg[0] * p0.g[0] + bg[1] * p0.g[1] + bg[2] * p0.g[2]);
    p1.S = p1.T * (bg[0] * p1.g[0] + bg[1] * p1.g[1] + bg[2] * p1.g[2]);
    const uint32_t warp_contrib = __reduce_max_sync(0xffffffffu, max(contrib0, contrib1));

    __syncthreads();   // a preceding tangent sweep may still be reading the ring
    if (c.tid == 0) ring.issue(rounds - 1, (rounds - 1) & 1);
    for (int b = rounds - 1; b >= 0; b--) {
        __syncthreads();   // batch b+1 consumed and written out; s_acc is zero
        if (c.tid == 0 && b > 0) ring.issue(b - 1, (b - 1) & 1);
        ring.wait(b & 1);
        const float4* rec = ring.stage(b & 1);
        const int n = min(DBATCH, c.total - b * DBATCH);
        const int jtop = min(n, (int)warp_contrib - b * DBATCH);   // warp-uniform upper bound
        for (int j = jtop - 1; j >= 0; j--) {
            const float4 q0 = rec[REC_F4 * j], q1 = rec[REC_F4 * j + 1];
            float dx, e, f, dy0, dy1, G0, G1, al0, al1;
            alpha_column(q0, c.pxf, dx, e, f);
            const uint32_t pos = (uint32_t)(b * DBATCH + j);
            const bool hit0 = pos < contrib0 && alpha_pixel(q0, q1, e, f, c.pyf0, dy0, G0, al0);
            const bool hit1 = pos < contrib1 && alpha_pixel(q0, q1, e, f, c.pyf1, dy1, G1, al1);
            if (!__any_sync(0xffffffffu, hit0 || hit1)) continue;
            float v[10];
#pragma unroll
            for (int i = 0; i < 10; i++) v[i] = 0.f;
            if (hit0 || hit1) {
                const float4 q2 = rec[REC_F4 * j + 2];
                if (hit0) adjoint_update<WITH_D>(p0, dx, dy0, G0, al0, q1, q2, v);
                if (hit1) adjoint_update<WITH_D>(p1, dx, dy1, G1, al1, q1, q2, v);
            }
            const float tot = warp_reduce10(v, c.lane);
            if (my_slot >= 0) atomicAdd(&s_acc[j * ACC_STRIDE + my_slot], tot);
        }
        __syncthreads();   // all warps' partials of batch b are in s_acc
        // moments -> gradient record, once per instance
#pragma unroll
        for (int hh = 0; hh < DBATCH / PAIR_THREADS; hh++) {
            const int slot = c.tid + hh * PAIR_THREADS;
            if (slot < n) {
                float* a = s_acc + slot * ACC_STRIDE;
                const float4 q0 = rec[REC_F4 * slot], q1 = rec[REC_F4 * slot + 1];
                const float A = q0.z, B = q0.w, C = q1.x, o = q1.y;
                const float m0 = a[0], mx = a[1], my = a[2], mxx = a[3], mxy = a[4], myy = a[5];
                float4* dst = out + (size_t)REC_F4 * (b * DBATCH + slot);
                // dL/dG = o dL/dalpha;  dG/ddelx = -G (A dx + B dy), ...   (backward.cu:606-635)
                dst[0] = make_float4(-half_W * o * (A * mx + B * my), -half_H * o * (C * my + B * mx),
                                     -0.5f * o * mxx, -0.5f * o * mxy);
                dst[1] = make_float4(-0.5f * o * myy, m0, a[6], a[7]);
                dst[2] = make_float4(a[8], a[9], 0.f, 0.f);
#pragma unroll
                for (int i = 0; i < 10; i++) a[i] = 0.f;
            }
        }
    }
}

}  // namespace gslm
