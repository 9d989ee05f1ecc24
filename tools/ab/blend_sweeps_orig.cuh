// =============================================================================
// blend_sweeps.cuh -- the tangent (forward-mode) and adjoint (reverse-mode) tile sweeps as
// device functions, shared by the stand-alone J v / J^T u kernels and the fused J^T J v kernel.
//
// Tangent sweep: DGR/cuda_rasterizer/forward_impl.h:373-539 (renderCUDAJvpSplit) restated
// without dual-number objects: all branches on the primal (forward_impl.h:477-492), a clamped
// alpha has zero tangent (helper_math.h:70-80).
// Adjoint sweep: DGR/cuda_rasterizer/backward.cu:452-638 (bwd renderCUDA), including its
// quirk that the 0.99 clamp is NOT masked in the gradient (backward.cu:569).
//
// Both sweeps run on the COMPACT per-tile lists (only instances some pixel blended).  Every
// per-(pixel, instance) decision of the primal pass -- power > 0, alpha < 1/255, "this pixel
// has already terminated" -- is read from the instance's 256-bit hit mask written by
// k_blend_fwd, so the sweeps contain no data-dependent branch and no threshold test.
//
// Mapping: ONE WARP per 16x16 tile.  Lane l owns image column (l & 15) and the eight rows
// 2k + (l >> 4), k = 0..7, i.e. exactly the pixels whose hit bit is bit l of mask word k
// (k_blend_fwd's warp k covers rows 2k, 2k+1 with lane = (row & 1) * 16 + column).  The eight
// pixels are processed as four float2 pairs (rows 4p + h, 4p + 2 + h) with packed FFMA2 /
// FMUL2 / FADD2, so one issue slot does the work of two pixels; record loads, the column part
// of the exponent, loop control and -- in the adjoint -- the warp reduction are paid once per
// (tile, instance) instead of once per (warp of 64 pixels, instance).  A warp never waits for
// another warp: there is no __syncthreads and no shared-memory atomic in these kernels.
//
// Algebra (same mathematics as the reference, fewer operations per (pixel, Gaussian) pair):
//  * the compact record stores the exponent in log2 units with the opacity folded in:
//      alpha_raw = o exp(power) = exp2(a dx^2 + b dx dy + c dy^2 + log2 o),
//      (a, b, c) = -log2(e) (A/2, B, C/2);  alpha = min(0.99, alpha_raw) only when o > 0.99;
//  * tangent: alpha_t = alpha_raw (power_t + o_t / o) with power_t a polynomial in (dx, dy),
//      power_t = dx (a3 dx + a5 dy + a1) + dy (a4 dy + a2)           (coefficients: preprocess.cu)
//  * adjoint: with q = alpha_raw dL/dalpha the six geometric partials are moments of q,
//      sum q, q dx, q dy, q dx^2, q dx dy, q dy^2,
//    turned into (dmean2D, dconic, dopacity) ONCE per instance after the reduction; and
//      dL/dalpha_j = T_j (c_j . g) - (S_j + T_final bg . g) / (1 - alpha_j),  S_j = sum_{k>j} (c_k . g) alpha_k T_k
//    which is backward.cu:573-603's accum_rec recurrence contracted with the pixel's cotangent g.
// =============================================================================
#pragma once
#include "blend_common.cuh"

namespace gslm {

#ifndef GSLM_CLAMP_BRANCH
#define GSLM_CLAMP_BRANCH 1
#endif
#ifndef GSLM_BAND_SKIP
#define GSLM_BAND_SKIP 1
#endif
#ifndef GSLM_DEFER_REDUCE
#define GSLM_DEFER_REDUCE 0
#endif
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;

struct WTile {
    int lane;
    uint32_t lbit;          // 1 << lane: this lane's bit in every mask word
    int px, pyb;            // pixel column; row of pixel k is pyb + 2 k
    float pxf, pyf;
    uint2 range;            // compact range of the tile
    int total;              // compact instances of the tile (every one was blended by some pixel)
    int W, H;
    __device__ __forceinline__ bool inside(int k) const { return px < W && pyb + 2 * k < H; }
    __device__ __forceinline__ size_t pix(int k) const { return (size_t)(pyb + 2 * k) * W + px; }
};

// Per-lane gather of the tangent record of instance (b, lane) into shared memory (cp.async, 3 x 16 B).
__device__ __forceinline__ void stage_tangents(const WTile& c, int b, const uint32_t* __restrict__ cid,
                                               const float4* __restrict__ tgeom, float4* dst) {
    const int k = b * WBATCH + c.lane;
    if (k < c.total) {
        const uint32_t id = __ldg(cid + c.range.x + k);
        const float4* src = tgeom + (size_t)REC_F4 * id;
        cp_async16(dst + REC_F4 * c.lane, src);
        cp_async16(dst + REC_F4 * c.lane + 1, src + 1);
        cp_async16(dst + REC_F4 * c.lane + 2, src + 2);
    }
    cp_async_commit();
}

// What one lane needs of one instance, shared by both sweeps.
struct InstCol {
    float dx, ex, fx, cc, dyb;
};
__device__ __forceinline__ InstCol inst_column(const WTile& c, const float4& q0, const float4& q1, const float4& q2) {
    InstCol i;
    i.dx = q0.x - c.pxf;
    i.ex = fmaf(q0.z * i.dx, i.dx, q1.y);   // a dx^2 + log2 o
    i.fx = q0.w * i.dx;                     // b dx
    i.cc = q1.x;
    i.dyb = q0.y - c.pyf;
    return i;
}
__device__ __forceinline__ uint32_t mask_word(const uint4& m0, const uint4& m1, int k) {
    switch (k) {
        case 0: return m0.x; case 1: return m0.y; case 2: return m0.z; case 3: return m0.w;
        case 4: return m1.x; case 5: return m1.y; case 6: return m1.z; default: return m1.w;
    }
}
// alpha_raw of pixel pair p (rows 4p + h, 4p + 2 + h) and its dy; 0 where the primal pass did not blend
__device__ __forceinline__ f2 pair_alpha(const WTile& c, const InstCol& i, const uint4& m0, const uint4& m1, int p, f2& dy) {
    dy = add2(splat(i.dyb), make_float2(-4.f * p, -4.f * p - 2.f));
    const f2 pl = fma2(dy, fma2(splat(i.cc), dy, splat(i.fx)), splat(i.ex));
    f2 a;
    a.x = (mask_word(m0, m1, 2 * p) & c.lbit) ? ex2_approx(pl.x) : 0.f;
    a.y = (mask_word(m0, m1, 2 * p + 1) & c.lbit) ? ex2_approx(pl.y) : 0.f;
    return a;
}

// ---------------------------------------------------------------------------------------------
// tangent sweep
// ---------------------------------------------------------------------------------------------
struct PairT {
    f2 T, Tt;               // transmittance and its tangent
    f2 C0, C1, C2, D;       // tangents of colour (without background) and inverse depth
};
__device__ __forceinline__ void pair_tangent_init(PairT& o) {
    o.T = splat(1.f);
    o.Tt = o.C0 = o.C1 = o.C2 = o.D = splat(0.f);
}

// Front-to-back tangent sweep over the tile's compact list. s_tan: [2][WBATCH * REC_F4] shared memory.
template <bool WITH_D>
__device__ __forceinline__ void tangent_sweep(const WTile& c, WarpRing& ring, float4* s_tan,
                                              const uint32_t* __restrict__ cid, const float4* __restrict__ tgeom,
                                              PairT (&st)[4]) {
#pragma unroll
    for (int p = 0; p < 4; p++) pair_tangent_init(st[p]);
    const int rounds = (c.total + WBATCH - 1) / WBATCH;
    if (rounds == 0) return;
    if (c.lane == 0) ring.issue(0, 0);
    stage_tangents(c, 0, cid, tgeom, s_tan);
    for (int b = 0; b < rounds; b++) {
        __syncwarp();   // batch b-1 fully consumed: stage (b+1)&1 is free
        if (b + 1 < rounds) {
            if (c.lane == 0) ring.issue(b + 1, (b + 1) & 1);
            stage_tangents(c, b + 1, cid, tgeom, s_tan + ((b + 1) & 1) * (WBATCH * REC_F4));
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        ring.wait(b & 1);
        __syncwarp();   // every lane's cp.async for batch b is visible
        const float4* rec = ring.rec_stage(b & 1);
        const uint4* msk = ring.msk_stage(b & 1);
        const float4* tan = s_tan + (b & 1) * (WBATCH * REC_F4);
        const int n = ring.batch_size(b);
        for (int j = 0; j < n; j++) {
            const float4 q0 = rec[REC_F4 * j], q1 = rec[REC_F4 * j + 1], q2 = rec[REC_F4 * j + 2];
            const uint4 m0 = msk[MSK_U4 * j], m1 = msk[MSK_U4 * j + 1];
            const float4 t0 = tan[REC_F4 * j], t1 = tan[REC_F4 * j + 1], t2 = tan[REC_F4 * j + 2];
            const InstCol i = inst_column(c, q0, q1, q2);
            // power_t + o_t / o = dy (a4 dy + kx) + hx
            const float hx = fmaf(fmaf(t0.z, i.dx, t0.x), i.dx, t1.y);
            const float kx = fmaf(t1.x, i.dx, t0.y);
#if GSLM_BAND_SKIP
            // a pair of rows-of-pixels (4-row band of the tile) that no pixel blended is skipped: the test is on
            // the broadcast mask words, hence warp-uniform (23 % of the bands at 1M Gaussians / 1080p)
#pragma unroll
            for (int p = 0; p < 4; p++) {
                if ((mask_word(m0, m1, 2 * p) | mask_word(m0, m1, 2 * p + 1)) == 0u) continue;
                f2 dy;
                f2 a = pair_alpha(c, i, m0, m1, p, dy);
                f2 pw = fma2(dy, fma2(splat(t0.w), dy, splat(kx)), splat(hx));
                if (q2.w > 0.99f) {
                    pw.x = a.x < 0.99f ? pw.x : 0.f;
                    pw.y = a.y < 0.99f ? pw.y : 0.f;
                    a.x = fminf(0.99f, a.x);
                    a.y = fminf(0.99f, a.y);
                }
                PairT& s = st[p];
                const f2 aT = mul2(a, s.T);
                const f2 aTt = mul2(a, fma2(pw, s.T, s.Tt));
                s.C0 = fma2(splat(t1.z), aT, fma2(splat(q1.z), aTt, s.C0));
                s.C1 = fma2(splat(t1.w), aT, fma2(splat(q1.w), aTt, s.C1));
                s.C2 = fma2(splat(t2.x), aT, fma2(splat(q2.x), aTt, s.C2));
                if (WITH_D) s.D = fma2(splat(t2.y), aT, fma2(splat(q2.y), aTt, s.D));
                s.Tt = add2(s.Tt, neg2(aTt));
                s.T = add2(s.T, neg2(aT));
            }
#else
            f2 a[4], pw[4];
#pragma unroll
            for (int p = 0; p < 4; p++) {
                f2 dy;
                a[p] = pair_alpha(c, i, m0, m1, p, dy);
                pw[p] = fma2(dy, fma2(splat(t0.w), dy, splat(kx)), splat(hx));
            }
            // alpha = min(0.99, alpha_raw); the constant branch has no tangent.  Branch-free on purpose (ALU-pipe work in
            // the shadow of the FMA pipe, and the whole instance stays one basic block for the scheduler).
#if GSLM_CLAMP_BRANCH
            if (q2.w > 0.99f)   // warp-uniform: alpha_raw <= o, so the clamp can only bind when o > 0.99
#endif
#pragma unroll
            for (int p = 0; p < 4; p++) {
                pw[p].x = a[p].x < 0.99f ? pw[p].x : 0.f;
                pw[p].y = a[p].y < 0.99f ? pw[p].y : 0.f;
                a[p].x = fminf(0.99f, a[p].x);
                a[p].y = fminf(0.99f, a[p].y);
            }
#pragma unroll
            for (int p = 0; p < 4; p++) {
                // with alpha_t = alpha pw:  (alpha T)_t = alpha (pw T + T_t) =: aTt,  T_t <- T_t (1 - alpha) - T alpha_t
                // = T_t - aTt,  T <- T - alpha T
                PairT& s = st[p];
                const f2 aT = mul2(a[p], s.T);
                const f2 aTt = mul2(a[p], fma2(pw[p], s.T, s.Tt));
                s.C0 = fma2(splat(t1.z), aT, fma2(splat(q1.z), aTt, s.C0));
                s.C1 = fma2(splat(t1.w), aT, fma2(splat(q1.w), aTt, s.C1));
                s.C2 = fma2(splat(t2.x), aT, fma2(splat(q2.x), aTt, s.C2));
                if (WITH_D) s.D = fma2(splat(t2.y), aT, fma2(splat(q2.y), aTt, s.D));
                s.Tt = add2(s.Tt, neg2(aTt));
                s.T = add2(s.T, neg2(aT));
            }
#endif
        }
    }
}

// ---------------------------------------------------------------------------------------------
// adjoint sweep
// ---------------------------------------------------------------------------------------------
struct PairA {
    f2 T;             // transmittance BEHIND the current instance, starts at T_final
    f2 S;             // sum over the instances behind of (c_k . g + invd_k gd) alpha_k T_k, plus T_final (bg . g)
    f2 g0, g1, g2, gd;   // the pixels' cotangent
};

// Back-to-front adjoint sweep.  sa[p] arrive with T = T_final and the pixels' cotangent (g, gd).
// Writes one 48-byte partial per compact instance of the tile, at the instance's slot in the per-Gaussian CSR
// (g_inv), so that the per-Gaussian adjoint stage reads a Gaussian's partials as one contiguous run:
// (dmean2D.xy | dconic xx,xy,yy | dopacity | dcolour rgb | dinvdepth | 0 0).
// s_acc: [(WBATCH + 1) * ACC_STRIDE] floats of this warp.
template <bool WITH_D>
__device__ __forceinline__ void adjoint_sweep(const WTile& c, WarpRing& ring, float* s_acc, const float bg[3],
                                              PairA (&sa)[4], const uint32_t* __restrict__ g_inv,
                                              float4* __restrict__ inst_grads) {
    const int rounds = (c.total + WBATCH - 1) / WBATCH;
    if (rounds == 0) return;
    const uint32_t* inv = g_inv + c.range.x;   // gradient records go to Gaussian-major (CSR) slots
    const int my_slot = slot10(c.lane);
    const float half_W = 0.5f * c.W, half_H = 0.5f * c.H;   // backward.cu:527-528
    // the background term of dL/dalpha, -T_final (bg . g) / (1 - alpha), rides in S
#pragma unroll
    for (int p = 0; p < 4; p++)
        sa[p].S = mul2(sa[p].T, fma2(splat(bg[0]), sa[p].g0, fma2(splat(bg[1]), sa[p].g1, mul2(splat(bg[2]), sa[p].g2))));

    __syncwarp();   // a preceding tangent sweep may still be reading the ring
    if (c.lane == 0) ring.issue(rounds - 1, (rounds - 1) & 1);
    for (int b = rounds - 1; b >= 0; b--) {
        __syncwarp();   // batch b+1 consumed and written out
        if (c.lane == 0 && b > 0) ring.issue(b - 1, (b - 1) & 1);
        ring.wait(b & 1);
        const float4* rec = ring.rec_stage(b & 1);
        const uint4* msk = ring.msk_stage(b & 1);
        const int n = ring.batch_size(b);
        // the warp reduction of instance j+1 is issued together with the arithmetic of instance j (software
        // pipelining by hand: the five dependent shuffle stages hide behind independent FMA work); the first
        // iteration reduces zeros into the spare row n
        float vp[10];
#pragma unroll
        for (int k = 0; k < 10; k++) vp[k] = 0.f;
        for (int j = n - 1; j >= 0; j--) {
            const float4 q0 = rec[REC_F4 * j], q1 = rec[REC_F4 * j + 1], q2 = rec[REC_F4 * j + 2];
            const uint4 m0 = msk[MSK_U4 * j], m1 = msk[MSK_U4 * j + 1];
            const InstCol i = inst_column(c, q0, q1, q2);
            // per-thread sums over its eight pixels; dx is common to them, so the moments with a dx factor
            // follow from Q0 = sum q, Q1 = sum q dy, Q2 = sum q dy^2 after the loop
            f2 Q0 = splat(0.f), Q1 = Q0, Q2 = Q0, G0 = Q0, G1 = Q0, G2 = Q0, GD = Q0;
#if GSLM_BAND_SKIP
#pragma unroll
            for (int p = 0; p < 4; p++) {
                if ((mask_word(m0, m1, 2 * p) | mask_word(m0, m1, 2 * p + 1)) == 0u) continue;
                f2 dy;
                const f2 araw = pair_alpha(c, i, m0, m1, p, dy);
                f2 a = araw;
                if (q2.w > 0.99f) { a.x = fminf(0.99f, a.x); a.y = fminf(0.99f, a.y); }
                PairA& s = sa[p];
                const f2 om = add2(splat(1.f), neg2(a));
                const f2 r = make_float2(rcp_approx(om.x), rcp_approx(om.y));
                s.T = mul2(s.T, r);
                f2 cg = fma2(splat(q1.z), s.g0, fma2(splat(q1.w), s.g1, mul2(splat(q2.x), s.g2)));
                if (WITH_D) cg = fma2(splat(q2.y), s.gd, cg);
                const f2 dL_dalpha = fma2(s.T, cg, neg2(mul2(r, s.S)));
                const f2 w = mul2(a, s.T);
                s.S = fma2(w, cg, s.S);
                const f2 q = mul2(araw, dL_dalpha);
                const f2 qy = mul2(q, dy);
                Q0 = add2(Q0, q);
                Q1 = add2(Q1, qy);
                Q2 = fma2(qy, dy, Q2);
                G0 = fma2(w, s.g0, G0);
                G1 = fma2(w, s.g1, G1);
                G2 = fma2(w, s.g2, G2);
                if (WITH_D) GD = fma2(w, s.gd, GD);
            }
#else
            f2 araw[4], a[4], dy[4];
#pragma unroll
            for (int p = 0; p < 4; p++) a[p] = araw[p] = pair_alpha(c, i, m0, m1, p, dy[p]);
            // backward.cu:569: T and the weights use the clamped alpha, dG does not
#if GSLM_CLAMP_BRANCH
            if (q2.w > 0.99f)
#endif
#pragma unroll
            for (int p = 0; p < 4; p++) { a[p].x = fminf(0.99f, a[p].x); a[p].y = fminf(0.99f, a[p].y); }
#pragma unroll
            for (int p = 0; p < 4; p++) {
                PairA& s = sa[p];
                const f2 om = add2(splat(1.f), neg2(a[p]));
                const f2 r = make_float2(rcp_approx(om.x), rcp_approx(om.y));
                s.T = mul2(s.T, r);                                  // T_j = T_{j+1} / (1 - alpha_j)   (backward.cu:573)
                f2 cg = fma2(splat(q1.z), s.g0, fma2(splat(q1.w), s.g1, mul2(splat(q2.x), s.g2)));
                if (WITH_D) cg = fma2(splat(q2.y), s.gd, cg);
                const f2 dL_dalpha = fma2(s.T, cg, neg2(mul2(r, s.S)));
                const f2 w = mul2(a[p], s.T);
                s.S = fma2(w, cg, s.S);
                const f2 q = mul2(araw[p], dL_dalpha);
                const f2 qy = mul2(q, dy[p]);
                Q0 = add2(Q0, q);
                Q1 = add2(Q1, qy);
                Q2 = fma2(qy, dy[p], Q2);
                G0 = fma2(w, s.g0, G0);
                G1 = fma2(w, s.g1, G1);
                G2 = fma2(w, s.g2, G2);
                if (WITH_D) GD = fma2(w, s.gd, GD);
            }
#endif
            float v[10];
            v[0] = Q0.x + Q0.y;            // sum q
            v[2] = Q1.x + Q1.y;            // sum q dy
            v[5] = Q2.x + Q2.y;            // sum q dy^2
            v[1] = v[0] * i.dx;            // sum q dx
            v[3] = v[1] * i.dx;            // sum q dx^2
            v[4] = v[2] * i.dx;            // sum q dx dy
            v[6] = G0.x + G0.y;
            v[7] = G1.x + G1.y;
            v[8] = G2.x + G2.y;
            v[9] = WITH_D ? GD.x + GD.y : 0.f;
#if GSLM_DEFER_REDUCE
            const float tot = warp_reduce10(vp, c.lane);
            if (my_slot >= 0) s_acc[(j + 1) * ACC_STRIDE + my_slot] = tot;
#pragma unroll
            for (int k = 0; k < 10; k++) vp[k] = v[k];
#else
            const float tot = warp_reduce10(v, c.lane);
            if (my_slot >= 0) s_acc[j * ACC_STRIDE + my_slot] = tot;
#endif
        }
#if GSLM_DEFER_REDUCE
        {
            const float tot = warp_reduce10(vp, c.lane);
            if (my_slot >= 0) s_acc[my_slot] = tot;
        }
#endif
        __syncwarp();   // the batch's moments are in s_acc
        // moments -> gradient record, once per instance (lane = slot of the batch)
        if (c.lane < n) {
            const float* m = s_acc + c.lane * ACC_STRIDE;
            const float4 q0 = rec[REC_F4 * c.lane], q1 = rec[REC_F4 * c.lane + 1], q2 = rec[REC_F4 * c.lane + 2];
            // conic back from log2 units: A = -2 ln2 a, B = -ln2 b, C = -2 ln2 c
            const float A = -2.f * LN2 * q0.z, B = -LN2 * q0.w, C = -2.f * LN2 * q1.x, o = q2.w;
            const float m0 = m[0], mx = m[1], my = m[2], mxx = m[3], mxy = m[4], myy = m[5];
            float4* dst = inst_grads + (size_t)REC_F4 * __ldg(inv + b * WBATCH + c.lane);
            // moments are of o G dL/dalpha:  dL/dG = o dL/dalpha;  dG/ddelx = -G (A dx + B dy), ...   (backward.cu:606-635)
            dst[0] = make_float4(-half_W * (A * mx + B * my), -half_H * (C * my + B * mx), -0.5f * mxx, -0.5f * mxy);
            dst[1] = make_float4(-0.5f * myy, o > 0.f ? m0 / o : 0.f, m[6], m[7]);
            dst[2] = make_float4(m[8], WITH_D ? m[9] : 0.f, 0.f, 0.f);
        }
    }
}

}  // namespace gslm
