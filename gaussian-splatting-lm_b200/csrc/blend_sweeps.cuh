// =============================================================================
// blend_sweeps.cuh -- the tangent (forward-mode) and adjoint (reverse-mode) tile sweeps as
// device functions, shared by the stand-alone J v / J^T u kernels and the fused J^T J v kernel.
//
// Tangent sweep: DGR/cuda_rasterizer/forward_impl.h:373-539 (renderCUDAJvpSplit) restated
// without dual-number objects: all branches on the primal (forward_impl.h:477-492), a clamped
// alpha has zero tangent (helper_math.h:70-80).
// Adjoint sweep: DGR/cuda_rasterizer/backward.cu:452-638 (bwd renderCUDA), including its
// quirk that the 0.99 clamp is NOT masked in the gradient (backward.cu:569).
// Both sweeps use the cached n_contrib / final_T of the primal pass, so they never test for
// termination: a hit at a position < n_contrib was blended by construction.
// =============================================================================
#pragma once
#include "blend_common.cuh"

namespace gslm {

struct TileCtx {
    int tid, lane, warp;
    int px, py;
    bool inside;
    float pxf, pyf;
    uint2 range;
    int total;          // instances in range
    int sweep_len;      // min(total, tile_max_contrib): instances any pixel of the tile blended
};

// Per-thread gather of the tangent records of batch b into shared memory (cp.async, 3 x 16 B).
__device__ __forceinline__ void stage_tangents(const TileCtx& c, int b, const uint32_t* __restrict__ point_list,
                                               const float4* __restrict__ tgeom, float4* dst) {
    const int k = b * BATCH + c.tid;
    if (k < c.sweep_len) {
        const uint32_t id = __ldg(point_list + c.range.x + k);
        const float4* src = tgeom + (size_t)REC_F4 * id;
        cp_async16(dst + REC_F4 * c.tid, src);
        cp_async16(dst + REC_F4 * c.tid + 1, src + 1);
        cp_async16(dst + REC_F4 * c.tid + 2, src + 2);
    }
    cp_async_commit();
}

struct PixTangent {
    float T, Tt;            // transmittance and its tangent at the end of the sweep
    float Ct[3], Dt;        // tangents of colour (without background) and inverse depth
};

// Front-to-back tangent sweep over the first `sweep_len` instances of the tile.
// s_tan: [2][BATCH*REC_F4] shared memory for tangent records.
__device__ __forceinline__ PixTangent tangent_sweep(const TileCtx& c, RecordRing& ring, float4* s_tan,
                                                    const uint32_t* __restrict__ point_list,
                                                    const float4* __restrict__ tgeom, uint32_t my_contrib) {
    PixTangent o;
    o.T = 1.f; o.Tt = 0.f; o.Ct[0] = o.Ct[1] = o.Ct[2] = 0.f; o.Dt = 0.f;
    const int rounds = (c.sweep_len + BATCH - 1) / BATCH;
    if (rounds == 0) return o;
    if (c.tid == 0) ring.issue(0, 0);
    stage_tangents(c, 0, point_list, tgeom, s_tan);
    for (int b = 0; b < rounds; b++) {
        __syncthreads();   // batch b-1 fully consumed: stage (b+1)&1 is free
        if (b + 1 < rounds) {
            if (c.tid == 0) ring.issue(b + 1, (b + 1) & 1);
            stage_tangents(c, b + 1, point_list, tgeom, s_tan + ((b + 1) & 1) * (BATCH * REC_F4));
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        ring.wait(b & 1);
        __syncthreads();   // everybody's cp.async for batch b is visible
        const float4* rec = ring.stage(b & 1);
        const float4* tan = s_tan + (b & 1) * (BATCH * REC_F4);
        const int n = min(BATCH, c.sweep_len - b * BATCH);
        const int mine = min(n, (int)my_contrib - b * BATCH);   // this pixel blends positions < n_contrib only
        for (int j = 0; j < mine; j++) {
            const float4 q0 = rec[REC_F4 * j], q1 = rec[REC_F4 * j + 1];
            float dx, dy, G, alpha;
            if (!eval_alpha(q0, q1, c.pxf, c.pyf, dx, dy, G, alpha)) continue;
            const float4 q2 = rec[REC_F4 * j + 2];
            const float4 t0 = tan[REC_F4 * j], t1 = tan[REC_F4 * j + 1], t2 = tan[REC_F4 * j + 2];
            // power = -0.5 (A dx^2 + C dy^2) - B dx dy ;  d(dx) = x_t, d(dy) = y_t
            const float pw_t = -0.5f * (t0.z * dx * dx + t1.x * dy * dy) - t0.w * dx * dy -
                               (q0.z * dx + q0.w * dy) * t0.x - (q1.x * dy + q0.w * dx) * t0.y;
            const float G_t = G * pw_t;
            // alpha = min(0.99, o G): the constant branch carries no tangent
            const float alpha_t = (0.99f < q1.y * G) ? 0.f : (t1.y * G + q1.y * G_t);
            const float aT = alpha * o.T;
            const float aT_t = alpha_t * o.T + alpha * o.Tt;
            o.Ct[0] += t1.z * aT + q1.z * aT_t;
            o.Ct[1] += t1.w * aT + q1.w * aT_t;
            o.Ct[2] += t2.x * aT + q2.x * aT_t;
            o.Dt += t2.y * aT + q2.y * aT_t;
            o.Tt = o.Tt * (1.f - alpha) - o.T * alpha_t;
            o.T = o.T * (1.f - alpha);
        }
    }
    return o;
}

// Back-to-front adjoint sweep. g[0..2] = dL/dcolour of this pixel, gd = dL/dinvdepth.
// Writes one 48-byte partial per instance of the tile: (dmean2D.xy | dconic xx,xy,yy | dopacity |
// dcolour rgb | dinvdepth | 0 0), positions >= sweep_len get zeros.
// s_acc: [BATCH * ACC_STRIDE] floats, zero on entry and on exit.
__device__ __forceinline__ void adjoint_sweep(const TileCtx& c, RecordRing& ring, float* s_acc, int W, int H,
                                              const float bg[3], const float g[3], float gd, float T_final,
                                              uint32_t my_contrib, float4* __restrict__ inst_grads) {
    const int rounds = (c.sweep_len + BATCH - 1) / BATCH;
    float4* out = inst_grads + (size_t)c.range.x * REC_F4;
    // instances no pixel reached: exact zeros (the batch write-outs below cover [0, sweep_len))
    for (int k = c.sweep_len + c.tid; k < c.total; k += TILE_PIX) {
        out[(size_t)REC_F4 * k] = make_float4(0.f, 0.f, 0.f, 0.f);
        out[(size_t)REC_F4 * k + 1] = make_float4(0.f, 0.f, 0.f, 0.f);
        out[(size_t)REC_F4 * k + 2] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (rounds == 0) return;
    const int my_slot = slot10(c.lane);
    const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;   // backward.cu:527-528
    const float bg_dot = bg[0] * g[0] + bg[1] * g[1] + bg[2] * g[2];
    const uint32_t warp_contrib = __reduce_max_sync(0xffffffffu, my_contrib);
    float T = T_final;
    float accum[3] = {0.f, 0.f, 0.f}, accum_d = 0.f, last_alpha = 0.f, last_c[3] = {0.f, 0.f, 0.f}, last_d = 0.f;

    __syncthreads();   // a preceding tangent sweep may still be reading the ring
    if (c.tid == 0) ring.issue(rounds - 1, (rounds - 1) & 1);
    for (int b = rounds - 1; b >= 0; b--) {
        __syncthreads();   // batch b+1 consumed and written out; s_acc is zero
        if (c.tid == 0 && b > 0) ring.issue(b - 1, (b - 1) & 1);
        ring.wait(b & 1);
        const float4* rec = ring.stage(b & 1);
        const int n = min(BATCH, c.sweep_len - b * BATCH);
        const int jtop = min(n, (int)warp_contrib - b * BATCH);   // warp-uniform upper bound
        for (int j = jtop - 1; j >= 0; j--) {
            const float4 q0 = rec[REC_F4 * j], q1 = rec[REC_F4 * j + 1];
            float dx, dy, G, alpha;
            bool hit = (uint32_t)(b * BATCH + j) < my_contrib;
            hit = hit && eval_alpha(q0, q1, c.pxf, c.pyf, dx, dy, G, alpha);
            if (!__any_sync(0xffffffffu, hit)) continue;
            float v[10];
#pragma unroll
            for (int i = 0; i < 10; i++) v[i] = 0.f;
            if (hit) {
                const float4 q2 = rec[REC_F4 * j + 2];
                T = T / (1.f - alpha);
                const float w = alpha * T;
                const float col[3] = {q1.z, q1.w, q2.x};
                float dL_dalpha = 0.f;
#pragma unroll
                for (int ch = 0; ch < 3; ch++) {
                    accum[ch] = last_alpha * last_c[ch] + (1.f - last_alpha) * accum[ch];
                    last_c[ch] = col[ch];
                    dL_dalpha += (col[ch] - accum[ch]) * g[ch];
                    v[6 + ch] = w * g[ch];
                }
                accum_d = last_alpha * last_d + (1.f - last_alpha) * accum_d;
                last_d = q2.y;
                dL_dalpha += (q2.y - accum_d) * gd;
                v[9] = w * gd;
                dL_dalpha *= T;
                last_alpha = alpha;
                dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot;
                const float dL_dG = q1.y * dL_dalpha;
                const float gdx = G * dx, gdy = G * dy;
                const float dG_ddelx = -gdx * q0.z - gdy * q0.w;
                const float dG_ddely = -gdy * q1.x - gdx * q0.w;
                v[0] = dL_dG * dG_ddelx * ddelx_dx;
                v[1] = dL_dG * dG_ddely * ddely_dy;
                v[2] = -0.5f * gdx * dx * dL_dG;
                v[3] = -0.5f * gdx * dy * dL_dG;
                v[4] = -0.5f * gdy * dy * dL_dG;
                v[5] = G * dL_dalpha;
            }
            const float tot = warp_reduce10(v, c.lane);
            if (my_slot >= 0) atomicAdd(&s_acc[j * ACC_STRIDE + my_slot], tot);
        }
        __syncthreads();   // all warps' partials of batch b are in s_acc
        if (c.tid < n) {
            float* a = s_acc + c.tid * ACC_STRIDE;
            float4* o = out + (size_t)REC_F4 * (b * BATCH + c.tid);
            o[0] = make_float4(a[0], a[1], a[2], a[3]);
            o[1] = make_float4(a[4], a[5], a[6], a[7]);
            o[2] = make_float4(a[8], a[9], 0.f, 0.f);
#pragma unroll
            for (int i = 0; i < 10; i++) a[i] = 0.f;
        }
    }
}

}  // namespace gslm
