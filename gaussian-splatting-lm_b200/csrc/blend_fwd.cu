// =============================================================================
// blend_fwd.cu -- primal tile sweep (a9): front-to-back alpha blending, plus the flagging that
// builds the compact linearisation cache.
// Reference: DGR/cuda_rasterizer/forward.cu:278-402 (renderCUDA).
//   C = sum c_i a_i T_i + T bg,  D = sum a_i T_i / z_i,  a_i = min(0.99, o_i exp(power_i));
//   skip if power > 0 or a < 1/255; a pixel stops when T (1 - a) < 1e-4 (that Gaussian is
//   not blended); the CTA stops when all 256 pixels have stopped.
// Like the reference, each thread fetches one Gaussian of the current batch through the
// sorted id list (forward.cu:337-345) -- but it fetches colour and depth too, so the inner
// loop never touches global memory (the reference reads them per hit, :377,380), and the
// loads of batch b+1 are in flight while batch b is blended.  The early exit means only
// ~10 % of the sorted list is ever fetched, which is why the records are NOT pre-gathered.
//
// New: for every instance the 256-bit mask of the pixels that blended it is collected (one
// ballot per warp); per batch the records with a non-empty mask are written, densely and in
// order, with their masks, to the tile's slice of a scratch array, and every pixel's
// n_contrib is also kept in that compact numbering.  The masks carry every hit / skip /
// termination decision of this pass to the derivative sweeps (blend_sweeps.cuh).  binning.cu turns the
// slices into the dense compact cache that all derivative kernels run on.
// Writes final_T, n_contrib (ImageState, rasterizer.h), n_contrib_c, ccount.
// =============================================================================
#include "blend_common.cuh"

namespace gslm {

struct FwdArgs {
    const uint2* ranges;
    const uint32_t* point_list;
    const float2* xy;
    const float4* conic_opacity;
    const float* features;
    const float* depths;
    int W, H;
    const float* bg;
    float* final_T;
    uint32_t* n_contrib;
    uint32_t* n_contrib_c;
    uint32_t* ccount;
    float4* records_tmp;
    uint4* masks_tmp;
    float* out_color;
    float* out_invdepth;
};

__global__ void __launch_bounds__(TILE_PIX) k_blend_fwd(FwdArgs a) {
    __shared__ __align__(16) float4 s_rec[2][BATCH * REC_F4];
    __shared__ __align__(16) uint32_t s_mask[BATCH][8];   // word w of instance j = ballot of warp w: which of its 32 pixels blended j
    __shared__ uint32_t s_pref[BATCH];     // exclusive prefix of s_flag inside each warp's 32-instance group
    __shared__ uint32_t s_wsum[TILE_PIX / 32];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int gx = (a.W + TILE_X - 1) / TILE_X;
    const int tile = blockIdx.x;
    const int px = (tile % gx) * TILE_X + (tid & (TILE_X - 1));
    const int py = (tile / gx) * TILE_Y + (tid >> 4);
    const bool inside = px < a.W && py < a.H;
    const float pxf = (float)px, pyf = (float)py;
    const uint2 range = a.ranges[tile];
    const int total = (int)(range.y - range.x);
    const int rounds = (total + BATCH - 1) / BATCH;

    float4 r0, r1, r2;
    bool have = false;
    auto fetch = [&](int b) {
        const int k = b * BATCH + tid;
        have = k < total;
        if (have) {
            const uint32_t id = __ldg(a.point_list + range.x + k);
            const float2 p = __ldg(a.xy + id);
            const float4 co = __ldg(a.conic_opacity + id);
            const float* f = a.features + 3 * (size_t)id;
            r0 = make_float4(p.x, p.y, co.x, co.y);
            r1 = make_float4(co.z, co.w, __ldg(f), __ldg(f + 1));
            r2 = make_float4(__ldg(f + 2), 1.f / __ldg(a.depths + id), __uint_as_float(id), 0.f);   // forward.cu:380
        }
    };
    if (rounds > 0) fetch(0);

    bool done = !inside;
    float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, D = 0.f;
    uint32_t last_contributor = 0, last_c = 0;
    uint32_t run = 0;   // compact instances of this tile so far (uniform over the CTA)

    for (int b = 0; b < rounds; b++) {
        // barrier: batch b-1 fully consumed (records, flags, prefixes); also the all-done vote
        if (__syncthreads_count(done) == TILE_PIX) break;
        float4* rec = s_rec[b & 1];
        if (have) {
            rec[REC_F4 * tid] = r0;
            rec[REC_F4 * tid + 1] = r1;
            rec[REC_F4 * tid + 2] = r2;
        }
        *reinterpret_cast<uint4*>(&s_mask[tid][0]) = make_uint4(0u, 0u, 0u, 0u);
        *reinterpret_cast<uint4*>(&s_mask[tid][4]) = make_uint4(0u, 0u, 0u, 0u);
        if (b + 1 < rounds) fetch(b + 1);   // in flight during the blend loop below
        __syncthreads();
        const int n = min(BATCH, total - b * BATCH);
        int last_j = -1;
        for (int j = 0; j < n; j++) {
            if ((j & 7) == 0 && __all_sync(0xffffffffu, done)) break;
            bool blended = false;
            if (!done) {
                const float4 q0 = rec[REC_F4 * j], q1 = rec[REC_F4 * j + 1];
                float dx, e, ff, dy, G, alpha;
                alpha_column(q0, pxf, dx, e, ff);
                if (alpha_pixel(q0, q1, e, ff, pyf, dy, G, alpha)) {
                    const float test_T = T * (1.f - alpha);
                    if (test_T < 0.0001f) {
                        done = true;
                    } else {
                        const float4 q2 = rec[REC_F4 * j + 2];
                        const float w = alpha * T;
                        C0 += q1.z * w;
                        C1 += q1.w * w;
                        C2 += q2.x * w;
                        D += q2.y * w;
                        T = test_T;
                        last_contributor = (uint32_t)(b * BATCH + j + 1);
                        last_j = j;
                        blended = true;
                    }
                }
            }
            const uint32_t bal = __ballot_sync(0xffffffffu, blended);
            if (bal != 0u && lane == 0) s_mask[j][warp] = bal;
        }
        __syncthreads();
        // block-wide exclusive prefix over the 256 flags
        const uint4 mk0 = *reinterpret_cast<const uint4*>(&s_mask[tid][0]), mk1 = *reinterpret_cast<const uint4*>(&s_mask[tid][4]);
        const uint32_t f = mk0.x | mk0.y | mk0.z | mk0.w | mk1.x | mk1.y | mk1.z | mk1.w;
        const uint32_t bal = __ballot_sync(0xffffffffu, f != 0);
        s_pref[tid] = __popc(bal & ((1u << lane) - 1u));
        if (lane == 0) s_wsum[warp] = __popc(bal);
        __syncthreads();
        uint32_t woff[TILE_PIX / 32 + 1];
        woff[0] = 0;
#pragma unroll
        for (int w = 0; w < TILE_PIX / 32; w++) woff[w + 1] = woff[w] + s_wsum[w];
        if (f) {
            const size_t pos = (size_t)range.x + run + woff[warp] + s_pref[tid];
            float4* dst = a.records_tmp + pos * REC_F4;
            dst[0] = rec[REC_F4 * tid];
            dst[1] = rec[REC_F4 * tid + 1];
            dst[2] = rec[REC_F4 * tid + 2];
            a.masks_tmp[2 * pos] = mk0;
            a.masks_tmp[2 * pos + 1] = mk1;
        }
        if (last_j >= 0) last_c = run + woff[last_j >> 5] + s_pref[last_j] + 1;   // the pixel's own last hit is flagged
        run += woff[TILE_PIX / 32];
    }

    if (inside) {
        const size_t pix = (size_t)py * a.W + px, HW = (size_t)a.H * a.W;
        a.final_T[pix] = T;
        a.n_contrib[pix] = last_contributor;
        a.n_contrib_c[pix] = last_c;
        a.out_color[pix] = C0 + T * a.bg[0];
        a.out_color[HW + pix] = C1 + T * a.bg[1];
        a.out_color[2 * HW + pix] = C2 + T * a.bg[2];
        if (a.out_invdepth) a.out_invdepth[pix] = D;
    }
    if (tid == 0) a.ccount[tile] = run;
}

// -----------------------------------------------------------------------------------------------------------------
// Warp-per-tile primal sweep (default).  Same arithmetic, same instance order and therefore the same bits as
// k_blend_fwd above, but with the pixel-to-thread mapping of the derivative sweeps (blend_sweeps.cuh): ONE warp owns
// the 16x16 tile, lane l owns column l % 16 and the eight rows 2k + l / 16.  Per (tile, instance) the three record
// loads, the column part of the exponent and the loop overhead are paid once instead of eight times, the eight
// pixels of a thread are eight independent dependency chains, and there is no block barrier anywhere: the batch of
// 32 instances is gathered by the warp itself (one instance per lane, the next batch in flight during the sweep),
// the hit mask of instance j is eight ballots (word k = pixel slot k of every lane: exactly the layout of the
// compact cache) collected by lane j, and the per-batch compaction is one ballot + popc.
// -----------------------------------------------------------------------------------------------------------------
#ifndef GSLM_FWD_MINB
#define GSLM_FWD_MINB 16
#endif
// RECORD = false is the render-only form (gslm_render: loss evaluations of the line search, inference): no hit
// masks, no compaction, no n_contrib -- only the image, the inverse depth and final_T.
template <bool RECORD>
__global__ void __launch_bounds__(32, GSLM_FWD_MINB) k_blend_fwd_warp(FwdArgs a) {
    __shared__ __align__(16) float4 s_rec[WBATCH * REC_F4];
    const int lane = threadIdx.x;
    const int gx = (a.W + TILE_X - 1) / TILE_X;
    const int tile = blockIdx.x;
    const int px = (tile % gx) * TILE_X + (lane & 15);
    const int pyb = (tile / gx) * TILE_Y + (lane >> 4);
    const float pxf = (float)px;
    const uint2 range = a.ranges[tile];
    const int total = (int)(range.y - range.x);
    const int rounds = (total + WBATCH - 1) / WBATCH;

    float4 r0, r1, r2;
    bool have = false;
    auto fetch = [&](int b) {
        const int k = b * WBATCH + lane;
        have = k < total;
        if (have) {
            const uint32_t id = __ldg(a.point_list + range.x + k);
            const float2 p = __ldg(a.xy + id);
            const float4 co = __ldg(a.conic_opacity + id);
            const float* f = a.features + 3 * (size_t)id;
            r0 = make_float4(p.x, p.y, co.x, co.y);
            r1 = make_float4(co.z, co.w, __ldg(f), __ldg(f + 1));
            r2 = make_float4(__ldg(f + 2), 1.f / __ldg(a.depths + id), __uint_as_float(id), 0.f);   // forward.cu:380
        }
    };
    if (rounds > 0) fetch(0);

    float T[8], C0[8], C1[8], C2[8], D[8], pyf[8];
    uint32_t last_contributor[8], last_c[8];
    uint32_t live = 0;                       // bit k: pixel slot k still accumulating
#pragma unroll
    for (int k = 0; k < 8; k++) {
        T[k] = 1.0f; C0[k] = C1[k] = C2[k] = D[k] = 0.f;
        last_contributor[k] = last_c[k] = 0;
        const int py = pyb + 2 * k;
        pyf[k] = (float)py;
        if (px < a.W && py < a.H) live |= 1u << k;
    }
    const uint32_t inside = live;
    uint32_t run = 0;                        // compact instances of this tile so far (warp-uniform)

    for (int b = 0; b < rounds; b++) {
        if (__all_sync(0xffffffffu, live == 0u)) break;
        __syncwarp();                        // batch b-1 fully consumed
        if (have) {
            s_rec[REC_F4 * lane] = r0;
            s_rec[REC_F4 * lane + 1] = r1;
            s_rec[REC_F4 * lane + 2] = r2;
        }
        if (b + 1 < rounds) fetch(b + 1);    // in flight during the sweep below
        __syncwarp();
        const int n = min(WBATCH, total - b * WBATCH);
        uint32_t mm[8];                      // lane j: hit mask of instance j of this batch
        int last_j[8];
#pragma unroll
        for (int k = 0; k < 8; k++) { mm[k] = 0u; last_j[k] = -1; }
        for (int j = 0; j < n; j++) {
            if (__all_sync(0xffffffffu, live == 0u)) break;
            const float4 q0 = s_rec[REC_F4 * j], q1 = s_rec[REC_F4 * j + 1], q2 = s_rec[REC_F4 * j + 2];
            float dx, e, ff;
            alpha_column(q0, pxf, dx, e, ff);
            // branch-free per pixel slot: a non-blending slot adds w = 0 (C + c 0 = C exactly) and keeps T, so the
            // bits equal the branching form while the eight slots stay independent straight-line chains
#pragma unroll
            for (int k = 0; k < 8; k++) {
                float dy, G = 0.f, alpha = 0.f;
                const bool hit = alpha_pixel(q0, q1, e, ff, pyf[k], dy, G, alpha) && (live & (1u << k));
                const float test_T = T[k] * (1.f - alpha);
                const bool term = hit && (test_T < 0.0001f);
                const bool blended = hit && !term;
                if (term) live &= ~(1u << k);
                const float w = blended ? alpha * T[k] : 0.f;
                C0[k] += q1.z * w;
                C1[k] += q1.w * w;
                C2[k] += q2.x * w;
                D[k] += q2.y * w;
                T[k] = blended ? test_T : T[k];
                if (RECORD) {
                    last_j[k] = blended ? j : last_j[k];
                    const uint32_t bal = __ballot_sync(0xffffffffu, blended);
                    mm[k] = (lane == j) ? bal : mm[k];
                }
            }
        }
        if (!RECORD) continue;
        // compaction of the batch: instances with a non-empty mask, densely and in order
        const uint32_t f = mm[0] | mm[1] | mm[2] | mm[3] | mm[4] | mm[5] | mm[6] | mm[7];
        const uint32_t bal = __ballot_sync(0xffffffffu, f != 0u);
        if (f) {
            const size_t pos = (size_t)range.x + run + __popc(bal & ((1u << lane) - 1u));
            float4* dst = a.records_tmp + pos * REC_F4;
            dst[0] = s_rec[REC_F4 * lane];
            dst[1] = s_rec[REC_F4 * lane + 1];
            dst[2] = s_rec[REC_F4 * lane + 2];
            a.masks_tmp[2 * pos] = make_uint4(mm[0], mm[1], mm[2], mm[3]);
            a.masks_tmp[2 * pos + 1] = make_uint4(mm[4], mm[5], mm[6], mm[7]);
        }
#pragma unroll
        for (int k = 0; k < 8; k++)          // the pixel's own last hit is flagged, so it is counted by the popc
            if (last_j[k] >= 0) {
                last_contributor[k] = (uint32_t)(b * WBATCH + last_j[k] + 1);
                last_c[k] = run + __popc(bal & ((2u << last_j[k]) - 1u));
            }
        run += __popc(bal);
    }

    const size_t HW = (size_t)a.H * a.W;
    const float bg0 = a.bg[0], bg1 = a.bg[1], bg2 = a.bg[2];
#pragma unroll
    for (int k = 0; k < 8; k++) {
        if (!(inside & (1u << k))) continue;
        const size_t pix = (size_t)(pyb + 2 * k) * a.W + px;
        a.final_T[pix] = T[k];
        if (RECORD) {
            a.n_contrib[pix] = last_contributor[k];
            a.n_contrib_c[pix] = last_c[k];
        }
        a.out_color[pix] = C0[k] + T[k] * bg0;
        a.out_color[HW + pix] = C1[k] + T[k] * bg1;
        a.out_color[2 * HW + pix] = C2[k] + T[k] * bg2;
        if (a.out_invdepth) a.out_invdepth[pix] = D[k];
    }
    if (RECORD && lane == 0) a.ccount[tile] = run;
}

int launch_blend_fwd(const GeomView& gv, const BinningTemp& bt, const ImageView& iv, int W, int H,
                     const float* features, const float* bg, float* out_color, float* out_invdepth, bool record,
                     cudaStream_t s) {
    const int tiles = ((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
    if (tiles == 0) return 0;
    FwdArgs a;
    a.ranges = iv.ranges; a.point_list = bt.point_list; a.xy = gv.means2D; a.conic_opacity = gv.conic_opacity;
    a.features = features; a.depths = gv.depths; a.W = W; a.H = H; a.bg = bg; a.final_T = iv.final_T;
    a.n_contrib = iv.n_contrib; a.n_contrib_c = iv.n_contrib_c; a.ccount = iv.ccount; a.records_tmp = bt.records_tmp; a.masks_tmp = bt.masks_tmp;
    a.out_color = out_color; a.out_invdepth = out_invdepth;
    StageScope ts(ST_BLEND_FWD, s);
#ifdef GSLM_FWD_CTA
    if (record) k_blend_fwd<<<tiles, TILE_PIX, 0, s>>>(a);
    else k_blend_fwd_warp<false><<<tiles, 32, 0, s>>>(a);
#else
    if (record) k_blend_fwd_warp<true><<<tiles, 32, 0, s>>>(a);
    else k_blend_fwd_warp<false><<<tiles, 32, 0, s>>>(a);
#endif
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

}  // namespace gslm
