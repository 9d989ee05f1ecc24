// =============================================================================
// blend_fwd.cu -- primal tile sweep (a9): front-to-back alpha blending.
// Reference: DGR/cuda_rasterizer/forward.cu:278-402 (renderCUDA).
//   C = sum c_i a_i T_i + T bg,  D = sum a_i T_i / z_i,  a_i = min(0.99, o_i exp(power_i));
//   skip if power > 0 or a < 1/255; a pixel stops when T (1 - a) < 1e-4 (that Gaussian is
//   not blended); the CTA stops when all 256 pixels have stopped.
// Writes final_T, n_contrib (ImageState, rasterizer.h) and the per-tile maximum of
// n_contrib, which bounds the tangent / adjoint sweeps of the cached CG path.
// =============================================================================
#include "blend_common.cuh"

namespace gslm {

__global__ void __launch_bounds__(TILE_PIX)
k_blend_fwd(const uint2* __restrict__ ranges, const float4* __restrict__ records, int W, int H,
            const float* __restrict__ bg, float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
            uint32_t* __restrict__ tile_max_contrib, float* __restrict__ out_color, float* __restrict__ out_invdepth) {
    __shared__ __align__(128) float4 s_rec[2 * BATCH * REC_F4];
    __shared__ __align__(8) uint64_t s_bar[2];
    __shared__ uint32_t s_max;

    const int tid = threadIdx.x;
    const int gx = (W + TILE_X - 1) / TILE_X;
    const int tile = blockIdx.x;
    const int px = (tile % gx) * TILE_X + (tid & (TILE_X - 1));
    const int py = (tile / gx) * TILE_Y + (tid >> 4);
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    const uint2 range = ranges[tile];

    RecordRing ring;
    ring.buf = s_rec; ring.bar = s_bar; ring.src = records + (size_t)range.x * REC_F4;
    ring.total = (int)(range.y - range.x); ring.phase = 0;
    const int rounds = (ring.total + BATCH - 1) / BATCH;

    if (tid == 0) {
        mbar_init(&s_bar[0], 1);
        mbar_init(&s_bar[1], 1);
        mbar_fence_init();
        s_max = 0;
    }
    __syncthreads();
    if (tid == 0 && rounds > 0) ring.issue(0, 0);

    bool done = !inside;
    float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, D = 0.f;
    uint32_t contributor = 0, last_contributor = 0;

    int b = 0;
    bool pending = rounds > 0;   // batch b has been issued and not yet waited for
    for (; b < rounds; b++) {
        // also the barrier that frees stage (b+1)&1: everybody has finished batch b-1
        const int num_done = __syncthreads_count(done);
        if (num_done == TILE_PIX) break;
        if (tid == 0 && b + 1 < rounds) ring.issue(b + 1, (b + 1) & 1);
        ring.wait(b & 1);
        const float4* rec = ring.stage(b & 1);
        const int n = ring.batch_size(b);
        for (int j = 0; !done && j < n; j++) {
            contributor++;
            const float4 q0 = rec[REC_F4 * j], q1 = rec[REC_F4 * j + 1];
            float dx, dy, G, alpha;
            if (!eval_alpha(q0, q1, pxf, pyf, dx, dy, G, alpha)) continue;
            const float test_T = T * (1.f - alpha);
            if (test_T < 0.0001f) {
                done = true;
                continue;
            }
            const float4 q2 = rec[REC_F4 * j + 2];
            const float w = alpha * T;
            C0 += q1.z * w;
            C1 += q1.w * w;
            C2 += q2.x * w;
            D += q2.y * w;
            T = test_T;
            last_contributor = contributor;
        }
        pending = false;
        if (b + 1 < rounds) pending = true;
    }
    // never leave the CTA with a bulk copy still in flight into its shared memory
    if (b < rounds && pending) ring.wait(b & 1);

    if (inside) {
        const size_t pix = (size_t)py * W + px, HW = (size_t)H * W;
        final_T[pix] = T;
        n_contrib[pix] = last_contributor;
        out_color[pix] = C0 + T * bg[0];
        out_color[HW + pix] = C1 + T * bg[1];
        out_color[2 * HW + pix] = C2 + T * bg[2];
        if (out_invdepth) out_invdepth[pix] = D;
    }
    const uint32_t wmax = __reduce_max_sync(0xffffffffu, inside ? last_contributor : 0u);
    if ((tid & 31) == 0) atomicMax(&s_max, wmax);
    __syncthreads();
    if (tid == 0) tile_max_contrib[tile] = s_max;
}

int launch_blend_fwd(const BinningView& bv, const ImageView& iv, int W, int H, const float* bg, float* out_color,
                     float* out_invdepth, cudaStream_t s) {
    const int tiles = ((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
    if (tiles == 0) return 0;
    StageScope ts(ST_BLEND_FWD, s);
    k_blend_fwd<<<tiles, TILE_PIX, 0, s>>>(iv.ranges, bv.records, W, H, bg, iv.final_T, iv.n_contrib,
                                          iv.tile_max_contrib, out_color, out_invdepth);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

}  // namespace gslm
