// =============================================================================
// loss_blocks.cu -- the residual side of the LM path on the device: the two per-pixel loss blocks
//     Ll1  = alpha_i sqrt(|x - y| + 1e-6),   ssim = beta_i sqrt(|1 - SSIM(x, y)| + 1e-6),   x = image * alpha_mask
// of compute_batch_loss_block (solver/batch_training_loss.py:10-31) with the 11x11 Gaussian-window SSIM map of
// _ssim_per_pixel (utils/loss_utils.py:98-117), and the forward-mode / reverse-mode derivatives the reference
// obtains from PyTorch (forward AD in LinearSolverFunctions.matvec, autograd in matvec_T;
// solver/solver_functions.py:75-113).  The reference spends 5 grouped 11x11 convolutions (121 taps each) plus ~20
// elementwise kernels per evaluation and again per JVP / VJP; here each direction is ONE kernel.
//
// Structure (HBM-bound): the window is separable, w2d = w w^T.  A CTA of 128 threads owns
// a 128-column x 64-row strip of one image plane and STREAMS over its rows: per row the threads stage the NIN
// pre-convolution quantities of 128 + 10 columns in shared memory (an 8-row ring: one barrier per row), each
// thread forms the horizontal 11-tap sums of its column and feeds them into 11 vertical partial sums held in
// REGISTERS (a systolic chain: the shift is the FMA's destination, no ring indexing), and from the 11th row on
// emits one finished output row per input row.  Every input element is read once per strip (+ 16 % vertical, 8 % horizontal halo,
// L2 hits), nothing intermediate touches HBM.
//
// Linearisation: for fixed y the SSIM map depends on x through mu1 = G*x, E11 = G*(x x), E12 = G*(x y), so
//     d ssim = pa (G*dx) + pb (G*(2 x dx)) + pc (G*(y dx)),
// with per-pixel pa, pb, pc.  k_loss_blocks stores q = (d1, d2 pa, 2 d2 pb, d2 pc) per pixel (16 B), where
// d1 = d Ll1 / dx and d2 = d ssim_block / d ssim; the JVP and VJP kernels are then three separable blurs each.
// =============================================================================
#include "gslm_common.cuh"

namespace gslm {
namespace {

constexpr int LW = 11, LR = 5;          // window size / radius (loss_utils.py:91 window_size=11)
#ifndef GSLM_LTW
#define GSLM_LTW 128
#endif
#ifndef GSLM_LTH
#define GSLM_LTH 64
#endif
#ifndef GSLM_VJP_PRE
#define GSLM_VJP_PRE 1
#endif
constexpr int L_TW = GSLM_LTW;          // columns per CTA = threads per CTA
constexpr int L_TH = GSLM_LTH;          // output rows per CTA
constexpr float L_C1 = 0.01f * 0.01f, L_C2 = 0.03f * 0.03f, L_EPS = 1e-6f;

// gaussian(11, 1.5) / sum, rounded to float32 exactly as loss_utils.py:49-51 produces it
__device__ __forceinline__ float win(int j)
{
    constexpr float w[LW] = {0x1.0d956cp-10f, 0x1.f1fe02p-8f, 0x1.26eb18p-5f, 0x1.bff0fep-4f, 0x1.b43c3ep-3f,
                             0x1.106560p-2f,  0x1.b43c3ep-3f, 0x1.bff0fep-4f, 0x1.26eb18p-5f, 0x1.f1fe02p-8f,
                             0x1.0d956cp-10f};
    return w[j];
}

struct Planes {          // one batch of image planes [B,3,H,W]; alpha_masks [B,1,H,W] or null
    int B, H, W;
    const float* images;
    const float* masks;
    const float* gt;
    const float* alphas;   // [B]
    const float* betas;    // [B]
    const float* tmask;    // [B,3,H,W] or null: d images / d (upstream image), applied to tangents in and gradients out
};

__device__ __forceinline__ float sgn(float v) { return (v > 0.f) - (v < 0.f); }

// The streaming separable-blur skeleton.  Op supplies
//   load(plane, y, x, float (&v)[NIN])   pre-convolution quantities at an IN-RANGE pixel
//   Pre pre(plane, y, x)                 the epilogue's own global operands, fetched one row ahead
//   store(plane, y, x, c, center, pre)   epilogue on the NIN blurred values c of an in-range pixel; center = the
//                                        pixel's own pre-convolution quantities (kept in the shared-memory history)
//   kCenter      true: keep 8 staged rows in shared memory so `center` is available (else 2 rows, center = 0)
//   kPrefetch    true: pre() is issued one row ahead (else at store time; measured per kernel, tools/variants_loss.sh)
template <int NIN, class Op>
__global__ void __launch_bounds__(L_TW) k_loss_stream(Op op, int H, int W)
{
    constexpr int L_HIST = Op::kCenter ? 8 : 2;
    __shared__ float row[L_HIST][NIN][L_TW + 2 * LR];
    const int t = threadIdx.x;
    const int x0 = blockIdx.x * L_TW, y0 = blockIdx.y * L_TH, plane = blockIdx.z;
    const int x = x0 + t;
    const int rows_out = min(L_TH, H - y0);
    const int nrows = rows_out + 2 * LR;
    // second element this thread stages: the 5 + 5 halo columns
    const int hx = (t < LR) ? x0 - LR + t : x0 + L_TW + (t - LR);
    const int hs = (t < LR) ? t : L_TW + t;          // shared-memory slot of that column
    // vertical pass as a systolic chain: acc[k] is the partial sum of output row (rr - 10 + k); a new horizontal
    // row h adds win(10 - k) h to each and the shift acc[k] <- acc[k + 1] rides in the FMA's destination register,
    // so there is no ring index, no register moves and the loop body stays small enough for the instruction cache
    float acc[LW][NIN];
#pragma unroll
    for (int k = 0; k < LW; ++k)
#pragma unroll
        for (int q = 0; q < NIN; ++q) acc[k][q] = 0.f;

    // software pipeline: the global loads of row rr + 1 are issued before the arithmetic of row rr
    float va[NIN], vb[NIN];
    typename Op::Pre pre_next = {}, pre_cur = {};
    auto fetch = [&](int rr) {
        const int y = y0 - LR + rr;
        const bool yin = (rr < nrows) && (y >= 0) && (y < H);
#pragma unroll
        for (int k = 0; k < NIN; ++k) va[k] = vb[k] = 0.f;
        if (yin && x < W) op.load(plane, y, x, va);
        if (t < 2 * LR && yin && hx >= 0 && hx < W) op.load(plane, y, hx, vb);
        if (Op::kPrefetch && rr >= 2 * LR && rr < nrows && x < W) pre_next = op.pre(plane, y - LR, x);
    };
    fetch(0);
#pragma unroll 1
    for (int rr = 0; rr < nrows; ++rr) {
        const int y = y0 - LR + rr;
        float(*buf)[L_TW + 2 * LR] = row[rr & (L_HIST - 1)];
        pre_cur = pre_next;
#pragma unroll
        for (int k = 0; k < NIN; ++k) buf[k][t + LR] = va[k];
        if (t < 2 * LR) {
#pragma unroll
            for (int k = 0; k < NIN; ++k) buf[k][hs] = vb[k];
        }
        fetch(rr + 1);
        __syncthreads();
#pragma unroll
        for (int q = 0; q < NIN; ++q) {
            float h = 0.f;
#pragma unroll
            for (int j = 0; j < LW; ++j) h = fmaf(win(j), buf[q][t + j], h);
#pragma unroll
            for (int k = 0; k < LW - 1; ++k) acc[k][q] = fmaf(win(LW - 1 - k), h, acc[k + 1][q]);
            acc[LW - 1][q] = win(0) * h;
        }
        if (rr >= 2 * LR && x < W) {
            float c[NIN], center[NIN];
#pragma unroll
            for (int q = 0; q < NIN; ++q) {
                c[q] = acc[0][q];
                center[q] = Op::kCenter ? row[(rr - LR) & (L_HIST - 1)][q][t + LR] : 0.f;
            }
            if (!Op::kPrefetch) pre_cur = op.pre(plane, y - LR, x);
            op.store(plane, y - LR, x, c, center, pre_cur);
        }
    }
}

__device__ __forceinline__ size_t pix(const Planes& p, int plane, int y, int x)
{
    return ((size_t)plane * p.H + y) * p.W + x;
}
__device__ __forceinline__ float mask_at(const Planes& p, int plane, int y, int x)
{
    return p.masks ? __ldg(p.masks + ((size_t)(plane / 3) * p.H + y) * p.W + x) : 1.f;
}

// ---- primal: both blocks + the linearisation record ----
struct OpBlocks {
    Planes p;
    float* r1;
    float* r2;
    float4* lin;     // may be null
    __device__ void load(int plane, int y, int x, float (&v)[5]) const
    {
        const size_t i = pix(p, plane, y, x);
        const float a = __ldg(p.images + i) * mask_at(p, plane, y, x), b = __ldg(p.gt + i);
        v[0] = a; v[1] = b; v[2] = a * a; v[3] = b * b; v[4] = a * b;
    }
    static constexpr bool kCenter = true, kPrefetch = false;
    struct Pre { float al, be; };
    __device__ Pre pre(int plane, int, int) const { return Pre{__ldg(p.alphas + plane / 3), __ldg(p.betas + plane / 3)}; }
    __device__ void store(int plane, int y, int x, const float (&c)[5], const float (&center)[5], const Pre& pr) const
    {
        const size_t i = pix(p, plane, y, x);
        const float xv = center[0], yv = center[1];
        const float al = pr.al, be = pr.be;
        // Individually rounded operations in the reference's order (loss_utils.py:102-114), no FMA contraction:
        // where render == target over the whole window the reference gets numerator == denominator bit for bit,
        // hence ssim == 1 exactly, |1 - ssim| == 0 and a zero derivative (sign(0) = 0); so must we.
        const float mu1 = c[0], mu2 = c[1];
        const float mu1s = __fmul_rn(mu1, mu1), mu2s = __fmul_rn(mu2, mu2), mu12 = __fmul_rn(mu1, mu2);
        const float s1 = __fsub_rn(c[2], mu1s), s2 = __fsub_rn(c[3], mu2s), s12 = __fsub_rn(c[4], mu12);
        const float A1 = __fadd_rn(__fmul_rn(2.f, mu12), L_C1), A2 = __fadd_rn(__fmul_rn(2.f, s12), L_C2);
        const float B1 = __fadd_rn(__fadd_rn(mu1s, mu2s), L_C1), B2 = __fadd_rn(__fadd_rn(s1, s2), L_C2);
        const float iB1 = 1.f / B1, iB2 = 1.f / B2;
        const float m = __fdiv_rn(__fmul_rn(A1, A2), __fmul_rn(B1, B2));
        const float e1 = xv - yv, e2 = 1.f - m;
        const float q1 = sqrtf(fabsf(e1) + L_EPS), q2 = sqrtf(fabsf(e2) + L_EPS);
        r1[i] = al * q1;
        r2[i] = be * q2;
        if (lin) {
            const float d1 = al * sgn(e1) / (2.f * q1);
            const float d2 = -be * sgn(e2) / (2.f * q2);
            const float pa = 2.f * mu2 * (A2 - A1) * iB1 * iB2 - 2.f * mu1 * m * (iB1 - iB2);
            const float pb = -m * iB2;
            const float pc = 2.f * A1 * iB1 * iB2;
            lin[i] = make_float4(d1, d2 * pa, 2.f * d2 * pb, d2 * pc);
        }
    }
};

// ---- forward mode: (t1, t2) = J_loss t_images ----
struct OpJvp {
    Planes p;
    const float4* lin;
    const float* t_images;
    float* t1;
    float* t2;
    __device__ void load(int plane, int y, int x, float (&v)[3]) const
    {
        const size_t i = pix(p, plane, y, x);
        const float mk = mask_at(p, plane, y, x);
        const float dx = __ldg(t_images + i) * (p.tmask ? mk * __ldg(p.tmask + i) : mk);
        v[0] = dx; v[1] = __ldg(p.images + i) * mk * dx; v[2] = __ldg(p.gt + i) * dx;
    }
    static constexpr bool kCenter = true, kPrefetch = true;
    struct Pre { float4 q; };
    __device__ Pre pre(int plane, int y, int x) const { return Pre{__ldg(lin + pix(p, plane, y, x))}; }
    __device__ void store(int plane, int y, int x, const float (&c)[3], const float (&center)[3], const Pre& pr) const
    {
        const size_t i = pix(p, plane, y, x);
        t1[i] = pr.q.x * center[0];
        t2[i] = pr.q.y * c[0] + pr.q.z * c[1] + pr.q.w * c[2];
    }
};

// ---- reverse mode: g_images = J_loss^T (u1, u2) ----
struct OpVjp {
    Planes p;
    const float4* lin;
    const float* u1;
    const float* u2;
    float* g;
    int accumulate;
    __device__ void load(int plane, int y, int x, float (&v)[3]) const
    {
        const size_t i = pix(p, plane, y, x);
        const float4 q = __ldg(lin + i);
        const float u = __ldg(u2 + i);
        v[0] = q.y * u; v[1] = q.z * u; v[2] = q.w * u;
    }
    static constexpr bool kCenter = false, kPrefetch = GSLM_VJP_PRE;
    struct Pre { float d1u1, xv, yv, mk, g0; };     // mk: the full output factor alpha_mask * tangent_mask
    __device__ Pre pre(int plane, int y, int x) const
    {
        const size_t i = pix(p, plane, y, x);
        const float mk = mask_at(p, plane, y, x);
        return Pre{__ldg(&lin[i].x) * __ldg(u1 + i), __ldg(p.images + i) * mk, __ldg(p.gt + i),
                   p.tmask ? mk * __ldg(p.tmask + i) : mk, accumulate ? g[i] : 0.f};
    }
    __device__ void store(int plane, int y, int x, const float (&c)[3], const float (&)[3], const Pre& pr) const
    {
        g[pix(p, plane, y, x)] = (pr.d1u1 + c[0] + pr.xv * c[1] + pr.yv * c[2]) * pr.mk + pr.g0;
    }
};

int check(const GslmLossBatch* b, const char* who)
{
    if (!b || b->B < 0 || b->H < 0 || b->W < 0) { set_error("%s: bad batch descriptor", who); return 1; }
    if ((size_t)b->B * 3 > 65535u) { set_error("%s: B * 3 = %d planes exceed the grid limit 65535", who, b->B * 3); return 1; }
    if (b->B && b->H && b->W && (!b->images || !b->gt || !b->alphas || !b->betas)) {
        set_error("%s: images, gt, alphas and betas must be non-null", who);
        return 1;
    }
    return 0;
}
Planes planes(const GslmLossBatch* b)
{
    return Planes{b->B, b->H, b->W, b->images, b->alpha_masks, b->gt, b->alphas, b->betas, b->tangent_masks};
}
dim3 grid_of(const GslmLossBatch* b)
{
    return dim3((b->W + L_TW - 1) / L_TW, (b->H + L_TH - 1) / L_TH, b->B * 3);
}

}  // namespace
}  // namespace gslm

using namespace gslm;

extern "C" size_t gslm_loss_lin_bytes(int32_t B, int32_t H, int32_t W)
{
    return (size_t)B * 3 * (size_t)H * (size_t)W * sizeof(float4);
}

extern "C" int32_t gslm_loss_blocks(const GslmLossBatch* b, float* r1, float* r2, void* lin, void* stream)
{
    if (check(b, "gslm_loss_blocks")) return 1;
    if ((size_t)b->B * b->H * b->W == 0) return 0;
    if (!r1 || !r2) { set_error("gslm_loss_blocks: r1 and r2 must be non-null"); return 1; }
    cudaStream_t s = (cudaStream_t)stream;
    StageScope sc(ST_LOSS, s);
    k_loss_stream<5, OpBlocks><<<grid_of(b), L_TW, 0, s>>>(OpBlocks{planes(b), r1, r2, (float4*)lin}, b->H, b->W);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int32_t gslm_loss_blocks_jvp(const GslmLossBatch* b, const void* lin, const float* t_images, float* t1,
                                        float* t2, void* stream)
{
    if (check(b, "gslm_loss_blocks_jvp")) return 1;
    if ((size_t)b->B * b->H * b->W == 0) return 0;
    if (!lin || !t_images || !t1 || !t2) { set_error("gslm_loss_blocks_jvp: null argument"); return 1; }
    cudaStream_t s = (cudaStream_t)stream;
    StageScope sc(ST_LOSS, s);
    k_loss_stream<3, OpJvp><<<grid_of(b), L_TW, 0, s>>>(OpJvp{planes(b), (const float4*)lin, t_images, t1, t2}, b->H,
                                                       b->W);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int32_t gslm_loss_blocks_vjp(const GslmLossBatch* b, const void* lin, const float* u1, const float* u2,
                                        float* g_images, int32_t accumulate, void* stream)
{
    if (check(b, "gslm_loss_blocks_vjp")) return 1;
    if ((size_t)b->B * b->H * b->W == 0) return 0;
    if (!lin || !u1 || !u2 || !g_images) { set_error("gslm_loss_blocks_vjp: null argument"); return 1; }
    cudaStream_t s = (cudaStream_t)stream;
    StageScope sc(ST_LOSS, s);
    k_loss_stream<3, OpVjp><<<grid_of(b), L_TW, 0, s>>>(OpVjp{planes(b), (const float4*)lin, u1, u2, g_images, accumulate},
                                                       b->H, b->W);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}
