// =============================================================================
// param_space.cu -- the activation layer between the solver's raw parameter vectors and the rasterizer inputs,
// differentiated: scaling -> exp, rotation -> normalize, opacity -> sigmoid, (features_dc, features_rest) -> cat
// (scene/gaussian_model.py:192-220 of the reference; PyTorch differentiates them there, one elementwise kernel
// per operation and mode).  Both directions are linear maps at the linearisation point and touch every parameter
// once per CG iteration, so each is ONE HBM-bound pass here:
//   k_activation_jvp : raw tangent (GaussianModelState layout)  -> tangents of (shs, scales, rotations, opacities)
//   k_activation_vjp : activated-space gradient (the allreduce payload) -> raw gradient, + damp * v fused
// 59 P floats in, 59 P floats out (+ the 8 P activations); the SH block is 48 of the 59 floats and is moved by a
// flat element-indexed loop so both sides stay coalesced.
// =============================================================================
#include "gslm_common.cuh"

namespace gslm {
namespace {

// quaternion rows are read / written as four scalars: inside a packed state the rotation block starts at a multiple
// of P floats, which is 16-byte aligned only for even P
__device__ __forceinline__ float4 ld4(const float* p, size_t i)
{
    return make_float4(__ldg(p + 4 * i), __ldg(p + 4 * i + 1), __ldg(p + 4 * i + 2), __ldg(p + 4 * i + 3));
}
__device__ __forceinline__ void st4(float* p, size_t i, float4 v)
{
    p[4 * i] = v.x; p[4 * i + 1] = v.y; p[4 * i + 2] = v.z; p[4 * i + 3] = v.w;
}

struct Act {
    int P, M;
    const float* scales;     // [P,3] exp(raw scaling)
    const float* rotations;  // [P,4] normalised quaternion q
    const float* rot_norm;   // [P]   |raw quaternion|
    const float* opacities;  // [P]   sigmoid(raw opacity)
};

__global__ void __launch_bounds__(256) k_activation_jvp(Act a, GslmRawVector v, GslmActVector t)
{
    const size_t n_sh = (size_t)a.P * 3 * a.M, stride = (size_t)gridDim.x * blockDim.x;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int row = 3 * a.M;
    if ((row & 3) == 0 && (reinterpret_cast<size_t>(t.shs) & 15u) == 0) {
        // rows of 3 M floats with M % 4 == 0: one float4 store per thread and iteration (the two sources are read by
        // scalars: inside a row they sit 3 floats off the 16-byte grid)
        const int qrow = row >> 2;
        const size_t n4 = n_sh >> 2;
        for (size_t c = tid; c < n4; c += stride) {
            const size_t i = c / qrow;
            const int r = (int)(c - i * qrow) << 2;
            const float* rest = v.features_rest + i * (row - 3) + (r - 3);
            float4 o;
            if (r == 0) {
                const float* dc = v.features_dc + 3 * i;
                o = make_float4(__ldg(dc), __ldg(dc + 1), __ldg(dc + 2), __ldg(rest + 3));
            } else {
                o = make_float4(__ldg(rest), __ldg(rest + 1), __ldg(rest + 2), __ldg(rest + 3));
            }
            reinterpret_cast<float4*>(t.shs)[c] = o;
        }
    } else {
        for (size_t e = tid; e < n_sh; e += stride) {
            const size_t i = e / row;
            const int r = (int)(e - i * row);
            t.shs[e] = r < 3 ? __ldg(v.features_dc + 3 * i + r) : __ldg(v.features_rest + i * (row - 3) + (r - 3));
        }
    }
    for (size_t i = tid; i < (size_t)a.P; i += stride) {
#pragma unroll
        for (int c = 0; c < 3; ++c) t.scales[3 * i + c] = __ldg(a.scales + 3 * i + c) * __ldg(v.scaling + 3 * i + c);
        const float4 q = ld4(a.rotations, i), vr = ld4(v.rotation, i);
        const float d = q.x * vr.x + q.y * vr.y + q.z * vr.z + q.w * vr.w, inv = 1.f / __ldg(a.rot_norm + i);
        st4(t.rotations, i, make_float4((vr.x - q.x * d) * inv, (vr.y - q.y * d) * inv, (vr.z - q.z * d) * inv,
                                        (vr.w - q.w * d) * inv));
        const float o = __ldg(a.opacities + i);
        t.opacities[i] = o * (1.f - o) * __ldg(v.opacity + i);
        if (t.means3D)
#pragma unroll
            for (int c = 0; c < 3; ++c) t.means3D[3 * i + c] = __ldg(v.xyz + 3 * i + c);
    }
}

struct Damp { float xyz, dc, rest, scaling, rotation, opacity; };

__global__ void __launch_bounds__(256) k_activation_vjp(Act a, GslmActVector g, GslmRawVector v, Damp d,
                                                        GslmRawVector out, int g_lo, int g_hi)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int row = 3 * a.M;
    const size_t sh_lo = (size_t)g_lo * row, n_sh = (size_t)g_hi * row;   // SH elements of the Gaussians [g_lo, g_hi)
    const bool dv = v.xyz != nullptr;
    if ((row & 3) == 0 && (reinterpret_cast<size_t>(g.shs) & 15u) == 0) {
        const int qrow = row >> 2;
        const size_t n4 = n_sh >> 2;
        for (size_t c = (sh_lo >> 2) + tid; c < n4; c += stride) {
            const size_t i = c / qrow;
            const int r = (int)(c - i * qrow) << 2;
            const float4 gq = __ldg(reinterpret_cast<const float4*>(g.shs) + c);
            const float gv[4] = {gq.x, gq.y, gq.z, gq.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (r + j < 3) {
                    const size_t k = 3 * i + r + j;
                    out.features_dc[k] = dv ? fmaf(d.dc, __ldg(v.features_dc + k), gv[j]) : gv[j];
                } else {
                    const size_t k = i * (row - 3) + (r + j - 3);
                    out.features_rest[k] = dv ? fmaf(d.rest, __ldg(v.features_rest + k), gv[j]) : gv[j];
                }
            }
        }
    } else {
        for (size_t e = sh_lo + tid; e < n_sh; e += stride) {
            const size_t i = e / row;
            const int r = (int)(e - i * row);
            const float gv = __ldg(g.shs + e);
            if (r < 3) {
                const size_t k = 3 * i + r;
                out.features_dc[k] = dv ? fmaf(d.dc, __ldg(v.features_dc + k), gv) : gv;
            } else {
                const size_t k = i * (row - 3) + (r - 3);
                out.features_rest[k] = dv ? fmaf(d.rest, __ldg(v.features_rest + k), gv) : gv;
            }
        }
    }
    for (size_t i = (size_t)g_lo + tid; i < (size_t)g_hi; i += stride) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const size_t k = 3 * i + c;
            const float gm = __ldg(g.means3D + k), gs = __ldg(a.scales + k) * __ldg(g.scales + k);
            out.xyz[k] = dv ? fmaf(d.xyz, __ldg(v.xyz + k), gm) : gm;
            out.scaling[k] = dv ? fmaf(d.scaling, __ldg(v.scaling + k), gs) : gs;
        }
        const float4 q = ld4(a.rotations, i), gr = ld4(g.rotations, i);
        const float dt = q.x * gr.x + q.y * gr.y + q.z * gr.z + q.w * gr.w, inv = 1.f / __ldg(a.rot_norm + i);
        float4 r = make_float4((gr.x - q.x * dt) * inv, (gr.y - q.y * dt) * inv, (gr.z - q.z * dt) * inv,
                               (gr.w - q.w * dt) * inv);
        const float o = __ldg(a.opacities + i);
        float go = o * (1.f - o) * __ldg(g.opacities + i);
        if (dv) {
            const float4 vr = ld4(v.rotation, i);
            r.x = fmaf(d.rotation, vr.x, r.x); r.y = fmaf(d.rotation, vr.y, r.y);
            r.z = fmaf(d.rotation, vr.z, r.z); r.w = fmaf(d.rotation, vr.w, r.w);
            go = fmaf(d.opacity, __ldg(v.opacity + i), go);
        }
        st4(out.rotation, i, r);
        out.opacity[i] = go;
    }
}

int check(const GslmActivations* a, const char* who)
{
    if (!a || a->P < 0 || a->M < 1) { set_error("%s: bad activation descriptor", who); return 1; }
    if (a->P && (!a->scales || !a->rotations || !a->rot_norm || !a->opacities)) {
        set_error("%s: scales, rotations, rot_norm and opacities must be non-null", who);
        return 1;
    }
    return 0;
}
bool raw_ok(const GslmRawVector* v)
{
    return v && v->xyz && v->features_dc && v->scaling && v->rotation && v->opacity;
}
int grid_for(int P, int M)
{
    const size_t n = (size_t)P * 3 * M;
    return (int)min((n + 255) / 256, (size_t)148 * 16);   // grid-stride: 16 CTAs of 256 threads per SM
}

}  // namespace
}  // namespace gslm

using namespace gslm;

extern "C" int32_t gslm_activation_jvp(const GslmActivations* a, const GslmRawVector* v, const GslmActVector* t,
                                       void* stream)
{
    if (check(a, "gslm_activation_jvp")) return 1;
    if (a->P == 0) return 0;
    if (!raw_ok(v) || (a->M > 1 && !v->features_rest)) { set_error("gslm_activation_jvp: bad raw tangent"); return 1; }
    if (!t || !t->shs || !t->scales || !t->rotations || !t->opacities) {
        set_error("gslm_activation_jvp: bad output vector");
        return 1;
    }
    cudaStream_t s = (cudaStream_t)stream;
    StageScope sc(ST_ACTIVATION, s);
    k_activation_jvp<<<grid_for(a->P, a->M), 256, 0, s>>>(
        Act{a->P, a->M, a->scales, a->rotations, a->rot_norm, a->opacities}, *v, *t);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int32_t gslm_activation_vjp(const GslmActivations* a, const GslmActVector* g, const GslmRawVector* v,
                                       const float* damp6, const GslmRawVector* out, void* stream)
{
    return gslm_activation_vjp_range(a, g, v, damp6, out, 0, a ? a->P : 0, stream);
}

extern "C" int32_t gslm_activation_vjp_range(const GslmActivations* a, const GslmActVector* g, const GslmRawVector* v,
                                             const float* damp6, const GslmRawVector* out, int32_t g_lo, int32_t g_hi,
                                             void* stream)
{
    if (check(a, "gslm_activation_vjp")) return 1;
    if (a->P == 0) return 0;
    if (g_lo < 0 || g_hi > a->P || g_lo > g_hi) { set_error("gslm_activation_vjp: bad Gaussian range"); return 1; }
    if (g_lo == g_hi) return 0;
    if (!g || !g->means3D || !g->shs || !g->scales || !g->rotations || !g->opacities) {
        set_error("gslm_activation_vjp: bad activated-space gradient");
        return 1;
    }
    if (!raw_ok(out) || (a->M > 1 && !out->features_rest)) { set_error("gslm_activation_vjp: bad output vector"); return 1; }
    GslmRawVector vin = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    Damp d = {0, 0, 0, 0, 0, 0};
    if (v) {
        if (!raw_ok(v) || (a->M > 1 && !v->features_rest) || !damp6) {
            set_error("gslm_activation_vjp: damping needs a full raw vector and 6 factors");
            return 1;
        }
        vin = *v;
        d = Damp{damp6[0], damp6[1], damp6[2], damp6[3], damp6[4], damp6[5]};
    }
    cudaStream_t s = (cudaStream_t)stream;
    StageScope sc(ST_ACTIVATION, s);
    k_activation_vjp<<<grid_for(g_hi - g_lo, a->M), 256, 0, s>>>(
        Act{a->P, a->M, a->scales, a->rotations, a->rot_norm, a->opacities}, *g, vin, d, *out, g_lo, g_hi);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

// ---- densification bookkeeping (train_jvp.py:311-316, scene/gaussian_model.py:561-563) in one pass ----
//   max_radii2D[i]        = max(max_radii2D[i], radii[i])                    for visible i (radii[i] > 0)
//   xyz_gradient_accum[i] += || dL_dmeans2D[i, :2] ||                         (the viewspace gradient of the batch)
//   denom[i]              += 1
// replacing three boolean-mask gathers / scatters and a norm over [P,3] in torch.
namespace gslm {
namespace {
__global__ void __launch_bounds__(256) k_densification_stats(int P, const int32_t* __restrict__ radii,
                                                             const float* __restrict__ viewspace_grad,
                                                             float* __restrict__ max_radii2D,
                                                             float* __restrict__ xyz_gradient_accum,
                                                             float* __restrict__ denom)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
        const int r = radii[i];
        if (r <= 0) continue;
        if (max_radii2D) max_radii2D[i] = fmaxf(max_radii2D[i], (float)r);
        if (viewspace_grad) {
            const float gx = viewspace_grad[3 * (size_t)i], gy = viewspace_grad[3 * (size_t)i + 1];
            xyz_gradient_accum[i] += sqrtf(__fadd_rn(__fmul_rn(gx, gx), __fmul_rn(gy, gy)));
            denom[i] += 1.f;
        }
    }
}
}  // namespace
}  // namespace gslm

extern "C" int32_t gslm_densification_stats(int32_t P, const int32_t* radii, const float* viewspace_grad,
                                            float* max_radii2D, float* xyz_gradient_accum, float* denom, void* stream)
{
    if (P < 0 || (P > 0 && (!radii || (viewspace_grad && (!xyz_gradient_accum || !denom))))) {
        set_error("gslm_densification_stats: bad argument");
        return 1;
    }
    if (P == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    const int grid = (int)min((size_t)(P + 255) / 256, (size_t)148 * 16);
    k_densification_stats<<<grid, 256, 0, s>>>(P, radii, viewspace_grad, max_radii2D, xyz_gradient_accum, denom);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}
