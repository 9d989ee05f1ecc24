// =============================================================================
// preprocess.cu -- per-Gaussian stages (HBM-bound, one thread per Gaussian).
//   k_preprocess          forward.cu:154-273   (a7)  3D->2D projection, conic, radius, SH->RGB
//   k_preprocess_tangent  forward_impl.h:208-367 (a10) same on dual numbers, tangent record only
//   k_mark_visible        rasterizer_impl.cu:54-66
//   k_preprocess_bwd      backward.cu:147-449  (a14) fused: instance-gradient gather-sum +
//                         cov2D / projective / SH / cov3D adjoints in ONE pass (the reference
//                         uses two kernels and ten zero-filled tensors per view).
// Citations are relative to DGR/cuda_rasterizer/.
// =============================================================================
#include "gslm_common.cuh"
#include "gslm_math.cuh"

namespace gslm {

struct CamConst {  // copied to shared memory by every CTA
    float V[16], PV[16], campos[3];
    float Vt[16], PVt[16], campos_t[3];
};

struct CamF {
    const CamConst* c;
    float fx, fy, tfx, tfy;
    __device__ __forceinline__ float V(int k) const { return c->V[k]; }
    __device__ __forceinline__ float PV(int k) const { return c->PV[k]; }
    __device__ __forceinline__ float focal_x() const { return fx; }
    __device__ __forceinline__ float focal_y() const { return fy; }
    __device__ __forceinline__ float tan_fovx() const { return tfx; }
    __device__ __forceinline__ float tan_fovy() const { return tfy; }
};
template <bool CAMT> struct CamD {
    const CamConst* c;
    float fx, fy, tfx, tfy, fx_t, fy_t, tfx_t, tfy_t;
    __device__ __forceinline__ dualf V(int k) const { return dualf(c->V[k], CAMT ? c->Vt[k] : 0.f); }
    __device__ __forceinline__ dualf PV(int k) const { return dualf(c->PV[k], CAMT ? c->PVt[k] : 0.f); }
    __device__ __forceinline__ dualf focal_x() const { return dualf(fx, CAMT ? fx_t : 0.f); }
    __device__ __forceinline__ dualf focal_y() const { return dualf(fy, CAMT ? fy_t : 0.f); }
    __device__ __forceinline__ dualf tan_fovx() const { return dualf(tfx, CAMT ? tfx_t : 0.f); }
    __device__ __forceinline__ dualf tan_fovy() const { return dualf(tfy, CAMT ? tfy_t : 0.f); }
};

__device__ __forceinline__ void load_cam(CamConst* sc, const float* V, const float* PV, const float* campos,
                                         const float* Vt, const float* PVt, const float* campos_t) {
    int t = threadIdx.x;
    if (t < 16) {
        sc->V[t] = V[t];
        sc->PV[t] = PV[t];
        sc->Vt[t] = Vt ? Vt[t] : 0.f;
        sc->PVt[t] = PVt ? PVt[t] : 0.f;
    }
    if (t < 3) {
        sc->campos[t] = campos ? campos[t] : 0.f;
        sc->campos_t[t] = campos_t ? campos_t[t] : 0.f;
    }
    __syncthreads();
}

struct PreArgs {
    int P, D, M, W, H;
    float scale_modifier, tan_fovx, tan_fovy, focal_x, focal_y;
    int antialiasing, prefiltered;
    const float *means3D, *opacities, *scales, *rotations, *cov3D_precomp, *shs, *colors_precomp;
    const float *viewmatrix, *projmatrix, *campos;
};

__global__ void __launch_bounds__(256) k_preprocess(PreArgs a, GeomView gv, int32_t* __restrict__ radii) {
    __shared__ CamConst sc;
    load_cam(&sc, a.viewmatrix, a.projmatrix, a.campos, nullptr, nullptr, nullptr);
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= a.P) return;
    CamF cam{&sc, a.focal_x, a.focal_y, a.tan_fovx, a.tan_fovy};
    const float* mp = a.means3D + 3 * (size_t)idx;
    v3<float> mean = {__ldg(mp), __ldg(mp + 1), __ldg(mp + 2)};
    float sx = 0.f, sy = 0.f, sz = 0.f, rot[4] = {1.f, 0.f, 0.f, 0.f};
    const float* cov_pre = nullptr;
    float cov_pre_v[6];
    if (a.cov3D_precomp) {
#pragma unroll
        for (int k = 0; k < 6; k++) cov_pre_v[k] = __ldg(a.cov3D_precomp + 6 * (size_t)idx + k);
        cov_pre = cov_pre_v;
    } else {
        const float* sp = a.scales + 3 * (size_t)idx;
        sx = __ldg(sp); sy = __ldg(sp + 1); sz = __ldg(sp + 2);
        float4 q = __ldg(reinterpret_cast<const float4*>(a.rotations) + idx);
        rot[0] = q.x; rot[1] = q.y; rot[2] = q.z; rot[3] = q.w;
    }
    PreOut<float> o;
    bool vis = project_gaussian<float>(mean, sx, sy, sz, a.scale_modifier, rot, cov_pre, __ldg(a.opacities + idx), cam,
                                       a.W, a.H, a.antialiasing != 0, false, o);
    if (!vis) {
        if (a.prefiltered && !(sc.V[2] * mean.x + sc.V[6] * mean.y + sc.V[10] * mean.z + sc.V[14] > 0.2f))
            __trap();  // auxiliary.h:211-215: "Point is filtered although prefiltered is set"
        radii[idx] = 0;
        gv.tiles_touched[idx] = 0;
        return;
    }
    if (!a.colors_precomp) {
        const float* sh = a.shs + (size_t)idx * a.M * 3;
        uint8_t cl;
        v3<float> campos = {sc.campos[0], sc.campos[1], sc.campos[2]};
        v3<float> col = color_from_sh<float>(
            a.D, mean, campos, [&](int k) { return v3<float>{__ldg(sh + 3 * k), __ldg(sh + 3 * k + 1), __ldg(sh + 3 * k + 2)}; },
            cl);
        gv.rgb[3 * (size_t)idx + 0] = col.x;
        gv.rgb[3 * (size_t)idx + 1] = col.y;
        gv.rgb[3 * (size_t)idx + 2] = col.z;
        gv.clamped[idx] = cl;
    } else {
        gv.clamped[idx] = 0;
    }
    if (!a.cov3D_precomp) {
#pragma unroll
        for (int k = 0; k < 6; k++) gv.cov3D[6 * (size_t)idx + k] = o.cov3D[k];
    }
    gv.depths[idx] = o.depth;
    radii[idx] = o.radius;
    gv.means2D[idx] = make_float2(o.x, o.y);
    gv.conic_opacity[idx] = make_float4(o.conx, o.cony, o.conz, o.opac);
    gv.tiles_touched[idx] = o.tiles;
}

int launch_preprocess(const GslmGaussians& g, const GslmCamera& cam, const GeomView& gv, int32_t* radii,
                      cudaStream_t s) {
    PreArgs a;
    a.P = g.P; a.D = g.sh_degree; a.M = g.M; a.W = cam.width; a.H = cam.height;
    a.scale_modifier = g.scale_modifier; a.tan_fovx = cam.tan_fovx; a.tan_fovy = cam.tan_fovy;
    a.focal_y = cam.height / (2.0f * cam.tan_fovy);  // rasterizer_impl.cu:265-266
    a.focal_x = cam.width / (2.0f * cam.tan_fovx);
    a.antialiasing = cam.antialiasing; a.prefiltered = cam.prefiltered;
    a.means3D = g.means3D; a.opacities = g.opacities; a.scales = g.scales; a.rotations = g.rotations;
    a.cov3D_precomp = g.cov3D_precomp; a.shs = g.shs; a.colors_precomp = g.colors_precomp;
    a.viewmatrix = cam.viewmatrix; a.projmatrix = cam.projmatrix; a.campos = cam.campos;
    if (g.P == 0) return 0;
    StageScope ts(ST_PREPROCESS, s);
    k_preprocess<<<(g.P + 255) / 256, 256, 0, s>>>(a, gv, radii);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------
// tangent preprocess
// ---------------------------------------------------------------------------------------------
struct PreTanArgs {
    PreArgs p;
    TangentPtrs t;
    float focal_x_t, focal_y_t;
};

__device__ __forceinline__ dualf ldd(const float* p, const float* t, size_t i) {
    return dualf(__ldg(p + i), t ? __ldg(t + i) : 0.f);
}

template <bool CAMT>
__global__ void __launch_bounds__(256)
k_preprocess_tangent(PreTanArgs a, const int32_t* __restrict__ radii, const uint32_t* __restrict__ g_off,
                     float4* __restrict__ tgeom) {
    __shared__ CamConst sc;
    load_cam(&sc, a.p.viewmatrix, a.p.projmatrix, a.p.campos, a.t.viewmatrix, a.t.projmatrix, a.t.campos);
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= a.p.P) return;
    float4* out = tgeom + (size_t)REC_F4 * idx;
    // only Gaussians with at least one compact instance are ever looked up by the tile sweeps
    if (radii[idx] <= 0 || g_off[idx] == g_off[idx + 1]) return;
    CamD<CAMT> cam{&sc, a.p.focal_x, a.p.focal_y, a.p.tan_fovx, a.p.tan_fovy,
                   a.focal_x_t, a.focal_y_t, a.t.tan_fovx, a.t.tan_fovy};
    v3<dualf> mean = {ldd(a.p.means3D, a.t.means3D, 3 * (size_t)idx), ldd(a.p.means3D, a.t.means3D, 3 * (size_t)idx + 1),
                      ldd(a.p.means3D, a.t.means3D, 3 * (size_t)idx + 2)};
    dualf sx(0.f), sy(0.f), sz(0.f), rot[4] = {dualf(1.f), dualf(0.f), dualf(0.f), dualf(0.f)};
    dualf cov_pre_v[6];
    const dualf* cov_pre = nullptr;
    if (a.p.cov3D_precomp) {
#pragma unroll
        for (int k = 0; k < 6; k++) cov_pre_v[k] = ldd(a.p.cov3D_precomp, a.t.cov3D_precomp, 6 * (size_t)idx + k);
        cov_pre = cov_pre_v;
    } else {
        sx = ldd(a.p.scales, a.t.scales, 3 * (size_t)idx);
        sy = ldd(a.p.scales, a.t.scales, 3 * (size_t)idx + 1);
        sz = ldd(a.p.scales, a.t.scales, 3 * (size_t)idx + 2);
#pragma unroll
        for (int k = 0; k < 4; k++) rot[k] = ldd(a.p.rotations, a.t.rotations, 4 * (size_t)idx + k);
    }
    dualf mod(a.p.scale_modifier, CAMT ? a.t.scale_modifier : 0.f);
    PreOut<dualf> o;
    project_gaussian<dualf>(mean, sx, sy, sz, mod, rot, cov_pre, ldd(a.p.opacities, a.t.opacities, idx), cam, a.p.W,
                            a.p.H, a.p.antialiasing != 0, true, o);
    float cr_t, cg_t, cb_t;
    if (!a.p.colors_precomp) {
        const float* sh = a.p.shs + (size_t)idx * a.p.M * 3;
        const float* sht = a.t.shs ? a.t.shs + (size_t)idx * a.p.M * 3 : nullptr;
        uint8_t cl;
        v3<dualf> campos = {dualf(sc.campos[0], CAMT ? sc.campos_t[0] : 0.f), dualf(sc.campos[1], CAMT ? sc.campos_t[1] : 0.f),
                            dualf(sc.campos[2], CAMT ? sc.campos_t[2] : 0.f)};
        v3<dualf> col = color_from_sh<dualf>(
            a.p.D, mean, campos,
            [&](int k) { return v3<dualf>{ldd(sh, sht, 3 * k), ldd(sh, sht, 3 * k + 1), ldd(sh, sht, 3 * k + 2)}; }, cl);
        cr_t = col.x.d; cg_t = col.y.d; cb_t = col.z.d;
    } else {
        const float* ct = a.t.colors_precomp;
        cr_t = ct ? __ldg(ct + 3 * (size_t)idx) : 0.f;
        cg_t = ct ? __ldg(ct + 3 * (size_t)idx + 1) : 0.f;
        cb_t = ct ? __ldg(ct + 3 * (size_t)idx + 2) : 0.f;
    }
    dualf invd = 1.f / o.depth;   // forward_impl.h:506: (1 / depth) on duals
    // tangent of the exponent, power = -0.5 (A dx^2 + C dy^2) - B dx dy with dx = x - px, as a polynomial in
    // (dx, dy):  power_t = dx (a3 dx + a5 dy + a1) + dy (a4 dy + a2)   (forward_impl.h:468-476 on duals)
    const float a1 = -(o.conx.v * o.x.d + o.cony.v * o.y.d);
    const float a2 = -(o.conz.v * o.y.d + o.cony.v * o.x.d);
    out[0] = make_float4(a1, a2, -0.5f * o.conx.d, -0.5f * o.conz.d);
    // the tile sweeps evaluate alpha_raw = exp2(... + log2 o): d alpha_raw = alpha_raw (power_t + o_t / o)
    out[1] = make_float4(-o.cony.d, o.opac.v > 0.f ? o.opac.d / o.opac.v : 0.f, cr_t, cg_t);
    out[2] = make_float4(cb_t, invd.d, 0.f, 0.f);
}

int launch_preprocess_tangent(const GslmGaussians& g, const TangentPtrs& t, const GslmCamera& cam,
                              const int32_t* radii, const uint32_t* g_off, float4* tgeom, cudaStream_t s) {
    PreTanArgs a;
    a.p.P = g.P; a.p.D = g.sh_degree; a.p.M = g.M; a.p.W = cam.width; a.p.H = cam.height;
    a.p.scale_modifier = g.scale_modifier; a.p.tan_fovx = cam.tan_fovx; a.p.tan_fovy = cam.tan_fovy;
    a.p.focal_y = cam.height / (2.0f * cam.tan_fovy);
    a.p.focal_x = cam.width / (2.0f * cam.tan_fovx);
    a.p.antialiasing = cam.antialiasing; a.p.prefiltered = cam.prefiltered;
    a.p.means3D = g.means3D; a.p.opacities = g.opacities; a.p.scales = g.scales; a.p.rotations = g.rotations;
    a.p.cov3D_precomp = g.cov3D_precomp; a.p.shs = g.shs; a.p.colors_precomp = g.colors_precomp;
    a.p.viewmatrix = cam.viewmatrix; a.p.projmatrix = cam.projmatrix; a.p.campos = cam.campos;
    a.t = t;
    // focal = S / (2 tan)  =>  d focal = -focal / tan * d tan   (rasterizer_jvp_impl.h:182-183 on duals)
    a.focal_x_t = -a.p.focal_x / cam.tan_fovx * t.tan_fovx;
    a.focal_y_t = -a.p.focal_y / cam.tan_fovy * t.tan_fovy;
    if (g.P == 0) return 0;
    StageScope ts(ST_PRE_TANGENT, s);
    bool camt = t.viewmatrix || t.projmatrix || t.campos || t.tan_fovx != 0.f || t.tan_fovy != 0.f ||
                t.scale_modifier != 0.f;
    if (camt) k_preprocess_tangent<true><<<(g.P + 255) / 256, 256, 0, s>>>(a, radii, g_off, tgeom);
    else k_preprocess_tangent<false><<<(g.P + 255) / 256, 256, 0, s>>>(a, radii, g_off, tgeom);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

// rasterizer_impl.cu:54-66 + auxiliary.h:194-219 (prefiltered = false)
__global__ void k_mark_visible(int P, const float* __restrict__ means3D, const float* __restrict__ V,
                               uint8_t* __restrict__ present) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float* p = means3D + 3 * (size_t)idx;
    float z = V[2] * p[0] + V[6] * p[1] + V[10] * p[2] + V[14];
    present[idx] = (z <= 0.2f) ? 0 : 1;
}
int launch_mark_visible(int P, const float* means3D, const float* viewmatrix, uint8_t* present, cudaStream_t s) {
    if (P == 0) return 0;
    k_mark_visible<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, viewmatrix, present);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

}  // namespace gslm
