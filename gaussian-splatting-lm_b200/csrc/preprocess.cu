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
#include <cstring>

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
        // scalar fetches on purpose: the image is compared bit for bit with the reference's kernels, and the float4
        // fetch used by the tangent / adjoint kernels (sh_elem4) changes how nvcc contracts this expression
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

// One launch covers up to MAX_BATCH_VIEWS linearised views (blockIdx.y) and, per view, only the Gaussians that own
// compact instances there (active list built at compaction): dense warps instead of ~3 live lanes in 32.
struct PreTanView {
    const float *viewmatrix, *projmatrix, *campos;
    const uint32_t* active_ids;
    const uint32_t* n_active;
    float4* tgeom;
    float tan_fovx, tan_fovy, focal_x, focal_y, focal_x_t, focal_y_t;
    int W, H;
};
struct PreTanBatch { PreTanView v[MAX_BATCH_VIEWS]; };

#ifndef GSLM_TAN_MINB
#define GSLM_TAN_MINB 3
#endif
template <bool CAMT>
__global__ void __launch_bounds__(256, GSLM_TAN_MINB)
k_preprocess_tangent(PreTanArgs a, const __grid_constant__ PreTanBatch vb) {
    const PreTanView& vw = vb.v[blockIdx.y];
    const uint32_t n_active = *vw.n_active;
    if (blockIdx.x * blockDim.x >= n_active) return;
    __shared__ CamConst sc;
    load_cam(&sc, vw.viewmatrix, vw.projmatrix, vw.campos, a.t.viewmatrix, a.t.projmatrix, a.t.campos);
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= n_active) return;
    const int idx = (int)__ldg(vw.active_ids + slot);
    float4* out = vw.tgeom + (size_t)REC_F4 * idx;
    a.p.W = vw.W; a.p.H = vw.H; a.p.focal_x = vw.focal_x; a.p.focal_y = vw.focal_y;
    a.p.tan_fovx = vw.tan_fovx; a.p.tan_fovy = vw.tan_fovy; a.focal_x_t = vw.focal_x_t; a.focal_y_t = vw.focal_y_t;
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
        auto dd4 = [&](int f) { return dualf(sh_elem4(sh, f), sht ? sh_elem4(sht, f) : 0.f); };
        v3<dualf> col = (sh_rows_aligned(a.p.shs, a.p.M) && (!sht || sh_rows_aligned(a.t.shs, a.p.M)))
            ? color_from_sh<dualf>(a.p.D, mean, campos,
                  [&](int k) { return v3<dualf>{dd4(3 * k), dd4(3 * k + 1), dd4(3 * k + 2)}; }, cl)
            : color_from_sh<dualf>(a.p.D, mean, campos,
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

int launch_preprocess_tangent_batch(const GslmGaussians& g, const TangentPtrs& t, int n_views, const GslmCamera* cams,
                                    const GeomView* gvs, float4* const* tgeoms, cudaStream_t s) {
    if (g.P == 0 || n_views == 0) return 0;
    if (n_views > MAX_BATCH_VIEWS) { set_error("too many views in one batch"); return 1; }
    PreTanArgs a;
    memset(&a, 0, sizeof(a));
    a.p.P = g.P; a.p.D = g.sh_degree; a.p.M = g.M; a.p.scale_modifier = g.scale_modifier;
    a.p.antialiasing = cams[0].antialiasing; a.p.prefiltered = cams[0].prefiltered;
    a.p.means3D = g.means3D; a.p.opacities = g.opacities; a.p.scales = g.scales; a.p.rotations = g.rotations;
    a.p.cov3D_precomp = g.cov3D_precomp; a.p.shs = g.shs; a.p.colors_precomp = g.colors_precomp;
    a.t = t;
    PreTanBatch vb;
    memset(&vb, 0, sizeof(vb));
    for (int i = 0; i < n_views; i++) {
        const GslmCamera& cam = cams[i];
        if (cam.antialiasing != cams[0].antialiasing) { set_error("views of one batch must agree on antialiasing"); return 1; }
        PreTanView& v = vb.v[i];
        v.viewmatrix = cam.viewmatrix; v.projmatrix = cam.projmatrix; v.campos = cam.campos;
        v.active_ids = gvs[i].active_ids; v.n_active = gvs[i].n_active; v.tgeom = tgeoms[i];
        v.tan_fovx = cam.tan_fovx; v.tan_fovy = cam.tan_fovy; v.W = cam.width; v.H = cam.height;
        v.focal_y = cam.height / (2.0f * cam.tan_fovy);
        v.focal_x = cam.width / (2.0f * cam.tan_fovx);
        // focal = S / (2 tan)  =>  d focal = -focal / tan * d tan   (rasterizer_jvp_impl.h:182-183 on duals)
        v.focal_x_t = -v.focal_x / cam.tan_fovx * t.tan_fovx;
        v.focal_y_t = -v.focal_y / cam.tan_fovy * t.tan_fovy;
    }
    StageScope ts(ST_PRE_TANGENT, s);
    const bool camt = t.viewmatrix || t.projmatrix || t.campos || t.tan_fovx != 0.f || t.tan_fovy != 0.f ||
                      t.scale_modifier != 0.f;
    const dim3 grid((g.P + 255) / 256, n_views);
    if (camt) k_preprocess_tangent<true><<<grid, 256, 0, s>>>(a, vb);
    else k_preprocess_tangent<false><<<grid, 256, 0, s>>>(a, vb);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

int launch_preprocess_tangent(const GslmGaussians& g, const TangentPtrs& t, const GslmCamera& cam, const GeomView& gv,
                              float4* tgeom, cudaStream_t s) {
    return launch_preprocess_tangent_batch(g, t, 1, &cam, &gv, &tgeom, s);
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
