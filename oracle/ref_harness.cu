// =============================================================================
// oracle/ref_harness.cu -- TEST INFRASTRUCTURE ONLY (GPU-side parity oracle + "kernel to beat").
//
// A small C ABI around the REFERENCE's own, unmodified CUDA rasterizer
// (CudaRasterizer::Rasterizer::{forward,backward,markVisible},
//  submodules/diff-gaussian-rasterization/cuda_rasterizer/rasterizer.h:26-121), compiled from
// the reference tree by oracle/Makefile (target `ref`) into oracle/_ref/libgsref.so.
// This file contains no rasterizer arithmetic: it allocates the three scratch buffers the
// reference asks for through its std::function allocators, forwards the arguments, and exposes
// the integer state (point_list, tile ranges, n_contrib, ...) for the bit-exactness tests.
// Only tests/ and bench.py's reference-kernel leg load it; the product never does.
// All pointers are DEVICE pointers unless noted.  Kernels run on the legacy default stream,
// like the reference.
// =============================================================================
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <functional>
#include <stdexcept>
#include <string>
#include "rasterizer.h"
#include "rasterizer_impl.h"

namespace {
struct Buf {
    char* p = nullptr;
    size_t cap = 0;
    char* get(size_t n) {
        if (n > cap) {
            if (p) cudaFree(p);
            cap = n + n / 4 + 4096;
            if (cudaMalloc(&p, cap) != cudaSuccess) { p = nullptr; cap = 0; }
        }
        return p;
    }
};
Buf g_geom, g_bin, g_img;
int g_P = 0, g_W = 0, g_H = 0, g_R = 0;
std::string g_err;
}  // namespace

extern "C" {

const char* gsref_last_error() { return g_err.c_str(); }

// returns num_rendered (>= 0) or -1 on error
int gsref_forward(int P, int D, int M, const float* background, int W, int H, const float* means3D,
                  const float* shs, const float* colors_precomp, const float* opacities, const float* scales,
                  float scale_modifier, const float* rotations, const float* cov3D_precomp,
                  const float* viewmatrix, const float* projmatrix, const float* campos, float tan_fovx,
                  float tan_fovy, int prefiltered, int antialiasing, float* out_color, float* out_invdepth,
                  int* radii, int debug) {
    try {
        std::function<char*(size_t)> ga = [](size_t n) { return g_geom.get(n); };
        std::function<char*(size_t)> ba = [](size_t n) { return g_bin.get(n); };
        std::function<char*(size_t)> ia = [](size_t n) { return g_img.get(n); };
        g_P = P; g_W = W; g_H = H;
        g_R = CudaRasterizer::Rasterizer::forward(ga, ba, ia, P, D, M, background, W, H, means3D, shs, colors_precomp,
                                                  opacities, scales, scale_modifier, rotations, cov3D_precomp,
                                                  viewmatrix, projmatrix, campos, tan_fovx, tan_fovy,
                                                  prefiltered != 0, out_color, out_invdepth, antialiasing != 0, radii,
                                                  debug != 0);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { g_err = cudaGetErrorString(e); return -1; }
        return g_R;
    } catch (const std::exception& ex) {
        g_err = ex.what();
        return -1;
    }
}

// gradients must be ZERO-FILLED by the caller (the reference accumulates with atomicAdd,
// rasterize_points.cu:339-348).  dL_dconic is [P,4] (float4 with .x,.y,.w used), dL_dinvdepth [P].
int gsref_backward(int P, int D, int M, const float* background, int W, int H, const float* means3D, const float* shs,
                   const float* colors_precomp, const float* opacities, const float* scales, float scale_modifier,
                   const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                   const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy, const int* radii,
                   const float* dL_dpix, const float* dL_dinvdepths, float* dL_dmean2D, float* dL_dconic,
                   float* dL_dopacity, float* dL_dcolor, float* dL_dinvdepth, float* dL_dmean3D, float* dL_dcov3D,
                   float* dL_dsh, float* dL_dscale, float* dL_drot, int antialiasing, int debug) {
    try {
        if (P != g_P || W != g_W || H != g_H) { g_err = "backward does not match the last forward"; return -1; }
        CudaRasterizer::Rasterizer::backward(P, D, M, g_R, background, W, H, means3D, shs, colors_precomp, opacities,
                                             scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix,
                                             campos, tan_fovx, tan_fovy, radii, g_geom.p, g_bin.p, g_img.p, dL_dpix,
                                             dL_dinvdepths, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor,
                                             dL_dinvdepth, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot,
                                             antialiasing != 0, debug != 0);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { g_err = cudaGetErrorString(e); return -1; }
        return 0;
    } catch (const std::exception& ex) {
        g_err = ex.what();
        return -1;
    }
}

// Copies the integer / per-pixel state of the last forward call into caller-provided DEVICE arrays
// (any may be NULL): point_list [R] u32, ranges [T,2] u32 (first T entries of the reference's
// per-pixel array), n_contrib [H*W] u32, final_T [H*W] f32, tiles_touched [P] u32,
// point_offsets [P] u32, depths [P] f32, means2D [P,2] f32.
int gsref_state(uint32_t* point_list, uint32_t* ranges, uint32_t* n_contrib, float* final_T, uint32_t* tiles_touched,
                uint32_t* point_offsets, float* depths, float* means2D) {
    using namespace CudaRasterizer;
    char* gp = g_geom.p;
    GeometryState geom = GeometryState::fromChunk(gp, g_P);
    char* ip = g_img.p;
    ImageState img = ImageState::fromChunk(ip, (size_t)g_W * g_H);
    char* bp = g_bin.p;
    ReducedBinningState bin = ReducedBinningState::fromChunk(bp, g_R);
    const size_t N = (size_t)g_W * g_H, T = (size_t)((g_W + 15) / 16) * ((g_H + 15) / 16);
    auto cp = [](void* dst, const void* src, size_t n) {
        return !dst || n == 0 || cudaMemcpy(dst, src, n, cudaMemcpyDeviceToDevice) == cudaSuccess;
    };
    bool ok = cp(point_list, bin.point_list, 4 * (size_t)g_R) && cp(ranges, img.ranges, 8 * T) &&
              cp(n_contrib, img.n_contrib, 4 * N) && cp(final_T, img.accum_alpha, 4 * N) &&
              cp(tiles_touched, geom.tiles_touched, 4 * (size_t)g_P) &&
              cp(point_offsets, geom.point_offsets, 4 * (size_t)g_P) && cp(depths, geom.depths, 4 * (size_t)g_P) &&
              cp(means2D, geom.means2D, 8 * (size_t)g_P);
    if (!ok) { g_err = "state copy failed"; return -1; }
    return 0;
}

#ifdef GSREF_WITH_JVP
// ---- forward mode: the reference's own JVP kernels (preprocessCUDAJvp, renderCUDAJvpSplit) driven through
// CudaRasterizer::Rasterizer::forwardJvp exactly as RasterizeGaussiansJvpCUDA does (rasterize_points.cu:244-297):
// every differentiable input travels as a FloatGradArray<float> / FloatGrad<float> (value, tangent).  Only built when
// the functional float_grad.h stand-in (oracle/ref_shim) is in use.  All tangent pointers must be valid device
// arrays of the primal's size (zeros for "no tangent"); colors_precomp / cov3D_precomp may be NULL (both halves).
int gsref_forward_jvp(int P, int D, int M, const float* background, const float* background_t, int W, int H,
                      const float* means3D, const float* means3D_t, const float* shs, const float* shs_t,
                      const float* colors_precomp, const float* colors_precomp_t, const float* opacities,
                      const float* opacities_t, const float* scales, const float* scales_t, float scale_modifier,
                      float scale_modifier_t, const float* rotations, const float* rotations_t,
                      const float* cov3D_precomp, const float* cov3D_precomp_t, const float* viewmatrix,
                      const float* viewmatrix_t, const float* projmatrix, const float* projmatrix_t,
                      const float* campos, const float* campos_t, float tan_fovx, float tan_fovx_t, float tan_fovy,
                      float tan_fovy_t, int prefiltered, int antialiasing, float* out_color, float* out_color_t,
                      float* out_invdepth, float* out_invdepth_t, int* radii, int debug) {
    try {
        std::function<char*(size_t)> ga = [](size_t n) { return g_geom.get(n); };
        std::function<char*(size_t)> ba = [](size_t n) { return g_bin.get(n); };
        std::function<char*(size_t)> ia = [](size_t n) { return g_img.get(n); };
        g_P = P; g_W = W; g_H = H;
        auto A = [](const float* d, const float* t) { return FloatGradArray<float>(const_cast<float*>(d), const_cast<float*>(t)); };
        FloatGradArray<float> bg = A(background, background_t), m3 = A(means3D, means3D_t), sh = A(shs, shs_t),
                              col = A(colors_precomp, colors_precomp_t), op = A(opacities, opacities_t),
                              sc = A(scales, scales_t), rot = A(rotations, rotations_t),
                              cov = A(cov3D_precomp, cov3D_precomp_t), vm = A(viewmatrix, viewmatrix_t),
                              pm = A(projmatrix, projmatrix_t), cp = A(campos, campos_t),
                              oc = A(out_color, out_color_t), od = A(out_invdepth, out_invdepth_t);
        FloatGrad<float> smod(scale_modifier, scale_modifier_t), tfx(tan_fovx, tan_fovx_t), tfy(tan_fovy, tan_fovy_t);
        const bool pre = prefiltered != 0, aa = antialiasing != 0, dbg = debug != 0;
        g_R = CudaRasterizer::Rasterizer::forwardJvp(ga, ba, ia, P, D, M, bg, W, H, m3, sh, col, op, sc, smod, rot, cov, vm,
                                                     pm, cp, tfx, tfy, pre, oc, od, aa, radii, dbg);
        cudaError_t e = cudaDeviceSynchronize();
        if (e == cudaSuccess) e = cudaGetLastError();
        if (e != cudaSuccess) { g_err = cudaGetErrorString(e); return -1; }
        return g_R;
    } catch (const std::exception& ex) {
        g_err = ex.what();
        return -1;
    }
}

// Per-Gaussian tangent records of the last gsref_forward_jvp call (the tail of GeometryStateJvp,
// rasterizer_impl.cu:172-194): depths_t [P], means2D_t [P,2], cov3D_t [P,6], conic_opacity_t [P,4], rgb_t [P,3];
// plus the primal conic_opacity [P,4] and rgb [P,3].  Any pointer may be NULL.
int gsref_jvp_state(float* depths_t, float* means2D_t, float* cov3D_t, float* conic_opacity_t, float* rgb_t,
                    float* conic_opacity, float* rgb) {
    using namespace CudaRasterizer;
    char* gp = g_geom.p;
    GeometryStateJvp geom = GeometryStateJvp::fromChunk(gp, g_P);
    const size_t P = (size_t)g_P;
    auto cp = [](void* dst, const void* src, size_t n) {
        return !dst || n == 0 || cudaMemcpy(dst, src, n, cudaMemcpyDeviceToDevice) == cudaSuccess;
    };
    bool ok = cp(depths_t, geom.depths.grad_ptr(), 4 * P) && cp(means2D_t, geom.means2D.grad_ptr(), 8 * P) &&
              cp(cov3D_t, geom.cov3D.grad_ptr(), 24 * P) && cp(conic_opacity_t, geom.conic_opacity.grad_ptr(), 16 * P) &&
              cp(rgb_t, geom.rgb.grad_ptr(), 12 * P) && cp(conic_opacity, geom.conic_opacity.data_ptr(), 16 * P) &&
              cp(rgb, geom.rgb.data_ptr(), 12 * P);
    if (!ok) { g_err = "jvp state copy failed"; return -1; }
    return 0;
}
int gsref_has_jvp() { return 1; }
#else
int gsref_has_jvp() { return 0; }
#endif

void gsref_mark_visible(int P, float* means3D, float* viewmatrix, float* projmatrix, bool* present) {
    CudaRasterizer::Rasterizer::markVisible(P, means3D, viewmatrix, projmatrix, present);
}

void gsref_release() {
    for (Buf* b : {&g_geom, &g_bin, &g_img}) {
        if (b->p) cudaFree(b->p);
        b->p = nullptr; b->cap = 0;
    }
}

}  // extern "C"
