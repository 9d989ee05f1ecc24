// =============================================================================
// gslm_math.cuh -- per-Gaussian projection math, written once for S = float (primal
// kernel) and S = dualf (forward-mode tangent kernel).
//
// What it computes is fixed by the reference (DGR/cuda_rasterizer/forward.cu:20-273 and
// its dual-number twin forward_impl.h:18-367); the expression trees of the S = float
// instantiation keep the reference's operand order so that nvcc's FMA contraction
// produces the same bits for radius / rect / depth (the inputs of the bit-exact
// tile|depth sort).  Dual semantics follow DGR/cuda_rasterizer/helper_math.h:70-80:
// comparisons look at the primal, min/max return the selected operand with its tangent.
// =============================================================================
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#define GS_HD __host__ __device__ __forceinline__

namespace gslm {

struct dualf {
    float v, d;
    GS_HD dualf() {}
    GS_HD dualf(float v_) : v(v_), d(0.f) {}
    GS_HD dualf(float v_, float d_) : v(v_), d(d_) {}
};
GS_HD float prim(float x) { return x; }
GS_HD float prim(const dualf& x) { return x.v; }
GS_HD float tang(float) { return 0.f; }
GS_HD float tang(const dualf& x) { return x.d; }

GS_HD dualf operator+(dualf a, dualf b) { return dualf(a.v + b.v, a.d + b.d); }
GS_HD dualf operator-(dualf a, dualf b) { return dualf(a.v - b.v, a.d - b.d); }
GS_HD dualf operator*(dualf a, dualf b) { return dualf(a.v * b.v, a.d * b.v + a.v * b.d); }
GS_HD dualf operator/(dualf a, dualf b) {
    float q = a.v / b.v;
    return dualf(q, (a.d - q * b.d) / b.v);
}
GS_HD dualf operator-(dualf a) { return dualf(-a.v, -a.d); }
GS_HD dualf operator+(dualf a, float b) { return dualf(a.v + b, a.d); }
GS_HD dualf operator+(float a, dualf b) { return dualf(a + b.v, b.d); }
GS_HD dualf operator-(dualf a, float b) { return dualf(a.v - b, a.d); }
GS_HD dualf operator-(float a, dualf b) { return dualf(a - b.v, -b.d); }
GS_HD dualf operator*(dualf a, float b) { return dualf(a.v * b, a.d * b); }
GS_HD dualf operator*(float a, dualf b) { return dualf(a * b.v, a * b.d); }
GS_HD dualf operator/(dualf a, float b) { return dualf(a.v / b, a.d / b); }
GS_HD dualf operator/(float a, dualf b) {
    float q = a / b.v;
    return dualf(q, -q * b.d / b.v);
}

GS_HD float s_sqrt(float x) { return sqrtf(x); }
GS_HD dualf s_sqrt(dualf x) {
    float r = sqrtf(x.v);
    return dualf(r, x.d / (2.f * r));
}
GS_HD float s_min(float a, float b) { return fminf(a, b); }
GS_HD float s_max(float a, float b) { return fmaxf(a, b); }
GS_HD dualf s_min(dualf a, dualf b) { return a.v < b.v ? a : b; }
GS_HD dualf s_max(dualf a, dualf b) { return a.v > b.v ? a : b; }

template <class S> struct v3 { S x, y, z; };
template <class S> GS_HD v3<S> operator+(v3<S> a, v3<S> b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
template <class S> GS_HD v3<S> operator-(v3<S> a, v3<S> b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
template <class S> GS_HD v3<S> operator*(S s, v3<S> a) { return {s * a.x, s * a.y, s * a.z}; }

// column-major 3x3 with glm's product (third_party/glm/glm/detail/type_mat3x3.inl:486-519)
template <class S> struct m3 { S m[3][3]; };
template <class S> GS_HD m3<S> mul(const m3<S>& A, const m3<S>& B) {
    m3<S> R;
#pragma unroll
    for (int c = 0; c < 3; c++)
#pragma unroll
        for (int r = 0; r < 3; r++)
            R.m[c][r] = A.m[0][r] * B.m[c][0] + A.m[1][r] * B.m[c][1] + A.m[2][r] * B.m[c][2];
    return R;
}
template <class S> GS_HD m3<S> transpose(const m3<S>& A) {
    m3<S> R;
#pragma unroll
    for (int c = 0; c < 3; c++)
#pragma unroll
        for (int r = 0; r < 3; r++) R.m[c][r] = A.m[r][c];
    return R;
}

// DGR/cuda_rasterizer/auxiliary.h:30-47
#define GS_SH_C0 0.28209479177387814f
#define GS_SH_C1 0.4886025119029199f
#define GS_SH_C2_0 1.0925484305920792f
#define GS_SH_C2_1 -1.0925484305920792f
#define GS_SH_C2_2 0.31539156525252005f
#define GS_SH_C2_3 -1.0925484305920792f
#define GS_SH_C2_4 0.5462742152960396f
#define GS_SH_C3_0 -0.5900435899266435f
#define GS_SH_C3_1 2.890611442640554f
#define GS_SH_C3_2 -0.4570457994644658f
#define GS_SH_C3_3 0.3731763325901154f
#define GS_SH_C3_4 -0.4570457994644658f
#define GS_SH_C3_5 1.445305721320277f
#define GS_SH_C3_6 -0.5900435899266435f

// auxiliary.h:49-52: ((v + 1.0) * S - 1.0) * 0.5 is evaluated in double.
GS_HD float ndc2pix(float v, int S) { return (float)((((double)v + 1.0) * S - 1.0) * 0.5); }
GS_HD dualf ndc2pix(dualf v, int S) { return dualf(ndc2pix(v.v, S), v.d * (float)S * 0.5f); }

// auxiliary.h:60-70 (int max_radius overload)
GS_HD void tile_rect(float px, float py, int max_radius, int gx, int gy, int& x0, int& y0, int& x1, int& y1) {
    x0 = min(gx, max(0, (int)((px - max_radius) / 16)));
    y0 = min(gy, max(0, (int)((py - max_radius) / 16)));
    x1 = min(gx, max(0, (int)((px + max_radius + 16 - 1) / 16)));
    y1 = min(gy, max(0, (int)((py + max_radius + 16 - 1) / 16)));
}

// forward.cu:117-151
template <class S> GS_HD void cov3d_from_scale_rot(S sx, S sy, S sz, S mod, S r, S x, S y, S z, S* cov3D) {
    m3<S> Sm, R;
    const S zero = S(0.f);
    Sm.m[0][0] = mod * sx; Sm.m[0][1] = zero; Sm.m[0][2] = zero;
    Sm.m[1][0] = zero; Sm.m[1][1] = mod * sy; Sm.m[1][2] = zero;
    Sm.m[2][0] = zero; Sm.m[2][1] = zero; Sm.m[2][2] = mod * sz;
    R.m[0][0] = 1.f - 2.f * (y * y + z * z); R.m[0][1] = 2.f * (x * y - r * z); R.m[0][2] = 2.f * (x * z + r * y);
    R.m[1][0] = 2.f * (x * y + r * z); R.m[1][1] = 1.f - 2.f * (x * x + z * z); R.m[1][2] = 2.f * (y * z - r * x);
    R.m[2][0] = 2.f * (x * z - r * y); R.m[2][1] = 2.f * (y * z + r * x); R.m[2][2] = 1.f - 2.f * (x * x + y * y);
    m3<S> M = mul(Sm, R);
    m3<S> Sigma = mul(transpose(M), M);
    cov3D[0] = Sigma.m[0][0]; cov3D[1] = Sigma.m[0][1]; cov3D[2] = Sigma.m[0][2];
    cov3D[3] = Sigma.m[1][1]; cov3D[4] = Sigma.m[1][2]; cov3D[5] = Sigma.m[2][2];
}

// forward.cu:77-112; V is the 16-float view matrix as S
template <class S, class CamT>
GS_HD void cov2d_from_cov3d(v3<S> mean, S focal_x, S focal_y, S tan_fovx, S tan_fovy, const S* cov3D, const CamT& cam,
                            S& a, S& b, S& c) {
    v3<S> t;
    t.x = cam.V(0) * mean.x + cam.V(4) * mean.y + cam.V(8) * mean.z + cam.V(12);
    t.y = cam.V(1) * mean.x + cam.V(5) * mean.y + cam.V(9) * mean.z + cam.V(13);
    t.z = cam.V(2) * mean.x + cam.V(6) * mean.y + cam.V(10) * mean.z + cam.V(14);
    const S limx = 1.3f * tan_fovx;
    const S limy = 1.3f * tan_fovy;
    const S txtz = t.x / t.z;
    const S tytz = t.y / t.z;
    t.x = s_min(limx, s_max(-limx, txtz)) * t.z;
    t.y = s_min(limy, s_max(-limy, tytz)) * t.z;
    const S zero = S(0.f);
    m3<S> J, W, Vrk;
    J.m[0][0] = focal_x / t.z; J.m[0][1] = zero; J.m[0][2] = -(focal_x * t.x) / (t.z * t.z);
    J.m[1][0] = zero; J.m[1][1] = focal_y / t.z; J.m[1][2] = -(focal_y * t.y) / (t.z * t.z);
    J.m[2][0] = zero; J.m[2][1] = zero; J.m[2][2] = zero;
    W.m[0][0] = cam.V(0); W.m[0][1] = cam.V(4); W.m[0][2] = cam.V(8);
    W.m[1][0] = cam.V(1); W.m[1][1] = cam.V(5); W.m[1][2] = cam.V(9);
    W.m[2][0] = cam.V(2); W.m[2][1] = cam.V(6); W.m[2][2] = cam.V(10);
    m3<S> T = mul(W, J);
    Vrk.m[0][0] = cov3D[0]; Vrk.m[0][1] = cov3D[1]; Vrk.m[0][2] = cov3D[2];
    Vrk.m[1][0] = cov3D[1]; Vrk.m[1][1] = cov3D[3]; Vrk.m[1][2] = cov3D[4];
    Vrk.m[2][0] = cov3D[2]; Vrk.m[2][1] = cov3D[4]; Vrk.m[2][2] = cov3D[5];
    m3<S> cov = mul(mul(transpose(T), transpose(Vrk)), T);
    a = cov.m[0][0];
    b = cov.m[0][1];
    c = cov.m[1][1];
}

// forward.cu:20-74 (vec3 indexing of sh; the reference's dual twin mis-indexes sh for
// degree > 0 -- SURVEY F7 -- the tangent here is the true derivative of forward.cu).
// SHF(k) returns the k-th coefficient as v3<S>.
template <class S, class SHF>
GS_HD v3<S> color_from_sh(int deg, v3<S> pos, v3<S> campos, SHF sh, uint8_t& clamped_bits) {
    v3<S> dir = pos - campos;
    S len = s_sqrt(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
    dir.x = dir.x / len; dir.y = dir.y / len; dir.z = dir.z / len;
    v3<S> result = S(GS_SH_C0) * sh(0);
    if (deg > 0) {
        S x = dir.x, y = dir.y, z = dir.z;
        result = result - (GS_SH_C1 * y) * sh(1) + (GS_SH_C1 * z) * sh(2) - (GS_SH_C1 * x) * sh(3);
        if (deg > 1) {
            S xx = x * x, yy = y * y, zz = z * z;
            S xy = x * y, yz = y * z, xz = x * z;
            result = result + (GS_SH_C2_0 * xy) * sh(4) + (GS_SH_C2_1 * yz) * sh(5) +
                     (GS_SH_C2_2 * (2.0f * zz - xx - yy)) * sh(6) + (GS_SH_C2_3 * xz) * sh(7) +
                     (GS_SH_C2_4 * (xx - yy)) * sh(8);
            if (deg > 2) {
                result = result + (GS_SH_C3_0 * y * (3.0f * xx - yy)) * sh(9) + (GS_SH_C3_1 * xy * z) * sh(10) +
                         (GS_SH_C3_2 * y * (4.0f * zz - xx - yy)) * sh(11) +
                         (GS_SH_C3_3 * z * (2.0f * zz - 3.0f * xx - 3.0f * yy)) * sh(12) +
                         (GS_SH_C3_4 * x * (4.0f * zz - xx - yy)) * sh(13) + (GS_SH_C3_5 * z * (xx - yy)) * sh(14) +
                         (GS_SH_C3_6 * x * (xx - 3.0f * yy)) * sh(15);
            }
        }
    }
    result.x = result.x + 0.5f; result.y = result.y + 0.5f; result.z = result.z + 0.5f;
    clamped_bits = (uint8_t)((prim(result.x) < 0.f ? 1 : 0) | (prim(result.y) < 0.f ? 2 : 0) |
                             (prim(result.z) < 0.f ? 4 : 0));
    const S zero = S(0.f);
    return {s_max(result.x, zero), s_max(result.y, zero), s_max(result.z, zero)};
}

template <class S> struct PreOut {
    int radius;
    uint32_t tiles;
    S x, y, depth, conx, cony, conz, opac;
    S cov3D[6];
};

// forward.cu:154-273 up to (not including) the colour; returns false when culled.
// When `force_visible` (cached tangent path) the culls are skipped: visibility was
// decided by the primal pass.
template <class S, class CamT>
GS_HD bool project_gaussian(v3<S> mean, S sx, S sy, S sz, S mod, const S* rot, const S* cov3D_pre, S opacity,
                            const CamT& cam, int W, int H, bool antialiasing, bool force_visible, PreOut<S>& o) {
    o.radius = 0;
    o.tiles = 0;
    // in_frustum (auxiliary.h:194-219): only the p_view.z <= 0.2 test culls
    S pvz = cam.V(2) * mean.x + cam.V(6) * mean.y + cam.V(10) * mean.z + cam.V(14);
    if (!force_visible && prim(pvz) <= 0.2f) return false;
    S hx = cam.PV(0) * mean.x + cam.PV(4) * mean.y + cam.PV(8) * mean.z + cam.PV(12);
    S hy = cam.PV(1) * mean.x + cam.PV(5) * mean.y + cam.PV(9) * mean.z + cam.PV(13);
    S hw = cam.PV(3) * mean.x + cam.PV(7) * mean.y + cam.PV(11) * mean.z + cam.PV(15);
    S p_w = 1.0f / (hw + 0.0000001f);
    S projx = hx * p_w, projy = hy * p_w;
    if (cov3D_pre) {
#pragma unroll
        for (int k = 0; k < 6; k++) o.cov3D[k] = cov3D_pre[k];
    } else {
        cov3d_from_scale_rot(sx, sy, sz, mod, rot[0], rot[1], rot[2], rot[3], o.cov3D);
    }
    S ca, cb, cc;
    cov2d_from_cov3d(mean, cam.focal_x(), cam.focal_y(), cam.tan_fovx(), cam.tan_fovy(), o.cov3D, cam, ca, cb, cc);
    const float h_var = 0.3f;
    const S det_cov = ca * cc - cb * cb;
    ca = ca + h_var;
    cc = cc + h_var;
    const S det = ca * cc - cb * cb;
    S h_scale = S(1.0f);
    if (antialiasing) h_scale = s_sqrt(s_max(S(0.000025f), det_cov / det));
    if (!force_visible && prim(det) == 0.0f) return false;
    S det_inv = 1.f / det;
    o.conx = cc * det_inv;
    o.cony = -cb * det_inv;
    o.conz = ca * det_inv;
    S mid = 0.5f * (ca + cc);
    S lambda1 = mid + s_sqrt(s_max(S(0.1f), mid * mid - det));
    S lambda2 = mid - s_sqrt(s_max(S(0.1f), mid * mid - det));
    float my_radius = ceilf(3.f * sqrtf(fmaxf(prim(lambda1), prim(lambda2))));
    o.x = ndc2pix(projx, W);
    o.y = ndc2pix(projy, H);
    int gx = (W + 15) / 16, gy = (H + 15) / 16;
    int x0, y0, x1, y1;
    tile_rect(prim(o.x), prim(o.y), (int)my_radius, gx, gy, x0, y0, x1, y1);
    if (!force_visible && (x1 - x0) * (y1 - y0) == 0) return false;
    o.depth = pvz;
    o.radius = (int)my_radius;
    o.opac = opacity * h_scale;
    o.tiles = (uint32_t)((y1 - y0) * (x1 - x0));
    return true;
}

}  // namespace gslm
