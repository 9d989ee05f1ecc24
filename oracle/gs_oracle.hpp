// =============================================================================
// oracle/gs_oracle.hpp -- CPU restatement of the reference rasterizer arithmetic.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under gslm_b200/ (the product) may include,
// link or call this.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs use it, as the checker / CPU baseline.
//
// Parity status: PINNED for preprocess primal, preprocess JVP (camera tangents)
// and binning by the reference's own golden vectors
// (submodules/diff-gaussian-rasterization/tests/{data,render_data}; see
// tests/golden/make_golden.py).  Rendered images, gradients and Gaussian-
// parameter tangents have NO reference golden vector ("parity unpinned" for
// those: self-consistency only -- finite differences, adjoint identity, float64
// instantiation of this same restatement, and on the GPU box the reference's
// own primal/VJP kernels built by oracle/Makefile into oracle/_ref/).
//
// All citations are file:line under
//   /root/reference/submodules/diff-gaussian-rasterization/cuda_rasterizer/ .
//
// One restatement, four instantiations: the scalar type S is float, double,
// Dual<float> or Dual<double>.  Dual<T> is a forward-mode dual number with the
// semantics the reference's (missing) float_grad.h is used with:
//   * every comparison / branch / ceil / integer conversion looks at the primal
//     only (forward_impl.h:285,325,337-342,477-492);
//   * fminf / fmaxf return the *selected operand* including its tangent; a
//     selected constant has zero tangent (helper_math.h:70-80).
// =============================================================================
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <algorithm>
#include <type_traits>

namespace gso {

// ----------------------------------------------------------------------------
// Dual numbers
// ----------------------------------------------------------------------------
template <class T> struct Dual {
    T v, d;
    Dual() : v(0), d(0) {}
    Dual(T v_) : v(v_), d(0) {}
    Dual(T v_, T d_) : v(v_), d(d_) {}
};
template <class T> struct is_dual : std::false_type {};
template <class T> struct is_dual<Dual<T>> : std::true_type {};
template <class S> struct base { using type = S; };
template <class T> struct base<Dual<T>> { using type = T; };
template <class S> using base_t = typename base<S>::type;

inline float  prim(float x) { return x; }
inline double prim(double x) { return x; }
template <class T> inline T prim(const Dual<T>& x) { return x.v; }
inline float  tang(float) { return 0.f; }
inline double tang(double) { return 0.0; }
template <class T> inline T tang(const Dual<T>& x) { return x.d; }

template <class T> inline Dual<T> operator+(Dual<T> a, Dual<T> b) { return {a.v + b.v, a.d + b.d}; }
template <class T> inline Dual<T> operator-(Dual<T> a, Dual<T> b) { return {a.v - b.v, a.d - b.d}; }
template <class T> inline Dual<T> operator*(Dual<T> a, Dual<T> b) { return {a.v * b.v, a.d * b.v + a.v * b.d}; }
template <class T> inline Dual<T> operator/(Dual<T> a, Dual<T> b) {
    T q = a.v / b.v;
    return {q, (a.d - q * b.d) / b.v};
}
template <class T> inline Dual<T> operator-(Dual<T> a) { return {-a.v, -a.d}; }
template <class T> inline Dual<T> operator+(Dual<T> a, T b) { return {a.v + b, a.d}; }
template <class T> inline Dual<T> operator+(T a, Dual<T> b) { return {a + b.v, b.d}; }
template <class T> inline Dual<T> operator-(Dual<T> a, T b) { return {a.v - b, a.d}; }
template <class T> inline Dual<T> operator-(T a, Dual<T> b) { return {a - b.v, -b.d}; }
template <class T> inline Dual<T> operator*(Dual<T> a, T b) { return {a.v * b, a.d * b}; }
template <class T> inline Dual<T> operator*(T a, Dual<T> b) { return {a * b.v, a * b.d}; }
template <class T> inline Dual<T> operator/(Dual<T> a, T b) { return {a.v / b, a.d / b}; }
template <class T> inline Dual<T> operator/(T a, Dual<T> b) {
    T q = a / b.v;
    return {q, -q * b.d / b.v};
}
template <class T> inline Dual<T>& operator+=(Dual<T>& a, Dual<T> b) { a = a + b; return a; }

inline float  s_sqrt(float x) { return std::sqrt(x); }
inline double s_sqrt(double x) { return std::sqrt(x); }
template <class T> inline Dual<T> s_sqrt(Dual<T> x) {
    T r = std::sqrt(x.v);
    return {r, x.d / (T(2) * r)};
}
inline float  s_exp(float x) { return std::exp(x); }
inline double s_exp(double x) { return std::exp(x); }
template <class T> inline Dual<T> s_exp(Dual<T> x) {
    T e = std::exp(x.v);
    return {e, e * x.d};
}
// helper_math.h:70-80: "a < b ? a : b" / "a > b ? a : b" on the primal,
// returning the selected operand (and its tangent).
template <class S> inline S s_min(S a, S b) { return prim(a) < prim(b) ? a : b; }
template <class S> inline S s_max(S a, S b) { return prim(a) > prim(b) ? a : b; }

// ----------------------------------------------------------------------------
// Tiny glm-compatible vec3 / mat3 (column-major m[c][r], glm product order)
// ----------------------------------------------------------------------------
template <class S> struct V3 { S x, y, z; };
template <class S> inline V3<S> operator+(V3<S> a, V3<S> b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
template <class S> inline V3<S> operator-(V3<S> a, V3<S> b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
template <class S> inline V3<S> operator*(S s, V3<S> a) { return {s * a.x, s * a.y, s * a.z}; }
template <class S> inline V3<S> operator/(V3<S> a, S s) { return {a.x / s, a.y / s, a.z / s}; }
template <class S> inline S dot3(V3<S> a, V3<S> b) {
    // glm::dot(vec3): tmp = a*b; tmp.x + tmp.y + tmp.z
    return a.x * b.x + a.y * b.y + a.z * b.z;
}

template <class S> struct M3 {
    S m[3][3];  // m[col][row]
    // glm::mat3(x0,y0,z0, x1,y1,z1, x2,y2,z2): columns in order
    static M3 cols(S a, S b, S c, S d, S e, S f, S g, S h, S i) {
        M3 r;
        r.m[0][0] = a; r.m[0][1] = b; r.m[0][2] = c;
        r.m[1][0] = d; r.m[1][1] = e; r.m[1][2] = f;
        r.m[2][0] = g; r.m[2][1] = h; r.m[2][2] = i;
        return r;
    }
};
template <class S> inline M3<S> transpose(const M3<S>& a) {
    M3<S> r;
    for (int c = 0; c < 3; c++) for (int q = 0; q < 3; q++) r.m[c][q] = a.m[q][c];
    return r;
}
// glm operator*(mat3, mat3): Result[c][r] = A[0][r]*B[c][0] + A[1][r]*B[c][1] + A[2][r]*B[c][2]
template <class S> inline M3<S> operator*(const M3<S>& A, const M3<S>& B) {
    M3<S> R;
    for (int c = 0; c < 3; c++)
        for (int r = 0; r < 3; r++)
            R.m[c][r] = A.m[0][r] * B.m[c][0] + A.m[1][r] * B.m[c][1] + A.m[2][r] * B.m[c][2];
    return R;
}

// ----------------------------------------------------------------------------
// Loading S from (data, tangent) pointer pairs
// ----------------------------------------------------------------------------
template <class S> struct Ld {
    using T = base_t<S>;
    static S get(const T* p, const T* /*t*/, size_t i) { return p[i]; }
    static S cst(T v, T /*t*/) { return v; }
};
template <class T> struct Ld<Dual<T>> {
    static Dual<T> get(const T* p, const T* t, size_t i) { return Dual<T>(p[i], t ? t[i] : T(0)); }
    static Dual<T> cst(T v, T t) { return Dual<T>(v, t); }
};

// SH constants, auxiliary.h:30-47 (float literals; promoted exactly for double)
static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                              -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                              0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                              -0.5900435899266435f};

constexpr int BLOCK_X = 16, BLOCK_Y = 16;  // config.h:17-19

// auxiliary.h:60-70 getRect (int max_radius overload); integer result.
template <class T>
inline void get_rect(T px, T py, int max_radius, int gx, int gy, int& x0, int& y0, int& x1, int& y1) {
    x0 = std::min(gx, std::max(0, (int)((px - max_radius) / BLOCK_X)));
    y0 = std::min(gy, std::max(0, (int)((py - max_radius) / BLOCK_Y)));
    x1 = std::min(gx, std::max(0, (int)((px + max_radius + BLOCK_X - 1) / BLOCK_X)));
    y1 = std::min(gy, std::max(0, (int)((py + max_radius + BLOCK_Y - 1) / BLOCK_Y)));
}

// auxiliary.h:49-52: ndc2Pix evaluates ((v + 1.0) * S - 1.0) * 0.5 in DOUBLE and
// the result is narrowed to float on store (forward.cu:244).
inline float  ndc2pix(float v, int S) { return (float)((((double)v + 1.0) * S - 1.0) * 0.5); }
inline double ndc2pix(double v, int S) { return ((v + 1.0) * S - 1.0) * 0.5; }
template <class T> inline Dual<T> ndc2pix(Dual<T> v, int S) {
    return Dual<T>(ndc2pix(v.v, S), v.d * T(S) * T(0.5));
}

template <class S> struct Geom {
    int radius = 0;
    int tiles = 0;
    S xy[2];
    S depth;
    S cov3D[6];
    S rgb[3];
    S conic_opacity[4];
    bool clamped[3] = {false, false, false};
};

// forward.cu:117-151 computeCov3D
template <class S>
inline void compute_cov3D(V3<S> scale, S mod, const S* q, S* cov3D) {
    using T = base_t<S>;
    S zero = S(T(0)), one = S(T(1));
    M3<S> Sm = M3<S>::cols(mod * scale.x, zero, zero, zero, mod * scale.y, zero, zero, zero, mod * scale.z);
    S r = q[0], x = q[1], y = q[2], z = q[3];
    T two = T(2);
    M3<S> R = M3<S>::cols(
        one - two * (y * y + z * z), two * (x * y - r * z), two * (x * z + r * y),
        two * (x * y + r * z), one - two * (x * x + z * z), two * (y * z - r * x),
        two * (x * z - r * y), two * (y * z + r * x), one - two * (x * x + y * y));
    M3<S> M = Sm * R;
    M3<S> Sigma = transpose(M) * M;
    cov3D[0] = Sigma.m[0][0];
    cov3D[1] = Sigma.m[0][1];
    cov3D[2] = Sigma.m[0][2];
    cov3D[3] = Sigma.m[1][1];
    cov3D[4] = Sigma.m[1][2];
    cov3D[5] = Sigma.m[2][2];
}

// auxiliary.h:85-93 / 110-117
template <class S> inline V3<S> transform4x3(V3<S> p, const S* m) {
    return {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
            m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
            m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]};
}

// forward.cu:77-112 computeCov2D
template <class S>
inline void compute_cov2D(V3<S> mean, S focal_x, S focal_y, S tan_fovx, S tan_fovy, const S* cov3D,
                          const S* V, S& a, S& b, S& c) {
    using T = base_t<S>;
    V3<S> t = transform4x3(mean, V);
    S limx = T(1.3f) * tan_fovx;
    S limy = T(1.3f) * tan_fovy;
    S txtz = t.x / t.z;
    S tytz = t.y / t.z;
    t.x = s_min(limx, s_max(-limx, txtz)) * t.z;
    t.y = s_min(limy, s_max(-limy, tytz)) * t.z;
    S zero = S(T(0));
    M3<S> J = M3<S>::cols(focal_x / t.z, zero, -(focal_x * t.x) / (t.z * t.z),
                          zero, focal_y / t.z, -(focal_y * t.y) / (t.z * t.z),
                          zero, zero, zero);
    M3<S> W = M3<S>::cols(V[0], V[4], V[8], V[1], V[5], V[9], V[2], V[6], V[10]);
    M3<S> Tm = W * J;
    M3<S> Vrk = M3<S>::cols(cov3D[0], cov3D[1], cov3D[2], cov3D[1], cov3D[3], cov3D[4], cov3D[2], cov3D[4], cov3D[5]);
    M3<S> cov = transpose(Tm) * transpose(Vrk) * Tm;
    a = cov.m[0][0];
    b = cov.m[0][1];
    c = cov.m[1][1];
}

// forward.cu:20-74 computeColorFromSH (vec3 form; SURVEY F7: the reference JVP
// indexes sh as floats for degree>0, which is a bug -- we follow forward.cu).
template <class S>
inline V3<S> color_from_sh(int deg, V3<S> pos, V3<S> campos, const V3<S>* sh, bool* clamped) {
    using T = base_t<S>;
    V3<S> dir = pos - campos;
    S len = s_sqrt(dot3(dir, dir));
    dir = dir / len;
    V3<S> result = S(T(SH_C0)) * sh[0];
    if (deg > 0) {
        S x = dir.x, y = dir.y, z = dir.z;
        result = result - (T(SH_C1) * y) * sh[1] + (T(SH_C1) * z) * sh[2] - (T(SH_C1) * x) * sh[3];
        if (deg > 1) {
            S xx = x * x, yy = y * y, zz = z * z;
            S xy = x * y, yz = y * z, xz = x * z;
            result = result + (T(SH_C2[0]) * xy) * sh[4] + (T(SH_C2[1]) * yz) * sh[5] +
                     (T(SH_C2[2]) * (T(2.0f) * zz - xx - yy)) * sh[6] + (T(SH_C2[3]) * xz) * sh[7] +
                     (T(SH_C2[4]) * (xx - yy)) * sh[8];
            if (deg > 2) {
                result = result + (T(SH_C3[0]) * y * (T(3.0f) * xx - yy)) * sh[9] +
                         (T(SH_C3[1]) * xy * z) * sh[10] +
                         (T(SH_C3[2]) * y * (T(4.0f) * zz - xx - yy)) * sh[11] +
                         (T(SH_C3[3]) * z * (T(2.0f) * zz - T(3.0f) * xx - T(3.0f) * yy)) * sh[12] +
                         (T(SH_C3[4]) * x * (T(4.0f) * zz - xx - yy)) * sh[13] +
                         (T(SH_C3[5]) * z * (xx - yy)) * sh[14] +
                         (T(SH_C3[6]) * x * (xx - T(3.0f) * yy)) * sh[15];
            }
        }
    }
    S half = S(T(0.5f));
    result = result + V3<S>{half, half, half};
    clamped[0] = prim(result.x) < 0;
    clamped[1] = prim(result.y) < 0;
    clamped[2] = prim(result.z) < 0;
    S zero = S(T(0));
    return {s_max(result.x, zero), s_max(result.y, zero), s_max(result.z, zero)};
}

struct PreCfg {
    int P, D, M, W, H;
    bool antialiasing, has_cov3D_precomp, has_colors_precomp;
};

// forward.cu:154-273 preprocessCUDA (and forward_impl.h:208-367 for S = Dual).
// Returns false if the Gaussian is culled (radius/tiles stay 0).
template <class S>
inline bool preprocess_one(const PreCfg& c, V3<S> mean, V3<S> scale, S mod, const S* rot, S opacity,
                           const V3<S>* sh, const S* cov3D_pre, const S* color_pre, const S* V, const S* PV,
                           V3<S> campos, S tan_fovx, S tan_fovy, S focal_x, S focal_y, Geom<S>& g) {
    using T = base_t<S>;
    g.radius = 0;
    g.tiles = 0;
    // in_frustum, auxiliary.h:194-219
    V3<S> p_view = transform4x3(mean, V);
    if (prim(p_view.z) <= T(0.2f)) return false;
    S hx = PV[0] * mean.x + PV[4] * mean.y + PV[8] * mean.z + PV[12];
    S hy = PV[1] * mean.x + PV[5] * mean.y + PV[9] * mean.z + PV[13];
    S hw = PV[3] * mean.x + PV[7] * mean.y + PV[11] * mean.z + PV[15];
    S p_w = T(1.0f) / (hw + T(0.0000001f));
    S projx = hx * p_w, projy = hy * p_w;

    if (c.has_cov3D_precomp) {
        for (int k = 0; k < 6; k++) g.cov3D[k] = cov3D_pre[k];
    } else {
        compute_cov3D(scale, mod, rot, g.cov3D);
    }
    S ca, cb, cc;
    compute_cov2D(mean, focal_x, focal_y, tan_fovx, tan_fovy, g.cov3D, V, ca, cb, cc);

    const T h_var = T(0.3f);
    S det_cov = ca * cc - cb * cb;
    ca = ca + h_var;
    cc = cc + h_var;
    S det_cov_plus_h = ca * cc - cb * cb;
    S h_scale = S(T(1.0f));
    if (c.antialiasing) h_scale = s_sqrt(s_max(S(T(0.000025f)), det_cov / det_cov_plus_h));
    S det = det_cov_plus_h;
    if (prim(det) == T(0)) return false;
    S det_inv = T(1.f) / det;
    S conx = cc * det_inv, cony = -cb * det_inv, conz = ca * det_inv;

    S mid = T(0.5f) * (ca + cc);
    S lam1 = mid + s_sqrt(s_max(S(T(0.1f)), mid * mid - det));
    S lam2 = mid - s_sqrt(s_max(S(T(0.1f)), mid * mid - det));
    T my_radius = std::ceil(T(3.f) * std::sqrt(std::max(prim(lam1), prim(lam2))));
    S px = ndc2pix(projx, c.W), py = ndc2pix(projy, c.H);
    int gx = (c.W + BLOCK_X - 1) / BLOCK_X, gy = (c.H + BLOCK_Y - 1) / BLOCK_Y;
    int x0, y0, x1, y1;
    get_rect<T>(prim(px), prim(py), (int)my_radius, gx, gy, x0, y0, x1, y1);
    if ((x1 - x0) * (y1 - y0) == 0) return false;

    if (!c.has_colors_precomp) {
        V3<S> col = color_from_sh(c.D, mean, campos, sh, g.clamped);
        g.rgb[0] = col.x; g.rgb[1] = col.y; g.rgb[2] = col.z;
    } else {
        g.rgb[0] = color_pre[0]; g.rgb[1] = color_pre[1]; g.rgb[2] = color_pre[2];
    }
    g.depth = p_view.z;
    g.radius = (int)my_radius;
    g.xy[0] = px; g.xy[1] = py;
    g.conic_opacity[0] = conx; g.conic_opacity[1] = cony; g.conic_opacity[2] = conz;
    g.conic_opacity[3] = opacity * h_scale;
    g.tiles = (y1 - y0) * (x1 - x0);
    return true;
}

// forward.cu:278-402 renderCUDA for one pixel (and forward_impl.h:373-539 for Dual).
// point list entries [r0, r1) index the per-Gaussian arrays.
template <class S, class GetXY, class GetCO, class GetRGB, class GetDepth>
inline void blend_pixel(int px, int py, uint32_t r0, uint32_t r1, const uint32_t* point_list, GetXY getxy,
                        GetCO getco, GetRGB getrgb, GetDepth getdepth, const S* bg, S* outC, S& outD,
                        S& final_T, uint32_t& n_contrib) {
    using T = base_t<S>;
    S Tr = S(T(1.0f));
    S C[3] = {S(T(0)), S(T(0)), S(T(0))};
    S D = S(T(0));
    uint32_t contributor = 0, last = 0;
    T pfx = (T)px, pfy = (T)py;
    for (uint32_t k = r0; k < r1; k++) {
        contributor++;
        uint32_t id = point_list[k];
        S xy[2]; getxy(id, xy);
        S dx = xy[0] - pfx, dy = xy[1] - pfy;
        S co[4]; getco(id, co);
        S power = T(-0.5f) * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
        if (prim(power) > T(0)) continue;
        S alpha = s_min(S(T(0.99f)), co[3] * s_exp(power));
        if (prim(alpha) < T(1.0f / 255.0f)) continue;
        S test_T = Tr * (T(1) - alpha);
        if (prim(test_T) < T(0.0001f)) break;  // done = true
        S rgb[3]; getrgb(id, rgb);
        for (int ch = 0; ch < 3; ch++) C[ch] = C[ch] + rgb[ch] * alpha * Tr;
        S depth = getdepth(id);
        D = D + (T(1) / depth) * alpha * Tr;
        Tr = test_T;
        last = contributor;
    }
    final_T = Tr;
    n_contrib = last;
    for (int ch = 0; ch < 3; ch++) outC[ch] = C[ch] + Tr * bg[ch];
    outD = D;
}

}  // namespace gso
