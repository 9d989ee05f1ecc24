// =============================================================================
// oracle/gs_oracle.cpp -- C entry points (ctypes) of the CPU oracle.
//
// TEST INFRASTRUCTURE ONLY (see gs_oracle.hpp header).  Restates, stage by stage,
//   preprocess fwd / JVP   forward.cu:154-273, forward_impl.h:208-367
//   binning                rasterizer_impl.cu:35-138,321-364
//   blend fwd / JVP        forward.cu:278-402, forward_impl.h:373-539
//   blend bwd              backward.cu:452-638
//   preprocess bwd         backward.cu:23-142,147-326,330-449
// Each stage exists as *_f32 (the reference's arithmetic type) and *_f64 (same
// formulas in double: derivative ground truth for small cases).
// Build: oracle/Makefile  ->  oracle/libgs_oracle.so
// =============================================================================
#include "gs_oracle.hpp"
#include <vector>
#include <numeric>
#include <cstdio>
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace gso;

extern "C" int gso_num_threads() {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
extern "C" void gso_set_num_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}

// ----------------------------------------------------------------------------
// preprocess (primal or JVP)
// ----------------------------------------------------------------------------
template <class T> struct PreArgs {
    int P, D, M, W, H, antialiasing, jvp;
    const T *means3D, *scales, *rotations, *opacities, *shs, *cov3D_precomp, *colors_precomp;
    const T *viewmatrix, *projmatrix, *campos;
    T scale_modifier, tan_fovx, tan_fovy;
    // tangents (any may be null => zero)
    const T *t_means3D, *t_scales, *t_rotations, *t_opacities, *t_shs, *t_cov3D_precomp, *t_colors_precomp;
    const T *t_viewmatrix, *t_projmatrix, *t_campos;
    T t_scale_modifier, t_tan_fovx, t_tan_fovy;
    // outputs
    int32_t* radii;
    uint32_t* tiles_touched;
    T *xy, *depths, *cov3D, *rgb, *conic_opacity;
    uint8_t* clamped;
    T *t_xy, *t_depths, *t_cov3D, *t_rgb, *t_conic_opacity;
};

template <class S, class T> static void store(T* p, T* tp, size_t i, const S& s) {
    p[i] = prim(s);
    if (tp) tp[i] = tang(s);
}

template <class S, class T> static void preprocess_impl(const PreArgs<T>& a) {
    PreCfg c;
    c.P = a.P; c.D = a.D; c.M = a.M; c.W = a.W; c.H = a.H;
    c.antialiasing = a.antialiasing != 0;
    c.has_cov3D_precomp = a.cov3D_precomp != nullptr;
    c.has_colors_precomp = a.colors_precomp != nullptr;
    using L = Ld<S>;
    S V[16], PV[16];
    for (int k = 0; k < 16; k++) {
        V[k] = L::get(a.viewmatrix, a.t_viewmatrix, k);
        PV[k] = L::get(a.projmatrix, a.t_projmatrix, k);
    }
    V3<S> campos = {L::get(a.campos, a.t_campos, 0), L::get(a.campos, a.t_campos, 1), L::get(a.campos, a.t_campos, 2)};
    S tfx = L::cst(a.tan_fovx, a.t_tan_fovx), tfy = L::cst(a.tan_fovy, a.t_tan_fovy);
    S mod = L::cst(a.scale_modifier, a.t_scale_modifier);
    // rasterizer_impl.cu:265-266 / rasterizer_jvp_impl.h:182-183
    S focal_y = T(a.H) / (T(2.0f) * tfy);
    S focal_x = T(a.W) / (T(2.0f) * tfx);

#pragma omp parallel for schedule(static)
    for (int i = 0; i < a.P; i++) {
        Geom<S> g;
        V3<S> mean = {L::get(a.means3D, a.t_means3D, 3 * (size_t)i), L::get(a.means3D, a.t_means3D, 3 * (size_t)i + 1),
                      L::get(a.means3D, a.t_means3D, 3 * (size_t)i + 2)};
        V3<S> scale = {S(T(0)), S(T(0)), S(T(0))};
        S rot[4] = {S(T(1)), S(T(0)), S(T(0)), S(T(0))};
        if (!c.has_cov3D_precomp) {
            scale = {L::get(a.scales, a.t_scales, 3 * (size_t)i), L::get(a.scales, a.t_scales, 3 * (size_t)i + 1),
                     L::get(a.scales, a.t_scales, 3 * (size_t)i + 2)};
            for (int k = 0; k < 4; k++) rot[k] = L::get(a.rotations, a.t_rotations, 4 * (size_t)i + k);
        }
        S opacity = L::get(a.opacities, a.t_opacities, i);
        V3<S> sh[16];
        S cov_pre[6], col_pre[3];
        if (!c.has_colors_precomp) {
            for (int k = 0; k < a.M && k < 16; k++) {
                size_t o = ((size_t)i * a.M + k) * 3;
                sh[k] = {L::get(a.shs, a.t_shs, o), L::get(a.shs, a.t_shs, o + 1), L::get(a.shs, a.t_shs, o + 2)};
            }
        } else {
            for (int k = 0; k < 3; k++) col_pre[k] = L::get(a.colors_precomp, a.t_colors_precomp, 3 * (size_t)i + k);
        }
        if (c.has_cov3D_precomp)
            for (int k = 0; k < 6; k++) cov_pre[k] = L::get(a.cov3D_precomp, a.t_cov3D_precomp, 6 * (size_t)i + k);
        bool vis = preprocess_one<S>(c, mean, scale, mod, rot, opacity, sh, cov_pre, col_pre, V, PV, campos, tfx, tfy,
                                     focal_x, focal_y, g);
        a.radii[i] = g.radius;
        a.tiles_touched[i] = (uint32_t)g.tiles;
        if (!vis) {
            // the reference leaves these uninitialised; the oracle writes zeros
            for (int k = 0; k < 2; k++) { a.xy[2 * (size_t)i + k] = 0; if (a.t_xy) a.t_xy[2 * (size_t)i + k] = 0; }
            a.depths[i] = 0; if (a.t_depths) a.t_depths[i] = 0;
            for (int k = 0; k < 6; k++) { a.cov3D[6 * (size_t)i + k] = 0; if (a.t_cov3D) a.t_cov3D[6 * (size_t)i + k] = 0; }
            for (int k = 0; k < 3; k++) { a.rgb[3 * (size_t)i + k] = 0; if (a.t_rgb) a.t_rgb[3 * (size_t)i + k] = 0; a.clamped[3 * (size_t)i + k] = 0; }
            for (int k = 0; k < 4; k++) { a.conic_opacity[4 * (size_t)i + k] = 0; if (a.t_conic_opacity) a.t_conic_opacity[4 * (size_t)i + k] = 0; }
            continue;
        }
        for (int k = 0; k < 2; k++) store(a.xy, a.t_xy, 2 * (size_t)i + k, g.xy[k]);
        store(a.depths, a.t_depths, i, g.depth);
        for (int k = 0; k < 6; k++) store(a.cov3D, a.t_cov3D, 6 * (size_t)i + k, g.cov3D[k]);
        for (int k = 0; k < 3; k++) {
            store(a.rgb, a.t_rgb, 3 * (size_t)i + k, g.rgb[k]);
            a.clamped[3 * (size_t)i + k] = g.clamped[k];
        }
        for (int k = 0; k < 4; k++) store(a.conic_opacity, a.t_conic_opacity, 4 * (size_t)i + k, g.conic_opacity[k]);
    }
}

extern "C" void gso_preprocess_f32(const PreArgs<float>* a) {
    if (a->jvp) preprocess_impl<Dual<float>, float>(*a); else preprocess_impl<float, float>(*a);
}
extern "C" void gso_preprocess_f64(const PreArgs<double>* a) {
    if (a->jvp) preprocess_impl<Dual<double>, double>(*a); else preprocess_impl<double, double>(*a);
}

// auxiliary.h:194-219 in_frustum via rasterizer_impl.cu:54-66 checkFrustum (markVisible)
extern "C" void gso_mark_visible_f32(int P, const float* means3D, const float* V, uint8_t* present) {
#pragma omp parallel for
    for (int i = 0; i < P; i++) {
        const float* p = means3D + 3 * (size_t)i;
        float z = V[2] * p[0] + V[6] * p[1] + V[10] * p[2] + V[14];
        present[i] = !(z <= 0.2f);
    }
}

// ----------------------------------------------------------------------------
// binning: rasterizer_impl.cu:35-50 (getHigherMsb), :70-111 (duplicateWithKeys),
// :116-138 (identifyTileRanges), :321-364 (scan, sort bits [0, 32+bit)).
// ----------------------------------------------------------------------------
extern "C" uint32_t gso_higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4;
    uint32_t step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

extern "C" uint32_t gso_scan(int P, const uint32_t* tiles_touched, uint32_t* point_offsets) {
    uint32_t s = 0;
    for (int i = 0; i < P; i++) { s += tiles_touched[i]; point_offsets[i] = s; }
    return P > 0 ? s : 0;
}

// keys_out / keys_unsorted_out may be null.  ranges is [T][2] and is zero-filled here
// (rasterizer_impl.cu:356).
extern "C" void gso_binning(int P, int W, int H, const float* xy, const float* depths, const int32_t* radii,
                            const uint32_t* point_offsets, uint32_t R, uint64_t* keys_out, uint32_t* point_list,
                            uint32_t* ranges) {
    int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
    std::vector<uint64_t> keys(R);
    std::vector<uint32_t> vals(R);
#pragma omp parallel for schedule(dynamic, 1024)
    for (int i = 0; i < P; i++) {
        if (radii[i] > 0) {
            uint32_t off = (i == 0) ? 0 : point_offsets[i - 1];
            int x0, y0, x1, y1;
            get_rect<float>(xy[2 * (size_t)i], xy[2 * (size_t)i + 1], radii[i], gx, gy, x0, y0, x1, y1);
            uint32_t dbits;
            std::memcpy(&dbits, &depths[i], 4);
            for (int y = y0; y < y1; y++)
                for (int x = x0; x < x1; x++) {
                    uint64_t key = (uint64_t)(y * gx + x);
                    key <<= 32;
                    key |= dbits;
                    keys[off] = key;
                    vals[off] = (uint32_t)i;
                    off++;
                }
        }
    }
    // Stable sort on bits [0, 32+bit): cub::DeviceRadixSort is a stable LSD sort, so
    // equal keys keep emission order (ascending Gaussian id within a tile).
    uint32_t bit = gso_higher_msb((uint32_t)(gx * gy));
    uint64_t mask = (32 + bit >= 64) ? ~0ull : ((1ull << (32 + bit)) - 1);
    std::vector<uint32_t> order(R);
    std::iota(order.begin(), order.end(), 0u);
    std::stable_sort(order.begin(), order.end(),
                     [&](uint32_t a, uint32_t b) { return (keys[a] & mask) < (keys[b] & mask); });
    for (uint32_t k = 0; k < R; k++) {
        point_list[k] = vals[order[k]];
        if (keys_out) keys_out[k] = keys[order[k]];
    }
    std::memset(ranges, 0, sizeof(uint32_t) * 2 * (size_t)gx * gy);
    for (uint32_t k = 0; k < R; k++) {
        uint32_t cur = (uint32_t)(keys[order[k]] >> 32);
        if (k == 0) ranges[2 * cur] = 0;
        else {
            uint32_t prev = (uint32_t)(keys[order[k - 1]] >> 32);
            if (cur != prev) { ranges[2 * prev + 1] = k; ranges[2 * cur] = k; }
        }
        if (k == R - 1) ranges[2 * cur + 1] = R;
    }
}

// ----------------------------------------------------------------------------
// blend forward (primal or JVP)
// ----------------------------------------------------------------------------
template <class T> struct BlendArgs {
    int W, H, jvp;
    const uint32_t* ranges;      // [T][2]
    const uint32_t* point_list;  // [R]
    const T *xy, *rgb, *conic_opacity, *depths, *bg;
    const T *t_xy, *t_rgb, *t_conic_opacity, *t_depths, *t_bg;
    T *out_color, *out_invdepth, *final_T;
    uint32_t* n_contrib;
    T *t_out_color, *t_out_invdepth, *t_final_T;
};

template <class S, class T> static void blend_impl(const BlendArgs<T>& a) {
    using L = Ld<S>;
    int gx = (a.W + BLOCK_X - 1) / BLOCK_X, gy = (a.H + BLOCK_Y - 1) / BLOCK_Y;
    size_t HW = (size_t)a.H * a.W;
    S bg[3] = {L::get(a.bg, a.t_bg, 0), L::get(a.bg, a.t_bg, 1), L::get(a.bg, a.t_bg, 2)};
#pragma omp parallel for schedule(dynamic, 1)
    for (int t = 0; t < gx * gy; t++) {
        int tx = t % gx, ty = t / gx;
        uint32_t r0 = a.ranges[2 * t], r1 = a.ranges[2 * t + 1];
        for (int ly = 0; ly < BLOCK_Y; ly++)
            for (int lx = 0; lx < BLOCK_X; lx++) {
                int px = tx * BLOCK_X + lx, py = ty * BLOCK_Y + ly;
                if (px >= a.W || py >= a.H) continue;
                S C[3], D, fT;
                uint32_t nc;
                blend_pixel<S>(
                    px, py, r0, r1, a.point_list,
                    [&](uint32_t id, S* o) { o[0] = L::get(a.xy, a.t_xy, 2 * (size_t)id); o[1] = L::get(a.xy, a.t_xy, 2 * (size_t)id + 1); },
                    [&](uint32_t id, S* o) { for (int k = 0; k < 4; k++) o[k] = L::get(a.conic_opacity, a.t_conic_opacity, 4 * (size_t)id + k); },
                    [&](uint32_t id, S* o) { for (int k = 0; k < 3; k++) o[k] = L::get(a.rgb, a.t_rgb, 3 * (size_t)id + k); },
                    [&](uint32_t id) { return L::get(a.depths, a.t_depths, id); }, bg, C, D, fT, nc);
                size_t pix = (size_t)py * a.W + px;
                for (int ch = 0; ch < 3; ch++) store(a.out_color, a.t_out_color, ch * HW + pix, C[ch]);
                store(a.out_invdepth, a.t_out_invdepth, pix, D);
                store(a.final_T, a.t_final_T, pix, fT);
                a.n_contrib[pix] = nc;
            }
    }
}
extern "C" void gso_blend_f32(const BlendArgs<float>* a) {
    if (a->jvp) blend_impl<Dual<float>, float>(*a); else blend_impl<float, float>(*a);
}
extern "C" void gso_blend_f64(const BlendArgs<double>* a) {
    if (a->jvp) blend_impl<Dual<double>, double>(*a); else blend_impl<double, double>(*a);
}

// ----------------------------------------------------------------------------
// blend backward, backward.cu:452-638.  Per-Gaussian sums are accumulated in
// double (the reference uses float atomicAdd in arbitrary order).
// ----------------------------------------------------------------------------
template <class T> struct BlendBwdArgs {
    int P, W, H;
    const uint32_t* ranges;
    const uint32_t* point_list;
    const T *bg, *xy, *conic_opacity, *rgb, *depths, *final_T;
    const uint32_t* n_contrib;
    const T* dL_dpix;       // [3][H][W]
    const T* dL_dinvdepth;  // [H][W] or null
    T* dL_dmean2D;          // [P][3]  (x,y used)
    T* dL_dconic;           // [P][4]  (x,y,w used)
    T* dL_dopacity;         // [P]
    T* dL_dcolors;          // [P][3]
    T* dL_dinvdepths;       // [P] or null
};

template <class T> static void blend_bwd_impl(const BlendBwdArgs<T>& a) {
    int gx = (a.W + BLOCK_X - 1) / BLOCK_X, gy = (a.H + BLOCK_Y - 1) / BLOCK_Y;
    size_t HW = (size_t)a.H * a.W;
    size_t P = a.P;
    std::vector<double> acc(P * 10, 0.0);  // mean2D.xy, conic.xyw, opacity, color[3], invdepth
    const T ddelx_dx = T(0.5) * a.W, ddely_dy = T(0.5) * a.H;
    const bool has_d = a.dL_dinvdepth != nullptr && a.dL_dinvdepths != nullptr;
#pragma omp parallel for schedule(dynamic, 1)
    for (int t = 0; t < gx * gy; t++) {
        int tx = t % gx, ty = t / gx;
        uint32_t r0 = a.ranges[2 * t], r1 = a.ranges[2 * t + 1];
        for (int ly = 0; ly < BLOCK_Y; ly++)
            for (int lx = 0; lx < BLOCK_X; lx++) {
                int px = tx * BLOCK_X + lx, py = ty * BLOCK_Y + ly;
                if (px >= a.W || py >= a.H) continue;
                size_t pix = (size_t)py * a.W + px;
                const T T_final = a.final_T[pix];
                T Tr = T_final;
                uint32_t contributor = r1 - r0;
                const uint32_t last_contributor = a.n_contrib[pix];
                T accum_rec[3] = {0, 0, 0}, dL_dpixel[3], last_color[3] = {0, 0, 0};
                for (int ch = 0; ch < 3; ch++) dL_dpixel[ch] = a.dL_dpix[ch * HW + pix];
                T dL_invdepth = has_d ? a.dL_dinvdepth[pix] : T(0);
                T accum_invdepth_rec = 0, last_alpha = 0, last_invdepth = 0;
                T pfx = (T)px, pfy = (T)py;
                for (uint32_t k = r1; k-- > r0;) {
                    contributor--;
                    if (contributor >= last_contributor) continue;
                    uint32_t id = a.point_list[k];
                    T dx = a.xy[2 * (size_t)id] - pfx, dy = a.xy[2 * (size_t)id + 1] - pfy;
                    const T* co = a.conic_opacity + 4 * (size_t)id;
                    T power = T(-0.5f) * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
                    if (power > T(0)) continue;
                    T G = std::exp(power);
                    T alpha = std::min(T(0.99f), co[3] * G);
                    if (alpha < T(1.0f / 255.0f)) continue;
                    Tr = Tr / (T(1) - alpha);
                    T dchannel_dcolor = alpha * Tr;
                    T dL_dalpha = 0;
                    double* ac = &acc[(size_t)id * 10];
                    for (int ch = 0; ch < 3; ch++) {
                        T c = a.rgb[3 * (size_t)id + ch];
                        accum_rec[ch] = last_alpha * last_color[ch] + (T(1) - last_alpha) * accum_rec[ch];
                        last_color[ch] = c;
                        dL_dalpha += (c - accum_rec[ch]) * dL_dpixel[ch];
#pragma omp atomic
                        ac[6 + ch] += (double)(dchannel_dcolor * dL_dpixel[ch]);
                    }
                    if (has_d) {
                        T invd = T(1) / a.depths[id];
                        accum_invdepth_rec = last_alpha * last_invdepth + (T(1) - last_alpha) * accum_invdepth_rec;
                        last_invdepth = invd;
                        dL_dalpha += (invd - accum_invdepth_rec) * dL_invdepth;
#pragma omp atomic
                        ac[9] += (double)(dchannel_dcolor * dL_invdepth);
                    }
                    dL_dalpha *= Tr;
                    last_alpha = alpha;
                    T bg_dot = 0;
                    for (int i = 0; i < 3; i++) bg_dot += a.bg[i] * dL_dpixel[i];
                    dL_dalpha += (-T_final / (T(1) - alpha)) * bg_dot;
                    T dL_dG = co[3] * dL_dalpha;
                    T gdx = G * dx, gdy = G * dy;
                    T dG_ddelx = -gdx * co[0] - gdy * co[1];
                    T dG_ddely = -gdy * co[2] - gdx * co[1];
#pragma omp atomic
                    ac[0] += (double)(dL_dG * dG_ddelx * ddelx_dx);
#pragma omp atomic
                    ac[1] += (double)(dL_dG * dG_ddely * ddely_dy);
#pragma omp atomic
                    ac[2] += (double)(T(-0.5f) * gdx * dx * dL_dG);
#pragma omp atomic
                    ac[3] += (double)(T(-0.5f) * gdx * dy * dL_dG);
#pragma omp atomic
                    ac[4] += (double)(T(-0.5f) * gdy * dy * dL_dG);
#pragma omp atomic
                    ac[5] += (double)(G * dL_dalpha);
                }
            }
    }
#pragma omp parallel for
    for (long long i = 0; i < (long long)P; i++) {
        const double* ac = &acc[(size_t)i * 10];
        a.dL_dmean2D[3 * i + 0] = (T)ac[0];
        a.dL_dmean2D[3 * i + 1] = (T)ac[1];
        a.dL_dmean2D[3 * i + 2] = 0;
        a.dL_dconic[4 * i + 0] = (T)ac[2];
        a.dL_dconic[4 * i + 1] = (T)ac[3];
        a.dL_dconic[4 * i + 2] = 0;
        a.dL_dconic[4 * i + 3] = (T)ac[4];
        a.dL_dopacity[i] = (T)ac[5];
        for (int ch = 0; ch < 3; ch++) a.dL_dcolors[3 * i + ch] = (T)ac[6 + ch];
        if (a.dL_dinvdepths) a.dL_dinvdepths[i] = (T)ac[9];
    }
}
extern "C" void gso_blend_bwd_f32(const BlendBwdArgs<float>* a) { blend_bwd_impl<float>(*a); }
extern "C" void gso_blend_bwd_f64(const BlendBwdArgs<double>* a) { blend_bwd_impl<double>(*a); }

// ----------------------------------------------------------------------------
// per-Gaussian backward: backward.cu:147-326 (computeCov2DCUDA),
// :398-449 (preprocessCUDA bwd), :23-142 (SH bwd), :330-393 (cov3D bwd).
// ----------------------------------------------------------------------------
template <class T> struct PreBwdArgs {
    int P, D, M, W, H, antialiasing;
    const T* means3D;
    const int32_t* radii;
    const T* shs;               // null if colours were precomputed
    const uint8_t* clamped;
    const T* opacities;
    const T *scales, *rotations;  // null if cov3D was precomputed
    T scale_modifier;
    const T* cov3D;             // [P][6] (geometry state or precomputed)
    const T *viewmatrix, *projmatrix, *campos;
    T tan_fovx, tan_fovy;
    const T* dL_dmean2D;        // [P][3]
    const T* dL_dconic;         // [P][4]
    const T* dL_dinvdepth;      // [P] or null
    T* dL_dopacity;             // [P] in/out (rescaled when antialiasing)
    const T* dL_dcolor;         // [P][3]
    T *dL_dmean3D, *dL_dcov3D, *dL_dsh, *dL_dscale, *dL_drot;  // outputs, caller-zeroed
};

template <class T> static inline T sq(T x) { return x * x; }

template <class T> static void pre_bwd_impl(const PreBwdArgs<T>& a) {
    const T h_x = T(a.W) / (T(2.0f) * a.tan_fovx);
    const T h_y = T(a.H) / (T(2.0f) * a.tan_fovy);
    const T* vm = a.viewmatrix;
    const T* proj = a.projmatrix;
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < a.P; idx++) {
        if (!(a.radii[idx] > 0)) continue;
        size_t i = idx;
        // ---------------- computeCov2DCUDA ----------------
        const T* cov3D = a.cov3D + 6 * i;
        V3<T> mean = {a.means3D[3 * i], a.means3D[3 * i + 1], a.means3D[3 * i + 2]};
        T dcx = a.dL_dconic[4 * i], dcy = a.dL_dconic[4 * i + 1], dcz = a.dL_dconic[4 * i + 3];
        V3<T> t = transform4x3(mean, vm);
        const T limx = T(1.3f) * a.tan_fovx, limy = T(1.3f) * a.tan_fovy;
        const T txtz = t.x / t.z, tytz = t.y / t.z;
        t.x = std::min(limx, std::max(-limx, txtz)) * t.z;
        t.y = std::min(limy, std::max(-limy, tytz)) * t.z;
        const T x_grad_mul = (txtz < -limx || txtz > limx) ? T(0) : T(1);
        const T y_grad_mul = (tytz < -limy || tytz > limy) ? T(0) : T(1);
        M3<T> J = M3<T>::cols(h_x / t.z, 0, -(h_x * t.x) / (t.z * t.z), 0, h_y / t.z, -(h_y * t.y) / (t.z * t.z), 0, 0, 0);
        M3<T> W = M3<T>::cols(vm[0], vm[4], vm[8], vm[1], vm[5], vm[9], vm[2], vm[6], vm[10]);
        M3<T> Vrk = M3<T>::cols(cov3D[0], cov3D[1], cov3D[2], cov3D[1], cov3D[3], cov3D[4], cov3D[2], cov3D[4], cov3D[5]);
        M3<T> Tm = W * J;
        M3<T> cov2D = transpose(Tm) * transpose(Vrk) * Tm;
        T c_xx = cov2D.m[0][0], c_xy = cov2D.m[0][1], c_yy = cov2D.m[1][1];
        const T h_var = T(0.3f);
        T d_inside_root = 0;
        if (a.antialiasing) {
            const T det_cov = c_xx * c_yy - c_xy * c_xy;
            c_xx += h_var; c_yy += h_var;
            const T det_cov_plus_h = c_xx * c_yy - c_xy * c_xy;
            const T hcs = std::sqrt(std::max(T(0.000025f), det_cov / det_cov_plus_h));
            const T dL_dop = a.dL_dopacity[i];
            const T d_hcs = dL_dop * a.opacities[i];
            a.dL_dopacity[i] = dL_dop * hcs;
            d_inside_root = (det_cov / det_cov_plus_h) <= T(0.000025f) ? T(0) : d_hcs / (2 * hcs);
        } else {
            c_xx += h_var; c_yy += h_var;
        }
        T dL_dc_xx = 0, dL_dc_xy = 0, dL_dc_yy = 0;
        if (a.antialiasing) {
            const T x = c_xx, y = c_yy, z = c_xy, w = h_var;
            const T denom_f = d_inside_root / sq(w * w + w * (x + y) + x * y - z * z);
            dL_dc_xx = w * (w * y + y * y + z * z) * denom_f;
            dL_dc_yy = w * (w * x + x * x + z * z) * denom_f;
            dL_dc_xy = T(-2.f) * w * z * (w + x + y) * denom_f;
        }
        T denom = c_xx * c_yy - c_xy * c_xy;
        T denom2inv = T(1.0f) / ((denom * denom) + T(0.0000001f));
        T* dcov = a.dL_dcov3D + 6 * i;
        auto Tt = [&](int c, int r) { return Tm.m[c][r]; };
        if (denom2inv != 0) {
            dL_dc_xx += denom2inv * (-c_yy * c_yy * dcx + 2 * c_xy * c_yy * dcy + (denom - c_xx * c_yy) * dcz);
            dL_dc_yy += denom2inv * (-c_xx * c_xx * dcz + 2 * c_xx * c_xy * dcy + (denom - c_xx * c_yy) * dcx);
            dL_dc_xy += denom2inv * 2 * (c_xy * c_yy * dcx - (denom + 2 * c_xy * c_xy) * dcy + c_xx * c_xy * dcz);
            dcov[0] = (Tt(0, 0) * Tt(0, 0) * dL_dc_xx + Tt(0, 0) * Tt(1, 0) * dL_dc_xy + Tt(1, 0) * Tt(1, 0) * dL_dc_yy);
            dcov[3] = (Tt(0, 1) * Tt(0, 1) * dL_dc_xx + Tt(0, 1) * Tt(1, 1) * dL_dc_xy + Tt(1, 1) * Tt(1, 1) * dL_dc_yy);
            dcov[5] = (Tt(0, 2) * Tt(0, 2) * dL_dc_xx + Tt(0, 2) * Tt(1, 2) * dL_dc_xy + Tt(1, 2) * Tt(1, 2) * dL_dc_yy);
            dcov[1] = 2 * Tt(0, 0) * Tt(0, 1) * dL_dc_xx + (Tt(0, 0) * Tt(1, 1) + Tt(0, 1) * Tt(1, 0)) * dL_dc_xy + 2 * Tt(1, 0) * Tt(1, 1) * dL_dc_yy;
            dcov[2] = 2 * Tt(0, 0) * Tt(0, 2) * dL_dc_xx + (Tt(0, 0) * Tt(1, 2) + Tt(0, 2) * Tt(1, 0)) * dL_dc_xy + 2 * Tt(1, 0) * Tt(1, 2) * dL_dc_yy;
            dcov[4] = 2 * Tt(0, 2) * Tt(0, 1) * dL_dc_xx + (Tt(0, 1) * Tt(1, 2) + Tt(0, 2) * Tt(1, 1)) * dL_dc_xy + 2 * Tt(1, 1) * Tt(1, 2) * dL_dc_yy;
        } else {
            for (int k = 0; k < 6; k++) dcov[k] = 0;
        }
        auto Vr = [&](int c, int r) { return Vrk.m[c][r]; };
        T dL_dT00 = 2 * (Tt(0, 0) * Vr(0, 0) + Tt(0, 1) * Vr(0, 1) + Tt(0, 2) * Vr(0, 2)) * dL_dc_xx + (Tt(1, 0) * Vr(0, 0) + Tt(1, 1) * Vr(0, 1) + Tt(1, 2) * Vr(0, 2)) * dL_dc_xy;
        T dL_dT01 = 2 * (Tt(0, 0) * Vr(1, 0) + Tt(0, 1) * Vr(1, 1) + Tt(0, 2) * Vr(1, 2)) * dL_dc_xx + (Tt(1, 0) * Vr(1, 0) + Tt(1, 1) * Vr(1, 1) + Tt(1, 2) * Vr(1, 2)) * dL_dc_xy;
        T dL_dT02 = 2 * (Tt(0, 0) * Vr(2, 0) + Tt(0, 1) * Vr(2, 1) + Tt(0, 2) * Vr(2, 2)) * dL_dc_xx + (Tt(1, 0) * Vr(2, 0) + Tt(1, 1) * Vr(2, 1) + Tt(1, 2) * Vr(2, 2)) * dL_dc_xy;
        T dL_dT10 = 2 * (Tt(1, 0) * Vr(0, 0) + Tt(1, 1) * Vr(0, 1) + Tt(1, 2) * Vr(0, 2)) * dL_dc_yy + (Tt(0, 0) * Vr(0, 0) + Tt(0, 1) * Vr(0, 1) + Tt(0, 2) * Vr(0, 2)) * dL_dc_xy;
        T dL_dT11 = 2 * (Tt(1, 0) * Vr(1, 0) + Tt(1, 1) * Vr(1, 1) + Tt(1, 2) * Vr(1, 2)) * dL_dc_yy + (Tt(0, 0) * Vr(1, 0) + Tt(0, 1) * Vr(1, 1) + Tt(0, 2) * Vr(1, 2)) * dL_dc_xy;
        T dL_dT12 = 2 * (Tt(1, 0) * Vr(2, 0) + Tt(1, 1) * Vr(2, 1) + Tt(1, 2) * Vr(2, 2)) * dL_dc_yy + (Tt(0, 0) * Vr(2, 0) + Tt(0, 1) * Vr(2, 1) + Tt(0, 2) * Vr(2, 2)) * dL_dc_xy;
        T dL_dJ00 = W.m[0][0] * dL_dT00 + W.m[0][1] * dL_dT01 + W.m[0][2] * dL_dT02;
        T dL_dJ02 = W.m[2][0] * dL_dT00 + W.m[2][1] * dL_dT01 + W.m[2][2] * dL_dT02;
        T dL_dJ11 = W.m[1][0] * dL_dT10 + W.m[1][1] * dL_dT11 + W.m[1][2] * dL_dT12;
        T dL_dJ12 = W.m[2][0] * dL_dT10 + W.m[2][1] * dL_dT11 + W.m[2][2] * dL_dT12;
        T tz = T(1.f) / t.z, tz2 = tz * tz, tz3 = tz2 * tz;
        T dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
        T dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
        T dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 + (2 * h_y * t.y) * tz3 * dL_dJ12;
        if (a.dL_dinvdepth) dL_dtz -= a.dL_dinvdepth[i] / (t.z * t.z);
        // transformVec4x3Transpose, auxiliary.h:134-142
        V3<T> dL_dmean = {vm[0] * dL_dtx + vm[1] * dL_dty + vm[2] * dL_dtz,
                          vm[4] * dL_dtx + vm[5] * dL_dty + vm[6] * dL_dtz,
                          vm[8] * dL_dtx + vm[9] * dL_dty + vm[10] * dL_dtz};
        // ---------------- preprocessCUDA (bwd) ----------------
        V3<T> m = mean;
        T m_hw = proj[3] * m.x + proj[7] * m.y + proj[11] * m.z + proj[15];
        T m_w = T(1.0f) / (m_hw + T(0.0000001f));
        T mul1 = (proj[0] * m.x + proj[4] * m.y + proj[8] * m.z + proj[12]) * m_w * m_w;
        T mul2 = (proj[1] * m.x + proj[5] * m.y + proj[9] * m.z + proj[13]) * m_w * m_w;
        T g2x = a.dL_dmean2D[3 * i], g2y = a.dL_dmean2D[3 * i + 1];
        dL_dmean.x += (proj[0] * m_w - proj[3] * mul1) * g2x + (proj[1] * m_w - proj[3] * mul2) * g2y;
        dL_dmean.y += (proj[4] * m_w - proj[7] * mul1) * g2x + (proj[5] * m_w - proj[7] * mul2) * g2y;
        dL_dmean.z += (proj[8] * m_w - proj[11] * mul1) * g2x + (proj[9] * m_w - proj[11] * mul2) * g2y;

        if (a.shs) {
            // backward.cu:23-142
            V3<T> campos = {a.campos[0], a.campos[1], a.campos[2]};
            V3<T> dir_orig = m - campos;
            V3<T> dir = dir_orig / std::sqrt(dot3(dir_orig, dir_orig));
            const T* shp = a.shs + (size_t)i * a.M * 3;
            auto sh = [&](int k) { return V3<T>{shp[3 * k], shp[3 * k + 1], shp[3 * k + 2]}; };
            V3<T> dL_dRGB = {a.dL_dcolor[3 * i], a.dL_dcolor[3 * i + 1], a.dL_dcolor[3 * i + 2]};
            dL_dRGB.x *= a.clamped[3 * i + 0] ? 0 : 1;
            dL_dRGB.y *= a.clamped[3 * i + 1] ? 0 : 1;
            dL_dRGB.z *= a.clamped[3 * i + 2] ? 0 : 1;
            V3<T> dRGBdx = {0, 0, 0}, dRGBdy = {0, 0, 0}, dRGBdz = {0, 0, 0};
            T x = dir.x, y = dir.y, z = dir.z;
            T* dsh = a.dL_dsh + (size_t)i * a.M * 3;
            auto put = [&](int k, T s) { dsh[3 * k] = s * dL_dRGB.x; dsh[3 * k + 1] = s * dL_dRGB.y; dsh[3 * k + 2] = s * dL_dRGB.z; };
            const T C0 = SH_C0, C1 = SH_C1;
            const T C2[5] = {SH_C2[0], SH_C2[1], SH_C2[2], SH_C2[3], SH_C2[4]};
            const T C3[7] = {SH_C3[0], SH_C3[1], SH_C3[2], SH_C3[3], SH_C3[4], SH_C3[5], SH_C3[6]};
            put(0, C0);
            if (a.D > 0) {
                put(1, -C1 * y); put(2, C1 * z); put(3, -C1 * x);
                dRGBdx = (-C1) * sh(3);
                dRGBdy = (-C1) * sh(1);
                dRGBdz = C1 * sh(2);
                if (a.D > 1) {
                    T xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                    put(4, C2[0] * xy); put(5, C2[1] * yz); put(6, C2[2] * (T(2) * zz - xx - yy));
                    put(7, C2[3] * xz); put(8, C2[4] * (xx - yy));
                    dRGBdx = dRGBdx + (C2[0] * y) * sh(4) + (C2[2] * T(2) * -x) * sh(6) + (C2[3] * z) * sh(7) + (C2[4] * T(2) * x) * sh(8);
                    dRGBdy = dRGBdy + (C2[0] * x) * sh(4) + (C2[1] * z) * sh(5) + (C2[2] * T(2) * -y) * sh(6) + (C2[4] * T(2) * -y) * sh(8);
                    dRGBdz = dRGBdz + (C2[1] * y) * sh(5) + (C2[2] * T(2) * T(2) * z) * sh(6) + (C2[3] * x) * sh(7);
                    if (a.D > 2) {
                        put(9, C3[0] * y * (T(3) * xx - yy));
                        put(10, C3[1] * xy * z);
                        put(11, C3[2] * y * (T(4) * zz - xx - yy));
                        put(12, C3[3] * z * (T(2) * zz - T(3) * xx - T(3) * yy));
                        put(13, C3[4] * x * (T(4) * zz - xx - yy));
                        put(14, C3[5] * z * (xx - yy));
                        put(15, C3[6] * x * (xx - T(3) * yy));
                        dRGBdx = dRGBdx + (C3[0] * T(3) * T(2) * xy) * sh(9) + (C3[1] * yz) * sh(10) + (C3[2] * T(-2) * xy) * sh(11) +
                                 (C3[3] * T(-3) * T(2) * xz) * sh(12) + (C3[4] * (T(-3) * xx + T(4) * zz - yy)) * sh(13) +
                                 (C3[5] * T(2) * xz) * sh(14) + (C3[6] * T(3) * (xx - yy)) * sh(15);
                        dRGBdy = dRGBdy + (C3[0] * T(3) * (xx - yy)) * sh(9) + (C3[1] * xz) * sh(10) +
                                 (C3[2] * (T(-3) * yy + T(4) * zz - xx)) * sh(11) + (C3[3] * T(-3) * T(2) * yz) * sh(12) +
                                 (C3[4] * T(-2) * xy) * sh(13) + (C3[5] * T(-2) * yz) * sh(14) + (C3[6] * T(-3) * T(2) * xy) * sh(15);
                        dRGBdz = dRGBdz + (C3[1] * xy) * sh(10) + (C3[2] * T(4) * T(2) * yz) * sh(11) +
                                 (C3[3] * T(3) * (T(2) * zz - xx - yy)) * sh(12) + (C3[4] * T(4) * T(2) * xz) * sh(13) +
                                 (C3[5] * (xx - yy)) * sh(14);
                    }
                }
            }
            V3<T> dL_ddir = {dot3(dRGBdx, dL_dRGB), dot3(dRGBdy, dL_dRGB), dot3(dRGBdz, dL_dRGB)};
            // dnormvdv, auxiliary.h:153-164
            V3<T> v = dir_orig, dv = dL_ddir;
            T sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
            T invsum32 = T(1.0f) / std::sqrt(sum2 * sum2 * sum2);
            dL_dmean.x += ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
            dL_dmean.y += (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
            dL_dmean.z += (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
        }
        a.dL_dmean3D[3 * i] = dL_dmean.x;
        a.dL_dmean3D[3 * i + 1] = dL_dmean.y;
        a.dL_dmean3D[3 * i + 2] = dL_dmean.z;

        if (a.scales) {
            // backward.cu:330-393
            const T* q = a.rotations + 4 * i;
            T r = q[0], x = q[1], y = q[2], z = q[3];
            M3<T> R = M3<T>::cols(T(1) - T(2) * (y * y + z * z), T(2) * (x * y - r * z), T(2) * (x * z + r * y),
                                  T(2) * (x * y + r * z), T(1) - T(2) * (x * x + z * z), T(2) * (y * z - r * x),
                                  T(2) * (x * z - r * y), T(2) * (y * z + r * x), T(1) - T(2) * (x * x + y * y));
            V3<T> s = {a.scale_modifier * a.scales[3 * i], a.scale_modifier * a.scales[3 * i + 1], a.scale_modifier * a.scales[3 * i + 2]};
            M3<T> Sm = M3<T>::cols(s.x, 0, 0, 0, s.y, 0, 0, 0, s.z);
            M3<T> M = Sm * R;
            const T* d = dcov;
            M3<T> dL_dSigma = M3<T>::cols(d[0], T(0.5f) * d[1], T(0.5f) * d[2], T(0.5f) * d[1], d[3], T(0.5f) * d[4],
                                          T(0.5f) * d[2], T(0.5f) * d[4], d[5]);
            M3<T> MdS = M * dL_dSigma;
            M3<T> dL_dM;
            for (int c = 0; c < 3; c++) for (int rr = 0; rr < 3; rr++) dL_dM.m[c][rr] = T(2.0f) * MdS.m[c][rr];
            M3<T> Rt = transpose(R);
            M3<T> dL_dMt = transpose(dL_dM);
            auto col = [](const M3<T>& A, int c) { return V3<T>{A.m[c][0], A.m[c][1], A.m[c][2]}; };
            a.dL_dscale[3 * i + 0] = dot3(col(Rt, 0), col(dL_dMt, 0));
            a.dL_dscale[3 * i + 1] = dot3(col(Rt, 1), col(dL_dMt, 1));
            a.dL_dscale[3 * i + 2] = dot3(col(Rt, 2), col(dL_dMt, 2));
            for (int k = 0; k < 3; k++) { dL_dMt.m[0][k] *= s.x; dL_dMt.m[1][k] *= s.y; dL_dMt.m[2][k] *= s.z; }
            auto D = [&](int c, int rr) { return dL_dMt.m[c][rr]; };
            a.dL_drot[4 * i + 0] = 2 * z * (D(0, 1) - D(1, 0)) + 2 * y * (D(2, 0) - D(0, 2)) + 2 * x * (D(1, 2) - D(2, 1));
            a.dL_drot[4 * i + 1] = 2 * y * (D(1, 0) + D(0, 1)) + 2 * z * (D(2, 0) + D(0, 2)) + 2 * r * (D(1, 2) - D(2, 1)) - 4 * x * (D(2, 2) + D(1, 1));
            a.dL_drot[4 * i + 2] = 2 * x * (D(1, 0) + D(0, 1)) + 2 * r * (D(2, 0) - D(0, 2)) + 2 * z * (D(1, 2) + D(2, 1)) - 4 * y * (D(2, 2) + D(0, 0));
            a.dL_drot[4 * i + 3] = 2 * r * (D(0, 1) - D(1, 0)) + 2 * x * (D(2, 0) + D(0, 2)) + 2 * y * (D(1, 2) + D(2, 1)) - 4 * z * (D(1, 1) + D(0, 0));
        }
    }
}
extern "C" void gso_preprocess_bwd_f32(const PreBwdArgs<float>* a) { pre_bwd_impl<float>(*a); }
extern "C" void gso_preprocess_bwd_f64(const PreBwdArgs<double>* a) { pre_bwd_impl<double>(*a); }
