// =============================================================================
// binning.cu -- tile binning (a8): scan, key generation, tile|depth radix sort, tile
// ranges; plus the compaction that turns the instances the forward pass actually blended
// into dense 48-byte records in tile|depth order (what the derivative tile kernels stage with
// one bulk copy per batch) and a per-Gaussian CSR over them (deterministic gradient gather).
//
// Reference: DGR/cuda_rasterizer/rasterizer_impl.cu:35-50 (getHigherMsb), :70-111
// (duplicateWithKeys), :116-138 (identifyTileRanges), :321-364 (orchestration).
// Integer work, bit-exact by construction: same key layout (tile << 32 | depth bits),
// same emission order (row-major over the rect, ascending Gaussian id), a stable LSD
// radix sort over bits [0, 32 + msb(tiles)), so equal keys keep emission order.
// The sort itself is cub::DeviceRadixSort (CUDA toolkit library, as in the reference).
// =============================================================================
#include "gslm_common.cuh"
#include "gslm_math.cuh"
#include <cub/cub.cuh>
#include <cstdarg>
#include <cstdio>

namespace gslm {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
const char* last_error() { return g_err; }

// rasterizer_impl.cu:35-50
static uint32_t higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4;
    uint32_t step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

size_t scan_temp_bytes(int P) {
    // cub::DeviceScan temp storage is O(number of tiles of the scan); bound it without
    // touching the device so that size queries work on a machine without a GPU.
    return align_up((size_t)(P / 128 + 1024) * 16 + 4096);
}
size_t sort_temp_bytes(int64_t n) {
    // cub::DeviceRadixSort (non-DoubleBuffer overload) on (u32 key, u32 value) pairs needs one alternate key and
    // value buffer (8 B / item) plus the onesweep look-back table: 256 counters x 4 B per tile of >= 2048 items
    // (<= 0.5 B / item) plus per-pass histograms.  The bound is checked against CUB's own answer at run time.
    return align_up((size_t)n * (2 * sizeof(uint32_t)) + (size_t)n / 2 + ((size_t)1 << 20));
}

template <class T> static T* take(char*& p, size_t n) {
    T* r = reinterpret_cast<T*>(p);
    p += align_up(n * sizeof(T));
    return r;
}

GeomView carve_geom(void* base, int P) {
    GeomView v;
    char* p = reinterpret_cast<char*>(base);
    size_t n = (size_t)P;
    v.depths = take<float>(p, n);
    v.means2D = take<float2>(p, n);
    v.conic_opacity = take<float4>(p, n);
    v.rgb = take<float>(p, 3 * n);
    v.cov3D = take<float>(p, 6 * n);
    v.clamped = take<uint8_t>(p, n);
    v.tiles_touched = take<uint32_t>(p, n);
    v.point_offsets = take<uint32_t>(p, n);
    v.num_rendered = take<uint32_t>(p, 32);
    v.g_off = take<uint32_t>(p, n + 1);
    v.g_cursor = take<uint32_t>(p, n);
    v.depth_sorted = take<uint32_t>(p, n);
    v.depth_perm = take<uint32_t>(p, n);
    v.active_ids = v.g_cursor;
    v.n_active = v.num_rendered + 1;
    v.act_range = take<uint32_t>(p, n / ACT_RANGE + 2);
    v.cull_mask = take<uint2>(p, n);
    v.scan_temp_bytes = scan_temp_bytes(P);
    v.scan_temp = take<char>(p, v.scan_temp_bytes);
    v.total = (size_t)(p - reinterpret_cast<char*>(base));
    return v;
}
ImageView carve_image(void* base, int W, int H) {
    ImageView v;
    char* p = reinterpret_cast<char*>(base);
    size_t N = (size_t)W * H, T = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
    v.final_T = take<float>(p, N);
    v.n_contrib = take<uint32_t>(p, N);
    v.n_contrib_c = take<uint32_t>(p, N);
    v.ranges = take<uint2>(p, T);
    v.ccount = take<uint32_t>(p, T);
    v.cstart = take<uint32_t>(p, T + 1);
    v.order = take<uint32_t>(p, T);
    v.total = (size_t)(p - reinterpret_cast<char*>(base));
    return v;
}
BinningView carve_binning(void* base, int64_t Rc) {
    BinningView v;
    char* p = reinterpret_cast<char*>(base);
    size_t n = (size_t)Rc;
    v.records = take<float4>(p, REC_F4 * n);
    v.masks = take<uint4>(p, 2 * n);
    v.cid = take<uint32_t>(p, n);
    v.g_list = take<uint32_t>(p, n);
    v.g_inv = take<uint32_t>(p, n);
    v.total = (size_t)(p - reinterpret_cast<char*>(base));
    return v;
}
BinningTemp carve_binning_temp(void* base, int64_t R, int P) {
    BinningTemp v;
    char* p = reinterpret_cast<char*>(base);
    size_t n = (size_t)R;
    v.keys_sorted = take<uint32_t>(p, n);
    v.keys_unsorted = take<uint32_t>(p, n);
    v.vals_unsorted = take<uint32_t>(p, n);
    v.point_list = take<uint32_t>(p, n);
    v.records_tmp = take<float4>(p, REC_F4 * n);
    v.masks_tmp = take<uint4>(p, 2 * n);
    v.sort_temp_bytes = sort_temp_bytes(R > P ? R : (int64_t)P);   // the same scratch serves the depth sort over P items
    v.sort_temp = take<char>(p, v.sort_temp_bytes);
    v.total = (size_t)(p - reinterpret_cast<char*>(base));
    return v;
}

int run_scan(const GeomView& gv, int P, cudaStream_t s) {
    size_t need = 0;
    GSLM_CUDA_OK(cub::DeviceScan::InclusiveSum(nullptr, need, gv.tiles_touched, gv.point_offsets, P, s));
    if (need > gv.scan_temp_bytes) {
        set_error("scan temp storage too small: need %zu have %zu", need, gv.scan_temp_bytes);
        return 1;
    }
    size_t have = gv.scan_temp_bytes;
    StageScope ts(ST_SCAN, s);
    GSLM_CUDA_OK(cub::DeviceScan::InclusiveSum(gv.scan_temp, have, gv.tiles_touched, gv.point_offsets, P, s));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// The reference sorts R (tile << 32 | depth bits) keys with a stable LSD radix sort over 32 + msb(T) bits (6 passes
// of 12 B per instance at 1080p).  The same order comes out of a two-level sort that moves a third of the bytes:
//   1. sort the P Gaussians by depth bits, stably from ascending id  (4 passes over P << R items);
//   2. emit the instances in THAT order with the tile id alone as key;
//   3. sort them stably by tile id                                  (2 passes of 8 B per instance).
// Within a tile the instances then appear by (depth, id) -- exactly the reference's order: its ties in (tile, depth)
// keep emission order, which is ascending Gaussian id (rasterizer_impl.cu:70-111, :346-354).  Bit-exact
// point_list and ranges are asserted against the reference fixture and the reference's own kernels in tests/.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_depth_keys(int P, const float* __restrict__ depths, const int32_t* __restrict__ radii, uint32_t* __restrict__ dkey,
             uint32_t* __restrict__ iota) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    dkey[idx] = radii[idx] > 0 ? __float_as_uint(depths[idx]) : 0xffffffffu;   // depths > 0.2: bits are monotone
    iota[idx] = (uint32_t)idx;
}
__global__ void __launch_bounds__(256)
k_gather_tiles(int P, const uint32_t* __restrict__ perm, const uint32_t* __restrict__ tiles_touched,
               uint32_t* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < P) out[i] = tiles_touched[perm[i]];
}

// rasterizer_impl.cu:70-111, in depth order. One thread per Gaussian; the rect is recomputed from the stored pixel
// centre and radius exactly as the reference does (:91).
__global__ void __launch_bounds__(256)
k_duplicate_keys(int P, const uint32_t* __restrict__ perm, const float2* __restrict__ xy,
                 const uint32_t* __restrict__ offsets, const int32_t* __restrict__ radii, int gx, int gy,
                 uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    // A warp takes 32 consecutive Gaussians of the depth order; each lane works out its Gaussian's tile rectangle,
    // then the WARP emits the rectangles one Gaussian at a time, lane t writing instance t, t + 32, ...: the stores
    // of a Gaussian's run are contiguous (the reference's one-thread-per-Gaussian loop, rasterizer_impl.cu:84-110,
    // writes 8-byte pieces at 32 unrelated addresses per instruction).  Same keys, same values, same positions.
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t idx = 0, off = 0;
    int x0 = 0, y0 = 0, w = 0, cnt = 0;
    if (i < P) {
        idx = perm[i];
        const int rad = radii[idx];
        if (rad > 0) {
            const float2 p = xy[idx];
            int x1, y1;
            tile_rect(p.x, p.y, rad, gx, gy, x0, y0, x1, y1);
            w = x1 - x0;
            cnt = w > 0 && y1 > y0 ? w * (y1 - y0) : 0;
            off = offsets[i];
        }
    }
    uint32_t todo = __ballot_sync(0xffffffffu, cnt > 0);
    while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const uint32_t g_idx = __shfl_sync(0xffffffffu, idx, src), g_off = __shfl_sync(0xffffffffu, off, src);
        const int g_x0 = __shfl_sync(0xffffffffu, x0, src), g_y0 = __shfl_sync(0xffffffffu, y0, src);
        const int g_w = __shfl_sync(0xffffffffu, w, src), g_cnt = __shfl_sync(0xffffffffu, cnt, src);
        for (int t = lane; t < g_cnt; t += 32) {
            const int ry = t / g_w, rx = t - ry * g_w;
            keys[g_off + t] = (uint32_t)((g_y0 + ry) * gx + g_x0 + rx);
            vals[g_off + t] = g_idx;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Exact tile culling (opt-in, GslmCamera::tile_culling; no reference counterpart).  The reference bins a Gaussian
// into every tile of the square that bounds its 3-sigma circle (auxiliary.h:60-70), whatever its opacity and
// orientation; most of those (tile, Gaussian) instances never reach alpha >= 1/255 at any pixel of the tile and are
// skipped pixel by pixel in renderCUDA (forward.cu:362-364).  Here such an instance is not emitted at all when a
// conservative bound proves it: with q(d) = (A dx^2 + C dy^2)/2 + B dx dy (so that power = -q), the blend test
// alpha = o exp(-q) >= 1/255 needs q <= ln(255 o); q is a convex quadratic whose minimum over the tile's pixel box,
// when the centre is outside the box, lies on one of the (at most two) box edges facing the centre, where it is a
// clamped 1-D parabola.  The bound is relaxed by the rounding error of the fp32 evaluations on both sides, so a
// culled instance is one renderCUDA would have skipped at every pixel: the image, final_T and the compact
// linearisation cache are bit-identical with and without culling; only num_rendered / point_list get shorter.
// Every operation is an explicitly rounded intrinsic: the count pass and the emission pass must agree bit for bit.
// ---------------------------------------------------------------------------------------------
struct CullGauss { float x, y, A, B, C, rA, rC, tau; };

__device__ __forceinline__ CullGauss cull_setup(float2 xy, float4 co) {
    CullGauss g;
    g.x = xy.x; g.y = xy.y; g.A = co.x; g.B = co.y; g.C = co.z;
    g.rA = __frcp_rn(co.x); g.rC = __frcp_rn(co.z);
    g.tau = co.w > 0.f ? logf(__fmul_rn(255.f, co.w)) : -INFINITY;
    return g;
}
__device__ __forceinline__ float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }
// min over t in [lo, hi] of (K t^2)/2 + L t, K > 0 (rK = 1/K; the value is flat around its minimiser, so the rounding
// of the minimiser itself is a second-order effect, far inside the slack below)
__device__ __forceinline__ float parabola_min(float K, float rK, float L, float lo, float hi) {
    const float t = clampf(__fmul_rn(-L, rK), lo, hi);
    return __fmul_rn(t, __fmaf_rn(__fmul_rn(0.5f, K), t, L));
}
__device__ __forceinline__ bool tile_can_blend(const CullGauss& g, int tx, int ty, int W, int H) {
    const float px0 = (float)(tx * TILE_X), px1 = (float)min(tx * TILE_X + TILE_X - 1, W - 1);
    const float py0 = (float)(ty * TILE_Y), py1 = (float)min(ty * TILE_Y + TILE_Y - 1, H - 1);
    const float dx_lo = __fsub_rn(g.x, px1), dx_hi = __fsub_rn(g.x, px0);     // d = centre - pixel (forward.cu:355)
    const float dy_lo = __fsub_rn(g.y, py1), dy_hi = __fsub_rn(g.y, py0);
    const float cx = clampf(0.f, dx_lo, dx_hi), cy = clampf(0.f, dy_lo, dy_hi);
    if (!(g.A > 0.f) || !(g.C > 0.f)) return true;                           // degenerate conic: leave it to the blend
    float qmin = 0.f;
    if (cx != 0.f || cy != 0.f) {
        qmin = INFINITY;
        if (cx != 0.f)   // edge dx = cx: q = (C dy^2)/2 + (B cx) dy + (A cx^2)/2
            qmin = __fadd_rn(parabola_min(g.C, g.rC, __fmul_rn(g.B, cx), dy_lo, dy_hi),
                             __fmul_rn(__fmul_rn(0.5f, g.A), __fmul_rn(cx, cx)));
        if (cy != 0.f)
            qmin = fminf(qmin, __fadd_rn(parabola_min(g.A, g.rA, __fmul_rn(g.B, cy), dx_lo, dx_hi),
                                         __fmul_rn(__fmul_rn(0.5f, g.C), __fmul_rn(cy, cy))));
    }
    // rounding slack: |fp32 error| of -power at any pixel of the box and of qmin is a few ulp of the term magnitudes
    const float DX = fmaxf(fabsf(dx_lo), fabsf(dx_hi)), DY = fmaxf(fabsf(dy_lo), fabsf(dy_hi));
    const float smax = __fmaf_rn(fabsf(g.B), __fmul_rn(DX, DY),
                                 __fmul_rn(0.5f, __fmaf_rn(g.A, __fmul_rn(DX, DX), __fmul_rn(g.C, __fmul_rn(DY, DY)))));
    const float slack = __fmaf_rn(4e-6f, smax, __fmaf_rn(1e-5f, fabsf(g.tau), 1e-3f));
    return !(__fsub_rn(qmin, slack) > g.tau);
}

// tiles_touched[idx] <- number of tiles of the rect that can blend (replaces the rect area before the scan), and the
// survivors of the first 64 tiles of the rect (row-major) as a bit mask for the emission pass.  Rects of up to 32
// tiles (the typical Gaussian has ~20) are walked by their own thread; the larger ones of a warp's 32 Gaussians are
// then tested by the whole warp, one tile per lane.
__global__ void __launch_bounds__(256)
k_cull_count(int P, const float2* __restrict__ xy, const float4* __restrict__ conic_opacity,
             const int32_t* __restrict__ radii, int gx, int gy, int W, int H, uint32_t* __restrict__ tiles_touched,
             uint2* __restrict__ cull_mask) {
    const int lane = threadIdx.x & 31;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    int x0 = 0, y0 = 0, x1 = 0, y1 = 0, w = 0, cnt = 0;
    float2 p = make_float2(0.f, 0.f);
    float4 co = make_float4(0.f, 0.f, 0.f, 0.f);
    if (idx < P && radii[idx] > 0) {
        p = xy[idx];
        co = conic_opacity[idx];
        tile_rect(p.x, p.y, radii[idx], gx, gy, x0, y0, x1, y1);
        w = x1 - x0;
        cnt = w > 0 && y1 > y0 ? w * (y1 - y0) : 0;
    }
    uint32_t n = 0, m0 = 0, m1 = 0;
    if (cnt > 0 && cnt <= 32) {
        // same bound as tile_can_blend with the row terms hoisted; its verdicts travel to the emission pass in the
        // mask, so this loop need not round like tile_can_blend, only be as conservative
        const CullGauss g = cull_setup(p, co);
        if (!(g.A > 0.f) || !(g.C > 0.f)) { n = cnt; m0 = cnt == 32 ? 0xffffffffu : (1u << cnt) - 1u; }
        else {
            const float hA = 0.5f * g.A, hC = 0.5f * g.C, aB = fabsf(g.B);
            const float thr = g.tau + fmaf(1e-5f, fabsf(g.tau), 1e-3f);
            int t = 0;
            for (int ty = y0; ty < y1; ++ty) {
                const float py0 = (float)(ty * TILE_Y), py1 = (float)min(ty * TILE_Y + TILE_Y - 1, H - 1);
                const float dy_lo = g.y - py1, dy_hi = g.y - py0;
                const float cy = clampf(0.f, dy_lo, dy_hi), DY = fmaxf(fabsf(dy_lo), fabsf(dy_hi));
                const float ty_c = hC * cy * cy, ty_L = g.B * cy, ty_s = hC * DY * DY;
                for (int tx = x0; tx < x1; ++tx, ++t) {
                    const float px0 = (float)(tx * TILE_X), px1 = (float)min(tx * TILE_X + TILE_X - 1, W - 1);
                    const float dx_lo = g.x - px1, dx_hi = g.x - px0;
                    const float cx = clampf(0.f, dx_lo, dx_hi), DX = fmaxf(fabsf(dx_lo), fabsf(dx_hi));
                    float qmin = 0.f;
                    if (cx != 0.f || cy != 0.f) {
                        // edge dx = cx (skipped when cx == 0: the other edge then holds the minimum) and edge dy = cy
                        const float L = g.B * cx, u = clampf(-L * g.rC, dy_lo, dy_hi);
                        const float e1 = cx != 0.f ? fmaf(u, fmaf(hC, u, L), hA * cx * cx) : INFINITY;
                        const float v = clampf(-ty_L * g.rA, dx_lo, dx_hi);
                        const float e2 = cy != 0.f ? fmaf(v, fmaf(hA, v, ty_L), ty_c) : INFINITY;
                        qmin = fminf(e1, e2);
                    }
                    const float smax = fmaf(aB * DX, DY, fmaf(hA * DX, DX, ty_s));
                    if (!(qmin - 4e-6f * smax > thr)) { n++; m0 |= 1u << t; }
                }
            }
        }
    }
    uint32_t todo = __ballot_sync(0xffffffffu, cnt > 32);
    while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const CullGauss g = cull_setup(make_float2(__shfl_sync(0xffffffffu, p.x, src), __shfl_sync(0xffffffffu, p.y, src)),
                                       make_float4(__shfl_sync(0xffffffffu, co.x, src), __shfl_sync(0xffffffffu, co.y, src),
                                                   __shfl_sync(0xffffffffu, co.z, src), __shfl_sync(0xffffffffu, co.w, src)));
        const int g_x0 = __shfl_sync(0xffffffffu, x0, src), g_y0 = __shfl_sync(0xffffffffu, y0, src);
        const int g_w = __shfl_sync(0xffffffffu, w, src), g_cnt = __shfl_sync(0xffffffffu, cnt, src);
        uint32_t gn = 0, b0 = 0, b1 = 0;
        for (int t0 = 0; t0 < g_cnt; t0 += 32) {
            const int t = t0 + lane, ry = t / g_w, rx = t - ry * g_w;
            const uint32_t bal = __ballot_sync(0xffffffffu, t < g_cnt && tile_can_blend(g, g_x0 + rx, g_y0 + ry, W, H));
            gn += __popc(bal);
            if (t0 == 0) b0 = bal;
            if (t0 == 32) b1 = bal;
        }
        if (lane == src) { n = gn; m0 = b0; m1 = b1; }
    }
    if (cnt > 0) {
        tiles_touched[idx] = n;
        cull_mask[idx] = make_uint2(m0, m1);
    }
}

// k_duplicate_keys with the culled counts: surviving tiles keep their row-major order inside the Gaussian's run.
// Rects of up to 64 tiles (nearly all) are emitted from the count pass's bit mask; larger ones repeat the test.
__global__ void __launch_bounds__(256)
k_duplicate_keys_culled(int P, const uint32_t* __restrict__ perm, const float2* __restrict__ xy,
                        const float4* __restrict__ conic_opacity, const uint2* __restrict__ cull_mask,
                        const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ tiles_sorted,
                        const int32_t* __restrict__ radii, int gx, int gy, int W, int H, uint32_t* __restrict__ keys,
                        uint32_t* __restrict__ vals) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t idx = 0, off = 0, kept = 0;
    int x0 = 0, y0 = 0, w = 0, cnt = 0;
    float2 p = make_float2(0.f, 0.f);
    float4 co = make_float4(0.f, 0.f, 0.f, 0.f);
    uint2 m = make_uint2(0u, 0u);
    if (i < P) {
        idx = perm[i];
        const int rad = radii[idx];
        kept = tiles_sorted[i];
        if (rad > 0 && kept > 0) {
            p = xy[idx];
            int x1, y1;
            tile_rect(p.x, p.y, rad, gx, gy, x0, y0, x1, y1);
            w = x1 - x0;
            cnt = w > 0 && y1 > y0 ? w * (y1 - y0) : 0;
            off = offsets[i];
            if (cnt > 64) co = conic_opacity[idx];
            else m = cull_mask[idx];
        }
    }
    uint32_t todo = __ballot_sync(0xffffffffu, cnt > 0);
    while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const uint32_t g_idx = __shfl_sync(0xffffffffu, idx, src), g_off = __shfl_sync(0xffffffffu, off, src);
        const int g_x0 = __shfl_sync(0xffffffffu, x0, src), g_y0 = __shfl_sync(0xffffffffu, y0, src);
        const int g_w = __shfl_sync(0xffffffffu, w, src), g_cnt = __shfl_sync(0xffffffffu, cnt, src);
        if (g_cnt <= 64) {
            const uint32_t m0 = __shfl_sync(0xffffffffu, m.x, src), m1 = __shfl_sync(0xffffffffu, m.y, src);
            const uint32_t below = (1u << lane) - 1u;
            // t / w for t < 64, w <= 64 without an integer division: (t + 1/2) / w is at least 1/(2w) away from
            // every integer, far more than the rounding of the reciprocal
            const float rw = __frcp_rn((float)g_w);
            if ((m0 >> lane) & 1u) {
                const int ry = (int)(((float)lane + 0.5f) * rw), rx = lane - ry * g_w;
                const uint32_t pos = __popc(m0 & below);
                keys[g_off + pos] = (uint32_t)((g_y0 + ry) * gx + g_x0 + rx);
                vals[g_off + pos] = g_idx;
            }
            if ((m1 >> lane) & 1u) {
                const int t = lane + 32, ry = (int)(((float)t + 0.5f) * rw), rx = t - ry * g_w;
                const uint32_t pos = __popc(m0) + __popc(m1 & below);
                keys[g_off + pos] = (uint32_t)((g_y0 + ry) * gx + g_x0 + rx);
                vals[g_off + pos] = g_idx;
            }
            continue;
        }
        const CullGauss g = cull_setup(make_float2(__shfl_sync(0xffffffffu, p.x, src), __shfl_sync(0xffffffffu, p.y, src)),
                                       make_float4(__shfl_sync(0xffffffffu, co.x, src), __shfl_sync(0xffffffffu, co.y, src),
                                                   __shfl_sync(0xffffffffu, co.z, src), __shfl_sync(0xffffffffu, co.w, src)));
        const uint32_t g_kept = __shfl_sync(0xffffffffu, kept, src);
        uint32_t base = 0;
        for (int t0 = 0; t0 < g_cnt; t0 += 32) {
            const int t = t0 + lane, ry = t / g_w, rx = t - ry * g_w;
            const bool keep = t < g_cnt && tile_can_blend(g, g_x0 + rx, g_y0 + ry, W, H);
            const uint32_t bal = __ballot_sync(0xffffffffu, keep);
            const uint32_t pos = base + __popc(bal & ((1u << lane) - 1u));
            if (keep && pos < g_kept) {           // pos < g_kept always holds (same test as the count pass)
                keys[g_off + pos] = (uint32_t)((g_y0 + ry) * gx + g_x0 + rx);
                vals[g_off + pos] = g_idx;
            }
            base += __popc(bal);
        }
    }
}

// rasterizer_impl.cu:116-138; ranges are zero-filled first (:356)
__global__ void __launch_bounds__(256)
k_tile_ranges(uint32_t L, const uint32_t* __restrict__ keys, uint2* __restrict__ ranges) {
    uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= L) return;
    uint32_t cur = keys[idx];
    if (idx == 0) ranges[cur].x = 0;
    else {
        uint32_t prev = keys[idx - 1];
        if (cur != prev) {
            ranges[prev].y = idx;
            ranges[cur].x = idx;
        }
    }
    if (idx == L - 1) ranges[cur].y = L;
}

static int sort_pairs_u32(const BinningTemp& bt, const uint32_t* kin, uint32_t* kout, const uint32_t* vin, uint32_t* vout,
                          int64_t n, int end_bit, cudaStream_t s) {
    size_t need = 0;
    GSLM_CUDA_OK(cub::DeviceRadixSort::SortPairs(nullptr, need, kin, kout, vin, vout, (int)n, 0, end_bit, s));
    if (need > bt.sort_temp_bytes) {
        set_error("sort temp storage too small: need %zu have %zu", need, bt.sort_temp_bytes);
        return 1;
    }
    size_t have = bt.sort_temp_bytes;
    GSLM_CUDA_OK(cub::DeviceRadixSort::SortPairs(bt.sort_temp, have, kin, kout, vin, vout, (int)n, 0, end_bit, s));
    return 0;
}

int run_tile_cull(const GeomView& gv, const int32_t* radii, int P, int W, int H, cudaStream_t s) {
    if (P == 0) return 0;
    const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    StageScope ts(ST_KEYS, s);
    k_cull_count<<<(P + 255) / 256, 256, 0, s>>>(P, gv.means2D, gv.conic_opacity, radii, gx, gy, W, H, gv.tiles_touched,
                                                  gv.cull_mask);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

int run_binning(const GeomView& gv, const int32_t* radii, const BinningTemp& bt, const ImageView& iv, int P, int W,
                int H, int64_t R, bool culled, cudaStream_t s) {
    const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    GSLM_CUDA_OK(cudaMemsetAsync(iv.ranges, 0, sizeof(uint2) * (size_t)gx * gy, s));
    if (R == 0 || P == 0) return 0;
    // g_off / g_cursor are free until compaction: depth keys and the identity permutation live there
    uint32_t* dkey = gv.g_off;
    uint32_t* iota = gv.g_cursor;
    stage_begin(ST_KEYS, s);
    k_depth_keys<<<(P + 255) / 256, 256, 0, s>>>(P, gv.depths, radii, dkey, iota);
    GSLM_CUDA_OK(cudaGetLastError());
    stage_end(ST_KEYS, s);
    stage_begin(ST_SORT, s);
    if (sort_pairs_u32(bt, dkey, gv.depth_sorted, iota, gv.depth_perm, P, 32, s)) return 1;
    stage_end(ST_SORT, s);
    stage_begin(ST_KEYS, s);
    uint32_t* tt_sorted = dkey;      // tiles_touched in depth order, then its exclusive scan = emission offsets
    uint32_t* toff = iota;
    k_gather_tiles<<<(P + 255) / 256, 256, 0, s>>>(P, gv.depth_perm, gv.tiles_touched, tt_sorted);
    GSLM_CUDA_OK(cudaGetLastError());
    size_t have = gv.scan_temp_bytes;
    GSLM_CUDA_OK(cub::DeviceScan::ExclusiveSum(gv.scan_temp, have, tt_sorted, toff, P, s));
    if (culled)
        k_duplicate_keys_culled<<<(P + 255) / 256, 256, 0, s>>>(P, gv.depth_perm, gv.means2D, gv.conic_opacity,
                                                                gv.cull_mask, toff,
                                                                tt_sorted, radii, gx, gy, W, H, bt.keys_unsorted,
                                                                bt.vals_unsorted);
    else
        k_duplicate_keys<<<(P + 255) / 256, 256, 0, s>>>(P, gv.depth_perm, gv.means2D, toff, radii, gx, gy,
                                                         bt.keys_unsorted, bt.vals_unsorted);
    stage_end(ST_KEYS, s);
    GSLM_CUDA_OK(cudaGetLastError());
    const int bit = (int)higher_msb((uint32_t)(gx * gy));
    stage_begin(ST_SORT, s);
    if (sort_pairs_u32(bt, bt.keys_unsorted, bt.keys_sorted, bt.vals_unsorted, bt.point_list, R, bit, s)) return 1;
    stage_end(ST_SORT, s);
    const uint32_t L = (uint32_t)R;
    StageScope ts(ST_RANGES, s);
    k_tile_ranges<<<(L + 255) / 256, 256, 0, s>>>(L, bt.keys_sorted, iv.ranges);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------
// compaction (no reference counterpart): dense compact arrays + per-Gaussian CSR
// ---------------------------------------------------------------------------------------------
// exclusive scan of the per-tile compact counts; one CTA (T is a few thousand)
__global__ void __launch_bounds__(1024) k_scan_tiles(int T, const uint32_t* __restrict__ ccount, uint32_t* __restrict__ cstart) {
    __shared__ uint32_t s_w[32];
    __shared__ uint32_t s_carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < T; base += 1024) {
        const int i = base + tid;
        const uint32_t v = i < T ? ccount[i] : 0u;
        uint32_t x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) s_w[warp] = x;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = s_w[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t y = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += y;
            }
            s_w[lane] = w;
        }
        __syncthreads();
        const uint32_t carry = s_carry;
        const uint32_t incl = x + (warp ? s_w[warp - 1] : 0u) + carry;
        if (i < T) cstart[i] = incl - v;
        __syncthreads();
        if (tid == 1023) s_carry = incl;
        __syncthreads();
    }
    if (tid == 0) cstart[T] = s_carry;
}

int run_tile_scan(const ImageView& iv, int W, int H, cudaStream_t s) {
    const int T = ((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
    k_scan_tiles<<<1, 1024, 0, s>>>(T, iv.ccount, iv.cstart);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

// one CTA per tile: tile-local compact records -> dense positions in derivative form (gslm_common.cuh); hit
// masks; ids; per-Gaussian counts
__global__ void __launch_bounds__(256)
k_compact_copy(const uint2* __restrict__ ranges, const uint32_t* __restrict__ cstart,
               const uint32_t* __restrict__ ccount, const float4* __restrict__ src, const uint4* __restrict__ msrc,
               float4* __restrict__ dst, uint4* __restrict__ mdst, uint32_t* __restrict__ cid,
               uint32_t* __restrict__ cpos, uint32_t* __restrict__ g_cnt) {
    const int tile = blockIdx.x;
    const uint32_t n = ccount[tile];
    if (n == 0) return;
    const float L = 1.4426950408889634f;
    const size_t s0 = (size_t)ranges[tile].x * REC_F4, d0 = (size_t)cstart[tile] * REC_F4;
    for (uint32_t i = threadIdx.x; i < n * REC_F4; i += blockDim.x) {
        float4 q = src[s0 + i];
        const uint32_t part = i % REC_F4;
        if (part == 0) {
            q.z *= -0.5f * L; q.w *= -L;
        } else if (part == 1) {
            q.x *= -0.5f * L; q.y = log2f(q.y);
        } else {
            const uint32_t id = __float_as_uint(q.z);
            q.w = src[s0 + i - 1].y;   // opacity
            const uint32_t c = cstart[tile] + i / REC_F4;
            cid[c] = id;
            cpos[c] = c;
            atomicAdd(&g_cnt[id], 1u);
        }
        dst[d0 + i] = q;
    }
    const size_t ms0 = (size_t)ranges[tile].x * 2, md0 = (size_t)cstart[tile] * 2;
    for (uint32_t i = threadIdx.x; i < n * 2; i += blockDim.x) mdst[md0 + i] = msrc[ms0 + i];
}
__global__ void __launch_bounds__(256)
k_invert_glist(uint32_t Rc, const uint32_t* __restrict__ g_list, uint32_t* __restrict__ g_inv) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < Rc) g_inv[g_list[s]] = s;
}

// act_range[r] = lower_bound(active_ids, ACT_RANGE * r): the slice of the active list a warp of the multi-view
// adjoint stage visits in this view
__global__ void __launch_bounds__(256)
k_active_ranges(int nranges, const uint32_t* __restrict__ active_ids, const uint32_t* __restrict__ n_active,
                uint32_t* __restrict__ act_range) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r > nranges) return;
    const uint32_t n = *n_active, key = (uint32_t)r * ACT_RANGE;
    uint32_t lo = 0, hi = n;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (active_ids[mid] < key) lo = mid + 1; else hi = mid;
    }
    act_range[r] = lo;
}

struct HasInstances {   // Gaussian i has at least one compact instance in this view
    const uint32_t* g_off;
    __device__ __forceinline__ bool operator()(uint32_t i) const { return g_off[i + 1] > g_off[i]; }
};

// Launch order of the derivative tile kernels: tiles by descending length class of their compact list (quarter-octave
// classes), so the long lists start first and the short ones fill the tail of the launch (measured: fused tile
// kernel 28.45 -> 27.7 ms per step at config C).  One CTA; the order inside a class is whatever the atomics give --
// it only schedules, every tile still writes its own outputs.
__device__ __forceinline__ int length_class(uint32_t c) {      // monotone in c, < 128
    if (c < 8) return (int)c;
    const int m = 31 - __clz(c);
    return 8 + (m - 3) * 4 + (int)((c >> (m - 2)) & 3u);
}
__global__ void __launch_bounds__(1024) k_tile_order(int T, const uint32_t* __restrict__ ccount, uint32_t* __restrict__ order) {
    // A stable counting sort of the T tiles on 128 length classes with all 32 warps: warp w owns the w-th contiguous
    // chunk of tile ids, counts its tiles per class, the per-class prefix over warps (+ the class base, longest class
    // first) gives every warp its own cursors, and each warp places its chunk 32 tiles per step (__match_any_sync
    // ranks the lanes of one class).  Inside a class the tiles keep their id order, so neighbours of one class (which
    // share Gaussians, i.e. tangent rows in L2) still run together.  (The first version placed all T tiles from ONE
    // warp, 255 dependent steps: 64 us per view, 4 % of a forward pass.)
    __shared__ uint32_t s_cnt[32][128];
    __shared__ uint32_t s_at[128];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    for (int k = tid; k < 32 * 128; k += blockDim.x) (&s_cnt[0][0])[k] = 0u;
    __syncthreads();
    const int chunk = (((T + 31) / 32) + 31) & ~31;            // tiles per warp, a multiple of 32
    const int t_lo = w * chunk, t_hi = min(T, t_lo + chunk);
    for (int t = t_lo + lane; t < t_hi; t += 32) atomicAdd(&s_cnt[w][length_class(ccount[t])], 1u);
    __syncthreads();
    if (tid < 128) {       // exclusive prefix over the warps of class tid; the total lands in s_at
        uint32_t run = 0;
        for (int k = 0; k < 32; k++) { const uint32_t v = s_cnt[k][tid]; s_cnt[k][tid] = run; run += v; }
        s_at[tid] = run;
    }
    __syncthreads();
    if (tid == 0) {        // class bases, longest class first
        uint32_t at = 0;
        for (int c = 127; c >= 0; --c) { const uint32_t n = s_at[c]; s_at[c] = at; at += n; }
    }
    __syncthreads();
    for (int t0 = t_lo; t0 < t_hi; t0 += 32) {
        const int t = t0 + lane;
        const int c = t < t_hi ? length_class(ccount[t]) : 127 + 1 + lane;    // out-of-range lanes: unique fake class
        const uint32_t peers = __match_any_sync(0xffffffffu, c);
        const int leader = __ffs(peers) - 1;
        uint32_t base = 0;
        if (lane == leader && t < t_hi) { base = s_at[c] + s_cnt[w][c]; s_cnt[w][c] += __popc(peers); }
        base = __shfl_sync(0xffffffffu, base, leader);
        if (t < t_hi) order[base + __popc(peers & ((1u << lane) - 1u))] = (uint32_t)t;
        __syncwarp();
    }
}

int run_compaction(const GeomView& gv, const BinningView& bv, const BinningTemp& bt, const ImageView& iv, int P,
                   int W, int H, int64_t Rc, cudaStream_t s) {
    const int T = ((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
    StageScope ts(ST_GATHER, s);
    GSLM_CUDA_OK(cudaMemsetAsync(gv.g_cursor, 0, sizeof(uint32_t) * (size_t)P, s));
    GSLM_CUDA_OK(cudaMemsetAsync(gv.g_off, 0, sizeof(uint32_t), s));
    if (Rc > 0) {
        k_compact_copy<<<T, 256, 0, s>>>(iv.ranges, iv.cstart, iv.ccount, bt.records_tmp, bt.masks_tmp, bv.records, bv.masks,
                                         bv.cid, bt.vals_unsorted, gv.g_cursor);
        GSLM_CUDA_OK(cudaGetLastError());
    }
    // g_off[1..P] = inclusive scan of the counts (the forward pass's scan scratch is free again)
    size_t have = gv.scan_temp_bytes;
    GSLM_CUDA_OK(cub::DeviceScan::InclusiveSum(gv.scan_temp, have, gv.g_cursor, gv.g_off + 1, P, s));
    if (Rc > 0) {
        // CSR payload: compact positions grouped by Gaussian, ascending inside a group = a STABLE sort of the positions
        // by Gaussian id (deterministic; the sort scratch of the forward pass is free again)
        int bits = 1;
        while (bits < 32 && (1u << bits) < (uint32_t)P) bits++;
        if (sort_pairs_u32(bt, bv.cid, bt.keys_unsorted, bt.vals_unsorted, bv.g_list, Rc, bits, s)) return 1;
        k_invert_glist<<<(unsigned)((Rc + 255) / 256), 256, 0, s>>>((uint32_t)Rc, bv.g_list, bv.g_inv);
        GSLM_CUDA_OK(cudaGetLastError());
    }
    // ascending list of the Gaussians that own compact instances (the only ones the per-Gaussian derivative stages
    // have work for: ~10 % of P per view at 1M Gaussians); it takes over the CSR cursors' storage
    GSLM_CUDA_OK(cudaMemsetAsync(gv.n_active, 0, sizeof(uint32_t), s));
    if (Rc > 0) {
        size_t need = 0;
        cub::CountingInputIterator<uint32_t> ids(0u);
        HasInstances pred{gv.g_off};
        GSLM_CUDA_OK(cub::DeviceSelect::If(nullptr, need, ids, gv.active_ids, gv.n_active, P, pred, s));
        if (need > gv.scan_temp_bytes) {
            set_error("select temp storage too small: need %zu have %zu", need, gv.scan_temp_bytes);
            return 1;
        }
        size_t have2 = gv.scan_temp_bytes;
        GSLM_CUDA_OK(cub::DeviceSelect::If(gv.scan_temp, have2, ids, gv.active_ids, gv.n_active, P, pred, s));
    }
    {
        const int nranges = (P + ACT_RANGE - 1) / ACT_RANGE;
        k_active_ranges<<<(nranges + 1 + 255) / 256, 256, 0, s>>>(nranges, gv.active_ids, gv.n_active, gv.act_range);
        GSLM_CUDA_OK(cudaGetLastError());
    }
    k_tile_order<<<1, 1024, 0, s>>>(T, iv.ccount, iv.order);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

}  // namespace gslm
