// =============================================================================
// binning.cu -- tile binning (a8): scan, key generation, tile|depth radix sort, tile
// ranges; plus the record gather that turns the sorted id list into contiguous 48-byte
// instance records (what the tile kernels stage with one bulk copy per batch).
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
size_t sort_temp_bytes(int64_t R) {
    // cub::DeviceRadixSort (non-DoubleBuffer overload) needs one alternate key and value buffer
    // (12 B / item) plus the onesweep look-back table: 256 counters x 4 B per tile of >= 2048 items
    // (<= 0.5 B / item; measured 0.18 B / item at R = 19.8 M) plus per-pass histograms.  The bound
    // is checked against CUB's own answer at run time.
    return align_up((size_t)R * (sizeof(uint64_t) + sizeof(uint32_t)) + (size_t)R / 2 + ((size_t)1 << 20));
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
    v.ranges = take<uint2>(p, T);
    v.tile_max_contrib = take<uint32_t>(p, T);
    v.total = (size_t)(p - reinterpret_cast<char*>(base));
    return v;
}
BinningView carve_binning(void* base, int64_t R) {
    BinningView v;
    char* p = reinterpret_cast<char*>(base);
    size_t n = (size_t)R;
    v.records = take<float4>(p, REC_F4 * n);
    v.point_list = take<uint32_t>(p, n);
    v.inv_perm = take<uint32_t>(p, n);
    v.total = (size_t)(p - reinterpret_cast<char*>(base));
    return v;
}
BinningTemp carve_binning_temp(void* base, int64_t R) {
    BinningTemp v;
    char* p = reinterpret_cast<char*>(base);
    size_t n = (size_t)R;
    v.keys_sorted = take<uint64_t>(p, n);
    v.keys_unsorted = take<uint64_t>(p, n);
    v.vals_unsorted = take<uint32_t>(p, n);
    v.sort_temp_bytes = sort_temp_bytes(R);
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

// rasterizer_impl.cu:70-111. One thread per Gaussian; the rect is recomputed from the
// stored pixel centre and radius exactly as the reference does (:91).
__global__ void __launch_bounds__(256)
k_duplicate_keys(int P, const float2* __restrict__ xy, const float* __restrict__ depths,
                 const uint32_t* __restrict__ offsets, const int32_t* __restrict__ radii, int gx, int gy,
                 uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    int rad = radii[idx];
    if (rad <= 0) return;
    uint32_t off = idx == 0 ? 0u : offsets[idx - 1];
    float2 p = xy[idx];
    int x0, y0, x1, y1;
    tile_rect(p.x, p.y, rad, gx, gy, x0, y0, x1, y1);
    const uint64_t dbits = (uint64_t)__float_as_uint(depths[idx]);
    for (int y = y0; y < y1; y++)
        for (int x = x0; x < x1; x++) {
            uint64_t key = (uint64_t)(uint32_t)(y * gx + x);
            key <<= 32;
            key |= dbits;
            keys[off] = key;
            vals[off] = (uint32_t)idx;
            off++;
        }
}

// rasterizer_impl.cu:116-138; ranges are zero-filled first (:356)
__global__ void __launch_bounds__(256)
k_tile_ranges(uint32_t L, const uint64_t* __restrict__ keys, uint2* __restrict__ ranges) {
    uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= L) return;
    uint32_t cur = (uint32_t)(keys[idx] >> 32);
    if (idx == 0) ranges[cur].x = 0;
    else {
        uint32_t prev = (uint32_t)(keys[idx - 1] >> 32);
        if (cur != prev) {
            ranges[prev].y = idx;
            ranges[cur].x = idx;
        }
    }
    if (idx == L - 1) ranges[cur].y = L;
}

// Gather: sorted position k -> 48-byte instance record, and the inverse permutation
// (Gaussian-major slot -> k) used by the deterministic gradient gather.
// slot = offsets[id-1] + (ty - rect.y0) * rect_w + (tx - rect.x0): the position at which
// k_duplicate_keys emitted this (tile, Gaussian) pair.
__global__ void __launch_bounds__(256)
k_gather_records(uint32_t L, const uint64_t* __restrict__ keys, const uint32_t* __restrict__ point_list,
                 const float2* __restrict__ xy, const float4* __restrict__ conic_opacity,
                 const float* __restrict__ features, const float* __restrict__ depths,
                 const uint32_t* __restrict__ offsets, const int32_t* __restrict__ radii, int gx, int gy,
                 float4* __restrict__ records, uint32_t* __restrict__ inv_perm) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= L) return;
    uint32_t id = point_list[k];
    uint32_t tile = (uint32_t)(keys[k] >> 32);
    float2 p = __ldg(xy + id);
    float4 co = __ldg(conic_opacity + id);
    float r = __ldg(features + 3 * (size_t)id), g = __ldg(features + 3 * (size_t)id + 1),
          b = __ldg(features + 3 * (size_t)id + 2);
    float invz = 1.f / __ldg(depths + id);   // forward.cu:380: (1 / depths[id])
    float4* rec = records + (size_t)REC_F4 * k;
    rec[0] = make_float4(p.x, p.y, co.x, co.y);
    rec[1] = make_float4(co.z, co.w, r, g);
    rec[2] = make_float4(b, invz, __uint_as_float(id), 0.f);
    int x0, y0, x1, y1;
    tile_rect(p.x, p.y, __ldg(radii + id), gx, gy, x0, y0, x1, y1);
    int tx = (int)(tile % (uint32_t)gx), ty = (int)(tile / (uint32_t)gx);
    uint32_t base = id == 0 ? 0u : __ldg(offsets + id - 1);
    inv_perm[base + (uint32_t)((ty - y0) * (x1 - x0) + (tx - x0))] = k;
}

int run_binning(const GeomView& gv, const int32_t* radii, const BinningView& bv, const BinningTemp& bt,
                const ImageView& iv, int P, int W, int H, int64_t R, const float* features, cudaStream_t s) {
    const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    GSLM_CUDA_OK(cudaMemsetAsync(iv.ranges, 0, sizeof(uint2) * (size_t)gx * gy, s));
    if (R == 0 || P == 0) return 0;
    stage_begin(ST_KEYS, s);
    k_duplicate_keys<<<(P + 255) / 256, 256, 0, s>>>(P, gv.means2D, gv.depths, gv.point_offsets, radii, gx, gy,
                                                     bt.keys_unsorted, bt.vals_unsorted);
    stage_end(ST_KEYS, s);
    GSLM_CUDA_OK(cudaGetLastError());
    const int bit = (int)higher_msb((uint32_t)(gx * gy));
    size_t need = 0;
    GSLM_CUDA_OK(cub::DeviceRadixSort::SortPairs(nullptr, need, bt.keys_unsorted, bt.keys_sorted, bt.vals_unsorted,
                                                 bv.point_list, (int)R, 0, 32 + bit, s));
    if (need > bt.sort_temp_bytes) {
        set_error("sort temp storage too small: need %zu have %zu", need, bt.sort_temp_bytes);
        return 1;
    }
    size_t have = bt.sort_temp_bytes;
    stage_begin(ST_SORT, s);
    GSLM_CUDA_OK(cub::DeviceRadixSort::SortPairs(bt.sort_temp, have, bt.keys_unsorted, bt.keys_sorted,
                                                 bt.vals_unsorted, bv.point_list, (int)R, 0, 32 + bit, s));
    stage_end(ST_SORT, s);
    const uint32_t L = (uint32_t)R;
    stage_begin(ST_RANGES, s);
    k_tile_ranges<<<(L + 255) / 256, 256, 0, s>>>(L, bt.keys_sorted, iv.ranges);
    stage_end(ST_RANGES, s);
    GSLM_CUDA_OK(cudaGetLastError());
    StageScope ts(ST_GATHER, s);
    k_gather_records<<<(L + 255) / 256, 256, 0, s>>>(L, bt.keys_sorted, bv.point_list, gv.means2D, gv.conic_opacity,
                                                     features, gv.depths, gv.point_offsets, radii, gx, gy, bv.records,
                                                     bv.inv_perm);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}

}  // namespace gslm
