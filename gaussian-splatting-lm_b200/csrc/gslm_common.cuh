// =============================================================================
// gslm_common.cuh -- shared declarations of the sm_100a rasterizer library.
//
// Buffer layouts replace GeometryState / ImageState / BinningState of the reference
// (DGR/cuda_rasterizer/rasterizer_impl.h:29-43, rasterizer_impl.cu:155-235); the
// layout is internal, only its total size crosses the C ABI.
// =============================================================================
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/gslm_b200.h"

namespace gslm {

constexpr int TILE_X = 16, TILE_Y = 16;          // DGR/cuda_rasterizer/config.h:17-19
constexpr int TILE_PIX = TILE_X * TILE_Y;        // 256 threads per tile CTA, one pixel each
constexpr int BATCH = 256;                       // instances staged in shared memory per round
constexpr int MAX_BATCH_VIEWS = 16;              // linearised views per batched launch (descriptors travel as kernel parameters)
#ifndef GSLM_ACT_RANGE
#define GSLM_ACT_RANGE 256
#endif
constexpr int ACT_RANGE = GSLM_ACT_RANGE;                   // Gaussian ids per warp in the multi-view adjoint stage
constexpr int REC_F4 = 3;                        // one instance record = 3 x float4 = 48 bytes

// Instance record (one per (tile, Gaussian) pair, stored in tile|depth order so that a
// tile's batch is ONE contiguous block that a single cp.async.bulk can stage):
//   primal form (k_blend_fwd, records_tmp):
//     q0 = (x, y, conic.x, conic.y)   q1 = (conic.z, opacity, r, g)   q2 = (b, 1/depth, id bits, 0)
//   derivative form (compact cache, written by k_compact_copy): exponent in log2 units, opacity folded in
//     q0 = (x, y, a, b)   q1 = (c, log2 o, r, g)   q2 = (b, 1/depth, id bits, o),   (a, b, c) = -log2(e) (A/2, B, C/2)
//   so that alpha_raw = o exp(power) = exp2(a dx^2 + b dx dy + c dy^2 + log2 o).
// The per-instance gradient record is (dmean2D.xy | dconic xx,xy,yy | dopacity | drgb |
// dinvdepth); the per-Gaussian tangent record holds the coefficients of the power tangent as a polynomial in
// (dx, dy): (a1, a2, a3, a4 | a5, opacity_t / opacity, r_t, g_t | b_t, invdepth_t, 0, 0), see preprocess.cu.
//
// COMPACT CACHE.  Only instances that at least one pixel of their tile actually BLENDED survive
// the forward pass (k_blend_fwd flags them): instances behind the point where every pixel of
// the tile has terminated, and instances of the 3-sigma bounding square that no pixel sees with
// alpha >= 1/255, carry exactly zero tangent / gradient and are dropped.  "Rc" = number of
// compact instances of a view (about 5 % of R at 1M Gaussians / 1080p); every derivative
// kernel works on the compact arrays only.

inline __host__ __device__ size_t align_up(size_t x, size_t a = 128) { return (x + a - 1) / a * a; }

struct GeomView {   // per-Gaussian state, SoA
    float* depths;            // [P]
    float2* means2D;          // [P]
    float4* conic_opacity;    // [P]
    float* rgb;               // [3P]
    float* cov3D;             // [6P]
    uint8_t* clamped;         // [P] bit c = channel c clamped (forward.cu:66-72)
    uint32_t* tiles_touched;  // [P]
    uint2* cull_mask;         // [P] tile_culling: survivors among the first 64 tiles of the rect (row-major bits)
    uint32_t* point_offsets;  // [P] inclusive scan
    uint32_t* num_rendered;   // [1] device copy of R (tail of the scan)
    uint32_t* g_off;          // [P+1] CSR offsets: compact instances of Gaussian i are g_list[g_off[i] .. g_off[i+1])
    uint32_t* g_cursor;       // [P]   build-time counters / cursors of the CSR
    uint32_t* active_ids;     // [n_active <= P] ascending ids of the Gaussians with >= 1 compact instance (aliases g_cursor)
    uint32_t* n_active;       // [1]
    uint32_t* depth_sorted;   // [P] binning scratch: depth keys in ascending order
    uint32_t* depth_perm;     // [P] binning scratch: Gaussian ids in (depth, id) order (invisible ones last)
    uint32_t* act_range;      // [ceil(P / ACT_RANGE) + 1] act_range[r] = first slot of active_ids with id >= ACT_RANGE r
    char* scan_temp;
    size_t scan_temp_bytes;
    size_t total;
};
struct ImageView {  // per-pixel / per-tile state
    float* final_T;           // [N]
    uint32_t* n_contrib;      // [N] last blended position in the FULL tile list (reference numbering, ImageState::n_contrib)
    uint32_t* n_contrib_c;    // [N] the same in the COMPACT tile list (what the derivative sweeps use)
    uint2* ranges;            // [T] full-list ranges (ImageState::ranges)
    uint32_t* ccount;         // [T] compact instances per tile
    uint32_t* cstart;         // [T+1] exclusive scan of ccount; cstart[T] = Rc
    uint32_t* order;          // [T] tile ids, longest compact lists first: the launch order of the derivative tile kernels
    size_t total;
};
struct BinningView {  // compact per-instance state that outlives the forward pass (92 B / compact instance)
    float4* records;          // [3 Rc] compact records in derivative form (see below), tile|depth order
    uint4* masks;             // [2 Rc] 256-bit hit mask: bit l of word k <=> pixel (row 2k + (l>>4), column l&15) of the tile blended it
    uint32_t* cid;            // [Rc] Gaussian id of each compact instance
    uint32_t* g_list;         // [Rc] CSR payload: compact positions grouped by Gaussian (ascending within a Gaussian)
    uint32_t* g_inv;          // [Rc] inverse of g_list: compact position -> CSR slot (where the tile kernels put its gradient record)
    size_t total;
};
struct BinningTemp {  // per-instance state that is dead once the forward pass returns (96 B / instance + sort temporaries)
    uint32_t* keys_sorted;    // [R] tile ids in sorted order
    uint32_t* keys_unsorted;  // [R] tile ids in emission order
    uint32_t* vals_unsorted;  // [R] Gaussian ids in emission order
    uint32_t* point_list;     // [R] sorted Gaussian ids (BinningState::point_list)
    float4* records_tmp;      // [3R] compact records at tile-local positions (range.x + j), before the global scan
    uint4* masks_tmp;         // [2R] their hit masks, same positions
    char* sort_temp;
    size_t sort_temp_bytes;
    size_t total;
};

size_t scan_temp_bytes(int P);
size_t sort_temp_bytes(int64_t n);
GeomView carve_geom(void* base, int P);
ImageView carve_image(void* base, int W, int H);
BinningView carve_binning(void* base, int64_t Rc);
BinningTemp carve_binning_temp(void* base, int64_t R, int P);

void set_error(const char* fmt, ...);
const char* last_error();

#define GSLM_CUDA_OK(expr)                                                                   \
    do {                                                                                     \
        cudaError_t e__ = (expr);                                                            \
        if (e__ != cudaSuccess) {                                                            \
            gslm::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
            return 1;                                                                        \
        }                                                                                    \
    } while (0)

// ---- optional per-stage device timing (gslm_timing_*; used by bench.py for the live roofline numbers) ----
enum Stage { ST_PREPROCESS = 0, ST_SCAN, ST_KEYS, ST_SORT, ST_RANGES, ST_GATHER, ST_BLEND_FWD, ST_PRE_TANGENT,
             ST_BLEND_JVP, ST_BLEND_BWD, ST_BLEND_JTJV, ST_PRE_BWD, ST_LOSS, ST_ACTIVATION, ST_COUNT };
void stage_begin(int stage, cudaStream_t s);   // no-ops unless timing is enabled
void stage_end(int stage, cudaStream_t s);
bool stage_timing_enabled();
struct StageScope {
    int st; cudaStream_t s;
    StageScope(int st_, cudaStream_t s_) : st(st_), s(s_) { stage_begin(st, s); }
    ~StageScope() { stage_end(st, s); }
};

// SH rows [M,3] of 12 M bytes are 16-byte aligned when M % 4 == 0 (M = 4, 16): element f of a row through a float4
// load.  With a compile-time f the compiler keeps one LDG.128 per four coefficients floats (identical read-only loads
// are merged) instead of four LDG.32 -- the per-Gaussian kernels are latency-bound gathers, bytes per instruction count.
#ifdef __CUDACC__
__device__ __forceinline__ float sh_elem4(const float* row, int f) {
    const float4 q = __ldg(reinterpret_cast<const float4*>(row) + (f >> 2));
    const int c = f & 3;
    return c == 0 ? q.x : (c == 1 ? q.y : (c == 2 ? q.z : q.w));
}
__host__ __device__ inline bool sh_rows_aligned(const float* shs, int M) {
    return shs != nullptr && (M & 3) == 0 && (reinterpret_cast<size_t>(shs) & 15u) == 0;
}
#endif

// ---- stage launchers (each one = one reference stage; see the .cu that defines it) ----
struct TangentPtrs {   // tangent inputs of the forward-mode preprocess; any pointer may be null
    const float *means3D, *opacities, *scales, *rotations, *cov3D_precomp, *shs, *colors_precomp;
    const float *viewmatrix, *projmatrix, *campos;
    float tan_fovx, tan_fovy, scale_modifier;
};

// preprocess.cu
int launch_preprocess(const GslmGaussians& g, const GslmCamera& cam, const GeomView& gv, int32_t* radii,
                      cudaStream_t s);
// writes the per-Gaussian tangent record tgeom[3 id] of every ACTIVE Gaussian (the others are never read)
int launch_preprocess_tangent(const GslmGaussians& g, const TangentPtrs& t, const GslmCamera& cam, const GeomView& gv,
                              float4* tgeom, cudaStream_t s);
// the same for up to MAX_BATCH_VIEWS linearised views in ONE launch (gvs / tgeoms: host arrays of n_views entries)
int launch_preprocess_tangent_batch(const GslmGaussians& g, const TangentPtrs& t, int n_views, const GslmCamera* cams,
                                    const GeomView* gvs, float4* const* tgeoms, cudaStream_t s);
int launch_mark_visible(int P, const float* means3D, const float* viewmatrix, uint8_t* present, cudaStream_t s);
int launch_preprocess_bwd(const GslmGaussians& g, const GslmCamera& cam, const int32_t* radii, const GeomView& gv,
                          const BinningView& bv, const float4* inst_grads, bool has_invdepth_grad,
                          const GslmGrads& grads, cudaStream_t s);
// accumulates the adjoints of up to MAX_BATCH_VIEWS views in ONE launch (deterministic, see preprocess_bwd.cu)
// Gaussian ranges [range_lo, range_hi) in units of ACT_RANGE ids (range_hi < 0: all).  Always accumulates: a caller
// with overwrite semantics clears the sinks first (cudaMemsetAsync; zero-filling inside this launch measured 0.06 ms
// slower per product at P = 1M)
int launch_preprocess_bwd_views(const GslmGaussians& g, int n_views, const GslmCamera* cams, const int32_t* const* radii,
                                const GeomView* gvs, const float4* const* inst_grads, const GslmGrads& grads,
                                cudaStream_t s, int range_lo = 0, int range_hi = -1);
// binning.cu
int run_scan(const GeomView& gv, int P, cudaStream_t s);
int run_binning(const GeomView& gv, const int32_t* radii, const BinningTemp& bt, const ImageView& iv, int P, int W,
                int H, int64_t R, bool culled, cudaStream_t s);
// exact tile culling (GslmCamera::tile_culling): rewrites tiles_touched before the scan; run_binning(culled = true)
// then emits only the surviving (tile, Gaussian) instances
int run_tile_cull(const GeomView& gv, const int32_t* radii, int P, int W, int H, cudaStream_t s);
// after k_blend_fwd: cstart = exclusive scan of ccount (device only)
int run_tile_scan(const ImageView& iv, int W, int H, cudaStream_t s);
// tile-local compact records -> dense compact arrays + per-Gaussian CSR
int run_compaction(const GeomView& gv, const BinningView& bv, const BinningTemp& bt, const ImageView& iv, int P,
                   int W, int H, int64_t Rc, cudaStream_t s);
// blend.cu
int launch_blend_fwd(const GeomView& gv, const BinningTemp& bt, const ImageView& iv, int W, int H,
                     const float* features, const float* bg, float* out_color, float* out_invdepth, bool record,
                     cudaStream_t s);
int launch_blend_jvp(const BinningView& bv, const ImageView& iv, int W, int H, const float4* tgeom, const float* bg,
                     const float* bg_tangent, float* out_color_t, float* out_invdepth_t, cudaStream_t s);
int launch_blend_bwd(const BinningView& bv, const ImageView& iv, int W, int H, const float* bg,
                     const float* dL_dcolor, const float* dL_dinvdepth, float4* inst_grads, cudaStream_t s);
int launch_blend_jtjv(const BinningView& bv, const ImageView& iv, int W, int H, const float* bg,
                      const float4* tgeom, const float* weight, float4* inst_grads, float* out_vJJv, cudaStream_t s);
int launch_blend_jvp_batch(int n_views, const BinningView* bvs, const ImageView* ivs, const GslmCamera* cams,
                           const float4* const* tgeoms, float* const* out_color_t, cudaStream_t s);
int launch_blend_bwd_batch(int n_views, const BinningView* bvs, const ImageView* ivs, const GslmCamera* cams,
                           const float* const* dL_dcolor, float4* const* inst_grads, cudaStream_t s);
// up to MAX_BATCH_VIEWS views in one launch; weights may be null (all ones) or hold null entries
int launch_blend_jtjv_batch(int n_views, const BinningView* bvs, const ImageView* ivs, const GslmCamera* cams,
                            const float4* const* tgeoms, const float* const* weights, float4* const* inst_grads,
                            float* out_vJJv, cudaStream_t s);

}  // namespace gslm
