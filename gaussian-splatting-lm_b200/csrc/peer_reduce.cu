// =============================================================================
// peer_reduce.cu -- the cross-GPU sum of the per-rank J^T J v partials over NVLink / NVSwitch, written against
// peer memory instead of calling a collective library (the reference has no multi-GPU path, F5).
//
// The activated-space accumulator of every rank lives in ONE symmetric allocation that is also mapped through an
// NVSwitch MULTICAST address (the caller sets it up: torch.distributed._symmetric_memory).  After a cross-rank barrier
// rank r owns the r-th slice of the 59 P floats:
//     multimem.ld_reduce.add.v4.f32   one load through the multicast address returns the SUM over all ranks -- the
//                                     switch reads the eight copies and adds them in the fabric (NVLS);
//     multimem.st.v4.f32              one store through the multicast address writes the sum into every rank's copy.
// Per GPU and direction the links carry (N-1)/N of the payload once (a ring or a two-shot over unicast peer pointers
// carries it twice), and no SM does any adding.  The kernel contains no cross-GPU synchronisation of its own: the
// caller brackets it with the symmetric-memory barrier on the same stream.
// =============================================================================
#include "gslm_common.cuh"
#include <cstdlib>

namespace gslm {
namespace {

__device__ __forceinline__ float4 mm_ld_reduce(float* p) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p)
                 : "memory");
    return v;
}
__device__ __forceinline__ void mm_st(float* p, const float4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                 :
                 : "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

// UNROLL independent load-reduces in flight per thread before the first store: the round trip through the switch is
// microseconds, the loads of one thread must not serialise behind their stores
template <int UNROLL>
__global__ void __launch_bounds__(512) k_multimem_allreduce(float* mc, size_t lo4, size_t hi4) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = lo4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (UNROLL - 1) * stride < hi4; i += UNROLL * stride) {
        float4 v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; u++) v[u] = mm_ld_reduce(mc + 4 * (i + u * stride));
#pragma unroll
        for (int u = 0; u < UNROLL; u++) mm_st(mc + 4 * (i + u * stride), v[u]);
    }
    for (; i < hi4; i += stride) mm_st(mc + 4 * i, mm_ld_reduce(mc + 4 * i));
}

}  // namespace
}  // namespace gslm

using namespace gslm;

extern "C" int32_t gslm_multimem_allreduce_f32(void* multicast_ptr, size_t n_floats, int32_t rank, int32_t world,
                                               void* stream) {
    if (!multicast_ptr || world < 1 || rank < 0 || rank >= world || (n_floats & 3u) ||
        (reinterpret_cast<size_t>(multicast_ptr) & 15u)) {
        set_error("gslm_multimem_allreduce_f32: bad argument (the buffer must be 16-byte aligned, n a multiple of 4)");
        return 1;
    }
    const size_t n4 = n_floats >> 2;
    const size_t lo = n4 * (size_t)rank / (size_t)world, hi = n4 * (size_t)(rank + 1) / (size_t)world;
    if (hi <= lo) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    // enough requests in flight to cover the switch round trip: 4 CTAs of 512 threads per SM, 16 B each
    static int ctas_per_sm = 0, unroll = 0;
    if (!ctas_per_sm) {      // tuning knobs for tools/bench_peer.py
        const char* e = getenv("GSLM_MM_CTAS");
        ctas_per_sm = e ? atoi(e) : 4;
        e = getenv("GSLM_MM_UNROLL");
        unroll = e ? atoi(e) : 1;    // measured on 8 x B200 (profiles/r2l_peer_reduce.txt): 0.537 ms at 1, 0.573 at 4
    }
    const size_t want = (hi - lo + 511) / 512;
    const size_t cap = (size_t)148 * (size_t)(ctas_per_sm > 0 ? ctas_per_sm : 4);
    const int grid = (int)(want < cap ? want : cap);
    float* mc = reinterpret_cast<float*>(multicast_ptr);
    if (unroll >= 8) k_multimem_allreduce<8><<<grid, 512, 0, s>>>(mc, lo, hi);
    else if (unroll >= 4) k_multimem_allreduce<4><<<grid, 512, 0, s>>>(mc, lo, hi);
    else if (unroll >= 2) k_multimem_allreduce<2><<<grid, 512, 0, s>>>(mc, lo, hi);
    else k_multimem_allreduce<1><<<grid, 512, 0, s>>>(mc, lo, hi);
    GSLM_CUDA_OK(cudaGetLastError());
    return 0;
}
