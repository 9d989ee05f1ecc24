// =============================================================================
// timing.cu -- optional per-stage device timing behind the C ABI (gslm_timing_enable /
// gslm_timing_collect).  When enabled, every stage launcher brackets its kernel(s) with a
// cudaEvent pair recorded on the launching stream; gslm_timing_collect synchronises, sums
// cudaEventElapsedTime per stage and recycles the events.  Used by bench.py to measure the
// dominant kernel's average launch duration live inside the timed region (roofline.achieved).
// Disabled (the default) it costs one predictable branch per launch.
// =============================================================================
#include "gslm_common.cuh"
#include <cstring>
#include <mutex>
#include <vector>
#include <nvtx3/nvToolsExt.h>   // header-only NVTX v3: a no-op function-pointer call unless a profiler is attached

namespace gslm {

static const char* kStageNames[ST_COUNT] = {"preprocess", "scan", "duplicate_keys", "radix_sort", "tile_ranges",
                                            "gather_records", "blend_fwd", "preprocess_tangent", "blend_jvp",
                                            "blend_bwd", "blend_jtjv", "preprocess_bwd", "loss_blocks", "activation_layer"};

namespace {
struct Pair { cudaEvent_t a, b; int stage; };
std::mutex g_mu;
bool g_on = false;
std::vector<Pair> g_used;
std::vector<Pair> g_free;
Pair g_open[ST_COUNT];
bool g_is_open[ST_COUNT] = {};
}  // namespace

// Every stage is also an NVTX range "gslm/<stage>" on the launching thread ("Tracing": the reference has
// only the cudaEvent pairs of train_jvp.py:89-90,120): nsys / ncu --nvtx --nvtx-include "gslm/blend_jtjv/" select a
// stage by name (tools/ncu_recipes.sh).
static thread_local int t_nvtx_depth = 0;
void stage_begin(int stage, cudaStream_t s) {
    {
        char name[48] = "gslm/";
        strncat(name, kStageNames[stage], sizeof(name) - 6);
        nvtxRangePushA(name);
        t_nvtx_depth++;
    }
    if (!g_on) return;
    std::lock_guard<std::mutex> lk(g_mu);
    Pair p;
    if (!g_free.empty()) { p = g_free.back(); g_free.pop_back(); }
    else { cudaEventCreate(&p.a); cudaEventCreate(&p.b); }
    p.stage = stage;
    cudaEventRecord(p.a, s);
    g_open[stage] = p;
    g_is_open[stage] = true;
}
bool stage_timing_enabled() { return g_on; }
void stage_end(int stage, cudaStream_t s) {
    if (t_nvtx_depth > 0) { nvtxRangePop(); t_nvtx_depth--; }
    if (!g_on) return;
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_is_open[stage]) return;
    cudaEventRecord(g_open[stage].b, s);
    g_used.push_back(g_open[stage]);
    g_is_open[stage] = false;
}

}  // namespace gslm

using namespace gslm;

// ---- FP32 peak probe: the denominator of the tile kernels' roofline, measured in the run that quotes it ----
// mode 0: scalar FFMA, 16 independent chains per thread (the SM's FP32 FLOP peak);
// mode 1: packed FFMA2 with three distinct 64-bit register operands (what the tile sweeps issue);
// mode 2: mode 1 + one MUFU.EX2 per 12 FFMA2 (the sweeps' instruction mix).
template <int MODE>
__global__ void __launch_bounds__(256) k_fp32_probe(float* out, int iters, float s) {
    float2 a[6], b[6], c[6];
#pragma unroll
    for (int i = 0; i < 6; i++) {
        a[i] = make_float2(threadIdx.x * 0.001f + i, i * 0.5f);
        b[i] = make_float2(s, s * 0.999f + i * 1e-4f);
        c[i] = make_float2(0.25f * i, 0.125f);
    }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int i = 0; i < 6; i++) {
                const int j = (i + r) % 6, k = (i + 2 * r + 1) % 6;
                if (MODE == 0) { a[i].x = fmaf(b[j].x, c[k].x, a[i].x); a[i].y = fmaf(b[j].y, c[k].y, a[i].y); }
                else a[i] = __ffma2_rn(b[j], c[k], a[i]);
            }
            if (MODE == 2) {
#pragma unroll
                for (int i = 0; i < 2; i++) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(c[i].x)); c[i].x = y; }
            }
        }
    }
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < 6; i++) acc += a[i].x + a[i].y + b[i].x + c[i].x;
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

extern "C" {

size_t gslm_fp32_probe_scratch_bytes(void) { return (size_t)148 * 8 * 256 * sizeof(float); }

int32_t gslm_fp32_peak_probe(int32_t mode, void* scratch, float* out_tflops, float* out_ms, void* stream) {
    if (!scratch || !out_tflops || mode < 0 || mode > 2) { set_error("fp32 probe: bad argument"); return 1; }
    cudaStream_t s = (cudaStream_t)stream;
    float* d = reinterpret_cast<float*>(scratch);
    const int blocks = 148 * 8, iters = 16384;      // ~9 ms at the FP32 peak: long enough for the clocks to settle
    auto launch = [&](int n) {
        if (mode == 0) k_fp32_probe<0><<<blocks, 256, 0, s>>>(d, n, 0.999f);
        else if (mode == 1) k_fp32_probe<1><<<blocks, 256, 0, s>>>(d, n, 0.999f);
        else k_fp32_probe<2><<<blocks, 256, 0, s>>>(d, n, 0.999f);
    };
    cudaEvent_t e0, e1;
    GSLM_CUDA_OK(cudaEventCreate(&e0));
    GSLM_CUDA_OK(cudaEventCreate(&e1));
    launch(iters / 4);                              // warm-up
    float best = 1e30f;
    for (int rep = 0; rep < 3; rep++) {
        GSLM_CUDA_OK(cudaEventRecord(e0, s));
        launch(iters);
        GSLM_CUDA_OK(cudaEventRecord(e1, s));
        GSLM_CUDA_OK(cudaEventSynchronize(e1));
        float ms = 0.f;
        GSLM_CUDA_OK(cudaEventElapsedTime(&ms, e0, e1));
        best = ms < best ? ms : best;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    const double fma = (double)blocks * 256 * iters * 4 * 12;      // lane-FMAs
    *out_tflops = (float)(2.0 * fma / best / 1e9);
    if (out_ms) *out_ms = best;
    return 0;
}

int32_t gslm_timing_num_stages(void) { return ST_COUNT; }
const char* gslm_timing_stage_name(int32_t stage) { return (stage >= 0 && stage < ST_COUNT) ? kStageNames[stage] : ""; }

void gslm_timing_enable(int32_t on) {
    std::lock_guard<std::mutex> lk(g_mu);
    g_on = on != 0;
}

// Synchronises the device, adds each finished event pair's elapsed time to ms[stage] and counts[stage]
// (both arrays of gslm_timing_num_stages() entries, zeroed by this call first), recycles the events.
int32_t gslm_timing_collect(float* ms, int32_t* counts) {
    if (cudaDeviceSynchronize() != cudaSuccess) { set_error("timing: device synchronize failed"); return 1; }
    std::lock_guard<std::mutex> lk(g_mu);
    for (int i = 0; i < ST_COUNT; i++) { ms[i] = 0.f; counts[i] = 0; }
    for (auto& p : g_used) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, p.a, p.b) == cudaSuccess) { ms[p.stage] += t; counts[p.stage]++; }
        g_free.push_back(p);
    }
    g_used.clear();
    return 0;
}

}  // extern "C"
