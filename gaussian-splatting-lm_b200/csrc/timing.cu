// =============================================================================
// timing.cu -- optional per-stage device timing behind the C ABI (gslm_timing_enable /
// gslm_timing_collect).  When enabled, every stage launcher brackets its kernel(s) with a
// cudaEvent pair recorded on the launching stream; gslm_timing_collect synchronises, sums
// cudaEventElapsedTime per stage and recycles the events.  Used by bench.py to measure the
// dominant kernel's average launch duration live inside the timed region (roofline.achieved).
// Disabled (the default) it costs one predictable branch per launch.
// =============================================================================
#include "gslm_common.cuh"
#include <mutex>
#include <vector>

namespace gslm {

namespace {
struct Pair { cudaEvent_t a, b; int stage; };
std::mutex g_mu;
bool g_on = false;
std::vector<Pair> g_used;
std::vector<Pair> g_free;
Pair g_open[ST_COUNT];
bool g_is_open[ST_COUNT] = {};
}  // namespace

void stage_begin(int stage, cudaStream_t s) {
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
void stage_end(int stage, cudaStream_t s) {
    if (!g_on) return;
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_is_open[stage]) return;
    cudaEventRecord(g_open[stage].b, s);
    g_used.push_back(g_open[stage]);
    g_is_open[stage] = false;
}

}  // namespace gslm

using namespace gslm;

extern "C" {

static const char* kStageNames[ST_COUNT] = {"preprocess", "scan", "duplicate_keys", "radix_sort", "tile_ranges",
                                            "gather_records", "blend_fwd", "preprocess_tangent", "blend_jvp",
                                            "blend_bwd", "blend_jtjv", "preprocess_bwd", "loss_blocks", "activation_layer"};

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
