// capi.cu -- error reporting and device probing for libgs2d.so.
#include <atomic>
#include <cstring>
#include <mutex>
#include <vector>

#include <cstdlib>
#include <nvtx3/nvToolsExt.h>

#include "gs2d_host.h"

namespace gs2d {

// ---- NVTX ranges per stage (GS2D_NVTX=1) ------------------------------------------------------
static const char* const kStageNames[GS2D_NUM_STAGES] = {
    "gs2d:project_fwd", "gs2d:memset", "gs2d:scan_map", "gs2d:sort", "gs2d:edges_pack", "gs2d:raster_fwd",
    "gs2d:raster_bwd", "gs2d:project_bwd", "gs2d:fit_tiles", "gs2d:adan", "gs2d:other"};
bool nvtx_enabled() {
    static const bool on = [] { const char* e = getenv("GS2D_NVTX"); return e && e[0] == '1'; }();
    return on;
}
void nvtx_push(int stage) { nvtxRangePushA(stage >= 0 && stage < GS2D_NUM_STAGES ? kStageNames[stage] : "gs2d"); }
void nvtx_pop() { nvtxRangePop(); }

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// ---- launch counter + per-stage event timing ------------------------------------------------
static std::atomic<long long> g_launches{0};
static std::atomic<bool> g_prof_on{false};
static std::mutex g_prof_mu;
struct ProfPair { cudaEvent_t a, b; int stage; };
static std::vector<ProfPair> g_pairs;        // recorded, not yet read
static std::vector<ProfPair> g_free;         // event pool
static std::vector<ProfPair> g_open;         // begin recorded, end pending (per stage nesting is not used)

void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
bool prof_enabled() { return g_prof_on.load(std::memory_order_relaxed); }

void prof_begin(int stage, cudaStream_t s) {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    ProfPair p;
    if (!g_free.empty()) { p = g_free.back(); g_free.pop_back(); }
    else { cudaEventCreate(&p.a); cudaEventCreate(&p.b); }
    p.stage = stage;
    cudaEventRecord(p.a, s);
    g_open.push_back(p);
}

void prof_end(int stage, cudaStream_t s) {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    for (size_t i = g_open.size(); i-- > 0;) {
        if (g_open[i].stage == stage) {
            ProfPair p = g_open[i];
            g_open.erase(g_open.begin() + (long)i);
            cudaEventRecord(p.b, s);
            g_pairs.push_back(p);
            return;
        }
    }
}

}  // namespace gs2d

extern "C" {

int gs2d_version(void) { return 101; }

long long gs2d_launch_count(void) { return gs2d::g_launches.load(); }

int gs2d_prof_enable(int on) {
    gs2d::g_prof_on.store(on != 0);
    return GS2D_OK;
}

int gs2d_prof_read(double* total_ms, int* spans, int num_stages) {
    using namespace gs2d;
    std::lock_guard<std::mutex> lk(g_prof_mu);
    for (int i = 0; i < num_stages; ++i) { total_ms[i] = 0.0; spans[i] = 0; }
    for (const ProfPair& p : g_pairs) {
        if (cudaEventSynchronize(p.b) != cudaSuccess) {
            set_error("prof_read: %s", cudaGetErrorString(cudaGetLastError()));
            return GS2D_ERR_CUDA;
        }
        float ms = 0.f;
        cudaEventElapsedTime(&ms, p.a, p.b);
        if (p.stage >= 0 && p.stage < num_stages) { total_ms[p.stage] += ms; spans[p.stage] += 1; }
        g_free.push_back(p);
    }
    g_pairs.clear();
    return GS2D_OK;
}

const char* gs2d_last_error(void) { return gs2d::g_err; }

int gs2d_device_arch(int device) {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, device) != cudaSuccess) {
        (void)cudaGetLastError();
        return -1;
    }
    return p.major * 10 + p.minor;
}

}  // extern "C"
