// Error plumbing and small pass-throughs of the C ABI (include/har_b200.h).
#include "har_common.cuh"

#include <atomic>

namespace har {

std::string& last_error() {
    static thread_local std::string e;
    return e;
}

int fail(int code, const char* fmt, ...) {
    char buf[2048];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    last_error() = buf;
    return code;
}

void pool_setup() {
    static bool done[64] = {false};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64 || done[dev]) return;
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
        unsigned long long thr = ~0ULL;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
    }
    done[dev] = true;
}

static std::atomic<long long> g_launches{0};
void count_launch(int n) { g_launches += n; }
long long launches_since_reset() { return g_launches.load(); }
void reset_launches() { g_launches = 0; }

}  // namespace har

HAR_API const char* har_last_error(void) { return har::last_error().c_str(); }
HAR_API int har_version(void) { return 100; }

HAR_API int har_device_count(int32_t* count) {
    int c = 0;
    HAR_CUDA(cudaGetDeviceCount(&c));
    *count = c;
    return HAR_OK;
}

HAR_API int har_set_device(int32_t device) {
    HAR_CUDA(cudaSetDevice(device));
    return HAR_OK;
}

HAR_API int har_stream_synchronize(har_stream_t stream) {
    HAR_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return HAR_OK;
}

// Kernel-launch counter for bench.py's "gpu_launches" (not part of the reference-facing surface).
HAR_API long long har_debug_launch_count(int reset) {
    long long v = har::launches_since_reset();
    if (reset) har::reset_launches();
    return v;
}

// ---- optional per-kernel timing (bench.py's live roofline): CUDA events around selected launches ----
namespace har {
namespace {
struct ProfPending { int id; cudaEvent_t a, b; double units; };
bool g_prof_on = false;
std::vector<ProfPending> g_pending;
std::vector<cudaEvent_t> g_free_events;
struct ProfAcc { double ms = 0, units = 0; long long launches = 0; } g_acc[PROF_NUM];
cudaEvent_t g_cur_start = nullptr;
cudaEvent_t take_event() {
    if (!g_free_events.empty()) { cudaEvent_t e = g_free_events.back(); g_free_events.pop_back(); return e; }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
}
}  // namespace
bool prof_enabled() { return g_prof_on; }
void prof_begin(cudaStream_t s) {
    if (!g_prof_on) return;
    g_cur_start = take_event();
    cudaEventRecord(g_cur_start, s);
}
void prof_end(int id, cudaStream_t s, double units) {
    if (!g_prof_on || !g_cur_start) return;
    cudaEvent_t b = take_event();
    cudaEventRecord(b, s);
    g_pending.push_back({id, g_cur_start, b, units});
    g_cur_start = nullptr;
}
void prof_resolve() {
    for (auto& p : g_pending) {
        float ms = 0;
        if (cudaEventSynchronize(p.b) == cudaSuccess && cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) {
            g_acc[p.id].ms += ms; g_acc[p.id].units += p.units; g_acc[p.id].launches++;
        }
        g_free_events.push_back(p.a); g_free_events.push_back(p.b);
    }
    g_pending.clear();
}
}  // namespace har

HAR_API void har_debug_profile_enable(int on) { har::g_prof_on = on != 0; }
HAR_API void har_debug_profile_reset(void) {
    har::prof_resolve();
    for (auto& a : har::g_acc) a = har::ProfAcc();
}
HAR_API int har_debug_profile_read(int id, double* ms, long long* launches, double* units) {
    if (id < 0 || id >= har::PROF_NUM) return HAR_E_INVALID;
    har::prof_resolve();
    if (ms) *ms = har::g_acc[id].ms;
    if (launches) *launches = har::g_acc[id].launches;
    if (units) *units = har::g_acc[id].units;
    return HAR_OK;
}
