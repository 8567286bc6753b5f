// Error plumbing and small pass-throughs of the C ABI (include/har_b200.h).
#include "har_common.cuh"

#include <atomic>
#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#include <thread>

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

// ---- per-thread device scratch arena (see har_common.cuh) ------------------------------------------------------
namespace har {
Arena& arena() { static thread_local Arena a; return a; }
ArenaScope::ArenaScope() {
    Arena& a = arena();
    if (a.active) return;                              // nested scope: the outer one owns the arena
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return;
    if (a.base && a.dev != dev) { cudaFree(a.base); a.base = nullptr; a.cap = 0; }
    a.dev = dev; a.off = 0; a.need = 0; a.active = true;
    owner = true;
}
ArenaScope::~ArenaScope() {
    if (!owner) return;
    Arena& a = arena();
    a.active = false;
    if (a.need > a.cap) {                              // grow for the next call; everything of this call has been freed
        if (a.base) cudaFree(a.base);
        a.base = nullptr; a.cap = 0;
        const size_t want = a.need + a.need / 16 + (1 << 20);
        void* q = nullptr;
        if (cudaMalloc(&q, want) == cudaSuccess) { a.base = (char*)q; a.cap = want; }
        else cudaGetLastError();                       // no slab: the stream-ordered allocator keeps being used
    }
}
}  // namespace har

// frees the calling thread's device scratch arena (it is otherwise kept between fits)
HAR_API void har_scratch_release(void) {
    har::Arena& a = har::arena();
    if (a.active) return;
    if (a.base) cudaFree(a.base);
    a.base = nullptr; a.cap = 0;
}

// ---- persistent host thread pool behind parallel_for ---------------------------------------------------
namespace har {
namespace {
struct PoolJob {
    std::function<void(int)> fn;
    int n = 0;
    std::atomic<int> next{0}, done{0};
};
struct HostPool {
    std::mutex mu;
    std::condition_variable cv;
    std::vector<std::thread> workers;
    std::shared_ptr<PoolJob> current;
    long long gen = 0;
    int want = 0;                  // workers 0..want-1 join the current job
    bool stop = false;
    ~HostPool() {
        { std::lock_guard<std::mutex> lk(mu); stop = true; }
        cv.notify_all();
        for (auto& t : workers) t.join();
    }
    void worker(int id) {
        long long seen = 0;
        for (;;) {
            std::shared_ptr<PoolJob> job;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return stop || (gen != seen && id < want); });
                if (stop) return;
                seen = gen;
                job = current;
            }
            if (!job) continue;
            for (int i; (i = job->next.fetch_add(1)) < job->n;) { job->fn(i); job->done.fetch_add(1); }
        }
    }
    void run(int n, int threads, const std::function<void(int)>& fn) {
        auto job = std::make_shared<PoolJob>();
        job->fn = fn; job->n = n;
        {
            std::lock_guard<std::mutex> lk(mu);
            while ((int)workers.size() < threads - 1) { int id = (int)workers.size(); workers.emplace_back([this, id] { worker(id); }); }
            current = job; gen++; want = threads - 1;
        }
        cv.notify_all();
        for (int i; (i = job->next.fetch_add(1)) < n;) { fn(i); job->done.fetch_add(1); }
        while (job->done.load() < n) std::this_thread::yield();
    }
};
HostPool& pool() { static HostPool* p = new HostPool(); return *p; }     // leaked on purpose: no join at process exit
std::atomic<int> g_ranks_on_host{1};
}  // namespace

int host_threads() {
    unsigned hw = std::thread::hardware_concurrency();
    int nt = (int)std::min<unsigned>(hw ? hw : 1, 8);
    const int ranks = std::max(1, g_ranks_on_host.load());
    if (hw) nt = std::min(nt, std::max(1, (int)hw / ranks));
    return nt;
}
void host_threads_set_ranks(int ranks_on_host) { g_ranks_on_host = std::max(1, ranks_on_host); }
void host_pool_run(int n, const std::function<void(int)>& fn) {
    static std::mutex one_job;                   // one fork-join at a time; nested or concurrent callers run inline
    std::unique_lock<std::mutex> lk(one_job, std::try_to_lock);
    if (!lk.owns_lock()) { for (int i = 0; i < n; i++) fn(i); return; }
    pool().run(n, std::min(host_threads(), n), fn);
}
}  // namespace har

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
cudaEvent_t g_cur_start = nullptr, g_span_start = nullptr;
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
void prof_span_begin(cudaStream_t s) {
    if (!g_prof_on) return;
    g_span_start = take_event();
    cudaEventRecord(g_span_start, s);
}
void prof_span_end(int id, cudaStream_t s, double units) {
    if (!g_prof_on || !g_span_start) return;
    cudaEvent_t b = take_event();
    cudaEventRecord(b, s);
    g_pending.push_back({id, g_span_start, b, units});
    g_span_start = nullptr;
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

// ---- shared-memory atomic micro-benchmark (the roofline denominator of k_node_hist) ------------------
// mode 0: every lane of a warp hits its own bank (conflict-free upper bound);
// mode 1: uniformly random counters inside a `words`-sized tile (what a histogram with well-spread bins does).
namespace {
__global__ void __launch_bounds__(256)
k_smem_atomic_peak(int mode, int words, int iters, unsigned* sink) {
    extern __shared__ unsigned sh_ctr[];
    for (int i = threadIdx.x; i < words; i += blockDim.x) sh_ctr[i] = 0;
    __syncthreads();
    unsigned s = (blockIdx.x * 256u + threadIdx.x) * 2654435761u + 12345u;
    const int lane = threadIdx.x & 31;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < 16; k++) {
            s = s * 1664525u + 1013904223u;
            // words is a power of two here: masks instead of modulo keep the loop at ~4 instructions per atomic
            unsigned idx = mode == 0 ? ((((s >> 8) & (unsigned)(words / 32 - 1)) << 5) + lane) : ((s >> 8) & (unsigned)(words - 1));
            atomicAdd(&sh_ctr[idx], 1u);
        }
    }
    __syncthreads();
    unsigned acc = 0;
    for (int i = threadIdx.x; i < words; i += blockDim.x) acc += sh_ctr[i];
    if (acc == 0xffffffffu) sink[0] = acc;
}
}  // namespace

// Returns shared-memory atomics per second (device-wide) measured with CUDA events; words = power of two <= 8192.
HAR_API int har_debug_smem_atomic_peak(int mode, int words, double* atomics_per_s, har_stream_t stream) {
    if (words < 32 || words > 8192 || (words & (words - 1)) || !atomics_per_s)
        return har::fail(HAR_E_INVALID, "har_debug_smem_atomic_peak: words must be a power of two in [32, 8192]");
    cudaStream_t s = (cudaStream_t)stream;
    har::DevBuf<unsigned> sink;
    HAR_CUDA(sink.alloc(1, s));
    const int iters = 2000, blocks = har::kNumSMs * 4 * 4;
    cudaEvent_t a, b;
    HAR_CUDA(cudaEventCreate(&a)); HAR_CUDA(cudaEventCreate(&b));
    HAR_CUDA(cudaFuncSetAttribute(k_smem_atomic_peak, cudaFuncAttributeMaxDynamicSharedMemorySize, 12288 * 4));
    k_smem_atomic_peak<<<blocks, 256, words * 4, s>>>(mode, words, 50, sink.p);      // warm-up
    HAR_CUDA(cudaEventRecord(a, s));
    k_smem_atomic_peak<<<blocks, 256, words * 4, s>>>(mode, words, iters, sink.p);
    HAR_CUDA(cudaEventRecord(b, s));
    HAR_CUDA(cudaEventSynchronize(b));
    float ms = 0;
    HAR_CUDA(cudaEventElapsedTime(&ms, a, b));
    cudaEventDestroy(a); cudaEventDestroy(b);
    *atomics_per_s = (double)blocks * 256.0 * iters * 16.0 / (ms * 1e-3);
    return HAR_OK;
}
