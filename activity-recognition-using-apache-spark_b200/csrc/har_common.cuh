// Shared host/device helpers for the sm_100a kernels behind include/har_b200.h.
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <algorithm>
#include <functional>
#include <string>
#include <thread>
#include <vector>

#include "../../include/har_b200.h"

#define HAR_API extern "C" __attribute__((visibility("default")))

namespace har {

constexpr int kNumSMs = 148;          // B200: 2 dies x 74 SMs

std::string& last_error();
int fail(int code, const char* fmt, ...);
void count_launch(int n = 1);
long long launches_since_reset();
void reset_launches();
// int32 sum all-reduce of a device buffer over the ranks of `d` (no-op for one rank); har_comm.cu
int dist_allreduce(const har_dist* d, void* buf, int64_t count, cudaStream_t s);
// true when the group's exchanges run on the library's own NCCL communicator (collective on first use)
bool dist_has_nccl(const har_dist* d, cudaStream_t s);
// node-sharded exchange (NCCL only): sums of int32 runs onto their owners / owners' byte runs to everybody, one group each
int dist_reduce_list(const har_dist* d, int n_ops, int32_t* const* ptr, const int64_t* cnt, const int* root, cudaStream_t s);
int dist_bcast_list(const har_dist* d, int n_ops, void* const* ptr, const int64_t* bytes, const int* root, cudaStream_t s);
int dist_scatter_exchange(const har_dist* d, int32_t* xbuf, int64_t seg, int n_all, int32_t* const* all_ptr,
                          const int64_t* all_cnt, cudaStream_t s);

// optional CUDA-event timing of selected kernels (har_debug_profile_*; off by default)
enum ProfId { PROF_NODE_HIST = 0, PROF_BEST_SPLIT, PROF_PART_COUNT, PROF_PART_SCATTER, PROF_WINDOW_FEATURES,
              PROF_SEG_TILES, PROF_PREDICT, PROF_BIN, PROF_FIND_SPLITS, PROF_BAG, PROF_WINDOW_SPAN, PROF_NUM };
bool prof_enabled();
void prof_begin(cudaStream_t s);
void prof_end(int id, cudaStream_t s, double units);
// a span over several launches (first kernel's start to last kernel's end, gaps included); may enclose prof_begin/prof_end pairs
void prof_span_begin(cudaStream_t s);
void prof_span_end(int id, cudaStream_t s, double units);
void prof_resolve();

#define HAR_CUDA(expr)                                                                         \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess)                                                                 \
            return har::fail(HAR_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                             __FILE__, __LINE__);                                              \
    } while (0)

#define HAR_LAUNCH_CHECK()                                                                     \
    do {                                                                                       \
        har::count_launch();                                                                   \
        cudaError_t _e = cudaGetLastError();                                                   \
        if (_e != cudaSuccess)                                                                 \
            return har::fail(HAR_E_CUDA, "kernel launch failed: %s (%s:%d)",                   \
                             cudaGetErrorString(_e), __FILE__, __LINE__);                      \
    } while (0)

// Keeps freed scratch cached in the device's default memory pool (release threshold = max) so that
// repeated calls do not go back to the driver for every buffer.  Once per device.
void pool_setup();

// Stream-ordered scratch buffer (cudaMallocAsync); freed on the same stream.
// Per-thread device scratch arena.  A fit allocates tens of buffers, some of them gigabytes; through the
// stream-ordered allocator that cost 20 ms per fit at 1e9 samples and, now and then, 100-700 ms (the pool mapping
// fresh memory).  While an ArenaScope is active, DevBuf carves its memory from one slab that is kept between calls
// (bump pointer, nothing is recycled inside a scope); the slab is regrown at the end of a scope that needed more.
// Buffers that outlive the call ask for persistent memory and bypass the arena.  har_scratch_release() frees it.
struct Arena { char* base = nullptr; size_t cap = 0, off = 0, need = 0; int dev = -1; bool active = false; };
Arena& arena();
struct ArenaScope {
    bool owner = false;
    ArenaScope();
    ~ArenaScope();
};

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    cudaStream_t s = nullptr;
    bool from_arena = false;
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    ~DevBuf() { release(); }
    cudaError_t alloc(size_t count, cudaStream_t stream, bool persistent = false) {
        release();
        s = stream;
        n = count;
        if (count == 0) return cudaSuccess;
        Arena& a = arena();
        if (a.active && !persistent) {
            const size_t bytes = (count * sizeof(T) + 255) & ~(size_t)255;
            a.need += bytes;
            if (a.off + bytes <= a.cap) {
                p = reinterpret_cast<T*>(a.base + a.off);
                a.off += bytes;
                from_arena = true;
                return cudaSuccess;
            }
        }
        pool_setup();
        return cudaMallocAsync((void**)&p, count * sizeof(T), stream);
    }
    void release() {
        if (p && !from_arena) cudaFreeAsync(p, s);
        p = nullptr;
        n = 0;
        from_arena = false;
    }
    void swap(DevBuf& o) { std::swap(p, o.p); std::swap(n, o.n); std::swap(s, o.s); std::swap(from_arena, o.from_arena); }
    T* detach() { T* q = p; p = nullptr; n = 0; from_arena = false; return q; }   // caller takes ownership (persistent buffers only)
};

inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// Fork-join helper for the host-side (driver) work; `go` = false runs inline.  The items go to a small persistent
// pool (har_api.cu): creating threads per call cost more than the work of a node group.  The number of threads is
// min(8, cores / ranks sharing the host) — see host_threads_set_ranks.
void host_pool_run(int n, const std::function<void(int)>& fn);
int host_threads();
void host_threads_set_ranks(int ranks_on_host);
template <typename Fn>
void parallel_for(int n, bool go, Fn fn) {
    if (!go || n < 2 || host_threads() <= 1) { for (int i = 0; i < n; i++) fn(i); return; }
    host_pool_run(n, std::function<void(int)>(std::ref(fn)));
}

// java.lang.Double.compare total order as a sortable unsigned key (NaN canonical, greatest).
__host__ __device__ inline uint64_t jd_key(double x) {
    if (x != x) return 0xfff8000000000000ULL;
#ifdef __CUDA_ARCH__
    uint64_t b = (uint64_t)__double_as_longlong(x);
#else
    uint64_t b;
    memcpy(&b, &x, 8);
#endif
    return (b & 0x8000000000000000ULL) ? ~b : (b | 0x8000000000000000ULL);
}
__host__ __device__ inline double jd_unkey(uint64_t k) {
    uint64_t b = (k & 0x8000000000000000ULL) ? (k & 0x7fffffffffffffffULL) : ~k;
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long)b);
#else
    double x;
    memcpy(&x, &b, 8);
    return x;
#endif
}

// java.util.Arrays.binarySearch(double[], key) + `if (idx >= 0) idx else -idx - 1` (TreePoint.findBin)
__host__ __device__ inline int java_find_bin(const double* thr, int n, double key) {
    uint64_t kk = jd_key(key);
    int low = 0, high = n - 1;
    while (low <= high) {
        int mid = (low + high) >> 1;
        uint64_t mk = jd_key(thr[mid]);
        if (mk < kk) low = mid + 1;
        else if (mk > kk) high = mid - 1;
        else return mid;
    }
    return low;
}

}  // namespace har
