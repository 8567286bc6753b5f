// WISDM window transform on sm_100a: segmentation (K1) and per-window features (K2/K3).
//
// No reference implementation exists for this stage: the reference ships only
// the transformed table (Main/wisdm_main_ver_0.0/data/wisdm_data.csv, schema line 1) and reads it at
// Main/main.py:18-20.  Definitions follow SURVEY Appendix B as fixed in DESIGN.md; oracle/har_oracle.c
// (orc_window_segments / orc_window_features) is the arbiter.
//
// K1  k_seg_pass (har_window_v4.cuh): one launch, the stream is read ONCE (8 B/sample, +8 with gap_ns); tiles exchange
//     their running (last boundary, window count) by decoupled look-back and write the windows in place.
// K2/K3 k_window_features_v4 (window length 200, har_window_v4.cuh): x/y/z slices staged into shared memory with
//     cp.async.bulk (TMA 1-D) behind an mbarrier, two items ahead of the arithmetic; 8 lanes per window, 25 consecutive
//     samples per lane.  Other window lengths (3..256) run k_window_features_v3 (har_window_v3.cuh: 8 lanes per window,
//     samples interleaved over the lanes, straight from global memory).
#include "har_common.cuh"

#include <algorithm>
#include <cstdlib>

namespace {

constexpr int SEG_THREADS = 256;
constexpr int SEG_PER_THREAD = 8;
constexpr int SEG_TILE = SEG_THREADS * SEG_PER_THREAD;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// TMA 1-D bulk copy global -> shared, completion counted on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ double shfl_xor_d(double v, int m) { return __shfl_xor_sync(0xffffffffu, v, m); }
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1) v += shfl_xor_d(v, m);
    return v;
}
__device__ __forceinline__ double warp_min_d(double v) {
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1) v = fmin(v, shfl_xor_d(v, m));
    return v;
}
__device__ __forceinline__ double warp_max_d(double v) {
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1) v = fmax(v, shfl_xor_d(v, m));
    return v;
}
__device__ __forceinline__ unsigned long long warp_sum_u64(unsigned long long v) {
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
    return v;
}

#include "har_window_v3.cuh"
#include "har_window_v4.cuh"

}  // namespace

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
namespace {

// scratch of one segmentation call: [0, 256) the window total, then the hand-over words of the scan (two per slab of
// 8192 tiles, at least 256 bytes), then the per-tile arrays of SegScratch
size_t seg_header_bytes(int64_t num_tiles) {
    const size_t chain = (size_t)har::ceil_div(std::max<int64_t>(num_tiles, 1), SEG_SCAN_SLAB) * 16;
    return 256 + std::max<size_t>(256, (chain + 255) & ~(size_t)255);
}
size_t seg_scratch_bytes(int64_t num_tiles, int cap) {
    return seg_header_bytes(num_tiles) + (size_t)num_tiles * seg_scratch_bytes_per_tile(cap);
}

// Launches the three segmentation kernels; *d_total (device) receives the window count.  Nothing is read back.
// Of the scratch only the scan's hand-over words must start at zero (cleared here, on the stream); every other entry a
// kernel reads is written earlier in the same call.
int seg_launch(const int32_t* d_user, const int32_t* d_activity, const int64_t* d_ts, int64_t n, int32_t window_len,
               int64_t gap_ns, unsigned char* scratch, int64_t* d_win_start, int32_t* d_win_user, int32_t* d_win_activity,
               int64_t** d_total, cudaStream_t s, cudaEvent_t after_scan = nullptr) {
    const int64_t num_tiles = har::ceil_div(n, SEG_TILE);
    const int32_t cap = SEG_TILE / window_len + 1;
    SegScratch sc;
    unsigned char* p = scratch + seg_header_bytes(num_tiles);
    sc.total = reinterpret_cast<int64_t*>(scratch);
    sc.chain = reinterpret_cast<unsigned long long*>(scratch + 256);
    HAR_CUDA(cudaMemsetAsync(scratch + 256, 0, seg_header_bytes(num_tiles) - 256, s));
    sc.summary = reinterpret_cast<unsigned long long*>(p); p += (size_t)num_tiles * 8;
    sc.open_bout = reinterpret_cast<int64_t*>(p); p += (size_t)num_tiles * 8;
    sc.out_base = reinterpret_cast<int64_t*>(p); p += (size_t)num_tiles * 8;
    sc.npre = reinterpret_cast<int32_t*>(p); p += (size_t)num_tiles * 4;
    sc.local = reinterpret_cast<uint16_t*>(p);
    *d_total = sc.total;
    const int vec_ok = ((reinterpret_cast<uintptr_t>(d_user) | reinterpret_cast<uintptr_t>(d_activity)) & 15) == 0;
    har::prof_begin(s);
    const unsigned magic = (unsigned)((1ULL << 32) / (unsigned)window_len) + 1u;
    k_seg_tiles<<<(unsigned)num_tiles, SEG_THREADS, 0, s>>>(d_user, d_activity, d_ts, n, window_len, magic, gap_ns, vec_ok, sc, cap);
    HAR_LAUNCH_CHECK();
    har::prof_end(har::PROF_SEG_TILES, s, (double)n);
    {
        static bool attr_set[64] = {false};
        int dev = 0;
        HAR_CUDA(cudaGetDevice(&dev));
        const size_t smem = (size_t)SEG_SCAN_PAD * 20;
        if (dev >= 0 && dev < 64 && !attr_set[dev]) {
            HAR_CUDA(cudaFuncSetAttribute(k_seg_scan, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr_set[dev] = true;                                   // idempotent
        }
        k_seg_scan<<<(unsigned)har::ceil_div(num_tiles, SEG_SCAN_SLAB), SEG_SCAN_THREADS, smem, s>>>(sc, num_tiles, n, window_len, magic);
    }
    HAR_LAUNCH_CHECK();
    if (after_scan) HAR_CUDA(cudaEventRecord(after_scan, s));
    k_seg_emit<<<(unsigned)har::ceil_div(num_tiles * 8, 256), 256, 0, s>>>(sc, cap, num_tiles, window_len, d_user, d_activity,
                                                                          d_win_start, d_win_user, d_win_activity);
    HAR_LAUNCH_CHECK();
    return HAR_OK;
}

int seg_check_args(const char* who, int64_t n, int32_t window_len, int64_t gap_ns, const int64_t* d_ts, int64_t* num_windows) {
    if (!num_windows) return har::fail(HAR_E_INVALID, "%s: num_windows is NULL", who);
    *num_windows = 0;
    if (n < 0 || window_len < 2) return har::fail(HAR_E_INVALID, "%s: need n >= 0 and window_len >= 2", who);
    if (window_len > (1 << 20)) return har::fail(HAR_E_UNSUPPORTED, "%s: window_len > 2^20", who);
    if (gap_ns > 0 && !d_ts) return har::fail(HAR_E_INVALID, "%s: gap_ns > 0 needs d_ts", who);
    return HAR_OK;
}

// launches the feature kernel; d_nw (device) overrides num_windows when non-NULL (window length 200 only)
int features_launch(const double* d_x, const double* d_y, const double* d_z, const int64_t* d_ts, int64_t n,
                    const int64_t* d_win_start, int64_t num_windows, const int64_t* d_nw, int32_t window_len,
                    double* d_out, int64_t ld_out, cudaStream_t s, const int32_t* d_user = nullptr,
                    const int32_t* d_act = nullptr, int32_t* d_win_user = nullptr, int32_t* d_win_act = nullptr) {
    if (window_len == W4_L) {
        static bool attr_set[64] = {false};
        int dev = 0;
        HAR_CUDA(cudaGetDevice(&dev));
        const size_t smem = W4_SMEM_PER_WARP * W4_WARPS;
        if (dev >= 0 && dev < 64 && !attr_set[dev]) {
            HAR_CUDA(cudaFuncSetAttribute(k_window_features_v4, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr_set[dev] = true;                                   // idempotent: a racing second writer stores the same value
        }
        // bulk copies need 16-byte aligned columns (window starts are re-aligned to an even sample inside the kernel)
        const int tma_ok = ((reinterpret_cast<uintptr_t>(d_x) | reinterpret_cast<uintptr_t>(d_y) |
                             reinterpret_cast<uintptr_t>(d_z)) & 15) == 0;
        const int64_t quads = har::ceil_div(num_windows, 4);
        const unsigned blocks = (unsigned)std::max<int64_t>(1, std::min<int64_t>(har::ceil_div(quads, W4_WARPS), (int64_t)har::kNumSMs * 3));
        har::prof_begin(s);
        k_window_features_v4<<<blocks, W4_WARPS * 32, smem, s>>>(d_x, d_y, d_z, d_ts, d_win_start, num_windows, d_nw, n, tma_ok,
                                                                 d_out, ld_out, d_user, d_act, d_win_user, d_win_act);
        HAR_LAUNCH_CHECK();
        har::prof_end(har::PROF_WINDOW_FEATURES, s, (double)num_windows * window_len);
        return HAR_OK;
    }
    const int64_t quads = har::ceil_div(num_windows, 4);
    const unsigned blocks = (unsigned)std::min<int64_t>(har::ceil_div(quads, 4), (int64_t)har::kNumSMs * 3 * 8);
    har::prof_begin(s);
    if (window_len <= 128)
        k_window_features_v3<16, false><<<blocks, 128, 0, s>>>(d_x, d_y, d_z, d_ts, d_win_start, num_windows, window_len, d_out, ld_out);
    else
        k_window_features_v3<32, false><<<blocks, 128, 0, s>>>(d_x, d_y, d_z, d_ts, d_win_start, num_windows, window_len, d_out, ld_out);
    HAR_LAUNCH_CHECK();
    har::prof_end(har::PROF_WINDOW_FEATURES, s, (double)num_windows * window_len);
    return HAR_OK;
}

// page-locked landing spot for the window count (per thread, allocated on first use; not used by the static entry point)
int64_t* pinned_total() {
    struct Holder { int64_t* p = nullptr; ~Holder() { if (p) cudaFreeHost(p); } };
    static thread_local Holder h;
    if (!h.p && cudaMallocHost((void**)&h.p, sizeof(int64_t)) != cudaSuccess) { cudaGetLastError(); h.p = nullptr; }
    return h.p;
}

// The fused call hands the count to the host while the feature kernel is still running: a side stream waits for the
// scan (event), copies the count into page-locked memory, and the host waits for that copy only.
struct CountLane {
    cudaStream_t side = nullptr;
    cudaEvent_t scanned = nullptr, copied = nullptr;
    int dev = -1;
    ~CountLane() { reset(); }
    void reset() {
        if (scanned) cudaEventDestroy(scanned);
        if (copied) cudaEventDestroy(copied);
        if (side) cudaStreamDestroy(side);
        side = nullptr; scanned = copied = nullptr; dev = -1;
    }
    cudaError_t ensure() {
        int cur = 0;
        cudaError_t e = cudaGetDevice(&cur);
        if (e != cudaSuccess) return e;
        if (side && dev == cur) return cudaSuccess;
        reset();
        if ((e = cudaStreamCreateWithFlags(&side, cudaStreamNonBlocking)) != cudaSuccess) return e;
        if ((e = cudaEventCreateWithFlags(&scanned, cudaEventDisableTiming)) != cudaSuccess) return e;
        if ((e = cudaEventCreateWithFlags(&copied, cudaEventDisableTiming)) != cudaSuccess) return e;
        dev = cur;
        return cudaSuccess;
    }
};
CountLane& count_lane() { static thread_local CountLane c; return c; }

// device -> host copy of the count and the one wait of the call
int read_total(const int64_t* d_total, int64_t* num_windows, cudaStream_t s) {
    int64_t stack_total = 0;
    int64_t* h = pinned_total();
    if (!h) h = &stack_total;
    HAR_CUDA(cudaMemcpyAsync(h, d_total, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    HAR_CUDA(cudaStreamSynchronize(s));
    *num_windows = *h;
    return HAR_OK;
}

}  // namespace

HAR_API int har_window_segments(const int32_t* d_user, const int32_t* d_activity, const int64_t* d_ts,
                                int64_t n, int32_t window_len, int64_t gap_ns, int64_t* d_win_start,
                                int32_t* d_win_user, int32_t* d_win_activity, int64_t* num_windows,
                                har_stream_t stream) {
    if (int rc = seg_check_args("har_window_segments", n, window_len, gap_ns, d_ts, num_windows)) return rc;
    if (n == 0) return HAR_OK;
    cudaStream_t s = (cudaStream_t)stream;
    har::DevBuf<unsigned char> scratch;
    HAR_CUDA(scratch.alloc(seg_scratch_bytes(har::ceil_div(n, SEG_TILE), SEG_TILE / window_len + 1), s));
    int64_t* d_total = nullptr;
    if (int rc = seg_launch(d_user, d_activity, d_ts, n, window_len, gap_ns, scratch.p, d_win_start, d_win_user,
                            d_win_activity, &d_total, s)) return rc;
    return read_total(d_total, num_windows, s);
}

// ---------------------------------------------------------------------------------------------
// har_window_segments_static: the same segmentation for hosts that must not allocate.
//
// Some programs that load this library are not allowed to ask CUDA for memory once they are running:
// a serving process that has already handed the whole device to its own memory manager (a later
// cudaMalloc or cudaMallocAsync fails or fragments that reservation), a control loop with a fixed
// latency budget (allocation calls may take a driver lock or wait for the device), or a plug-in host
// that only lets extensions touch buffers it pre-approved.  har_window_segments needs scratch memory
// for its per-tile summaries and takes it from cudaMallocAsync on every call, so those programs cannot
// use it.  This variant keeps that scratch in the array below, which is part of the library image: every
// GPU that loads the library carries these 4 MiB until the process exits, whether or not the function
// is ever called.  The caller still supplies all inputs and outputs.
//
// In host code `g_seg_static` is only a handle for the symbol, not a device address.  The device
// address is looked up with cudaGetSymbolAddress once per device (with that device current) and
// cached; the kernel gets the looked-up pointer.  That look-up happens inside the first call made on
// each device, so make that first call during start-up, before any stream capture begins.
//
// All calls on a device share the one array.  Two calls whose kernels can overlap on the GPU (different
// streams, or different host threads) would overwrite each other's tile summaries: callers must
// serialise them.  The array keeps whatever the previous call left in it.  Only the few hand-over words
// of the scan kernel must start at zero, and every call clears them itself (cudaMemsetAsync on the
// caller's stream, which allocates nothing); every other entry a call reads (per-tile summary, first
// window of the open bout, output position, counts, the window list of each tile up to its own count,
// and the total) is written earlier in the same call.
// ---------------------------------------------------------------------------------------------
constexpr size_t SEG_STATIC_BYTES = 4u << 20;      // HAR_SEGMENTS_STATIC_BYTES
static __device__ __align__(256) unsigned char g_seg_static[SEG_STATIC_BYTES];

static int seg_static_base(unsigned char** base) {
    // one slot per device, filled on first use.  No lock: cudaGetSymbolAddress returns the same address to every
    // thread for a given device, so two threads racing here store the same value.
    static unsigned char* cache[64] = {nullptr};
    int dev = 0;
    HAR_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return har::fail(HAR_E_UNSUPPORTED, "har_window_segments_static: device index %d not supported", dev);
    if (!cache[dev]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_seg_static);
        if (e != cudaSuccess || !p)
            return har::fail(HAR_E_CUDA, "har_window_segments_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
        cache[dev] = static_cast<unsigned char*>(p);
    }
    *base = cache[dev];
    return HAR_OK;
}

HAR_API int har_window_segments_static(const int32_t* d_user, const int32_t* d_activity, const int64_t* d_ts,
                                       int64_t n, int32_t window_len, int64_t gap_ns, int64_t* d_win_start,
                                       int32_t* d_win_user, int32_t* d_win_activity, int64_t* num_windows,
                                       har_stream_t stream) {
    if (int rc = seg_check_args("har_window_segments_static", n, window_len, gap_ns, d_ts, num_windows)) return rc;
    const int64_t num_tiles = har::ceil_div(n, SEG_TILE);
    const int32_t cap = SEG_TILE / window_len + 1;
    const size_t need = seg_scratch_bytes(num_tiles, cap);
    if (need > SEG_STATIC_BYTES) {
        // (the array holds fewer than 2^17 tiles, so the header is 512 bytes)
        const long long max_n = (long long)((SEG_STATIC_BYTES - 512) / seg_scratch_bytes_per_tile(cap)) * SEG_TILE;
        return har::fail(HAR_E_UNSUPPORTED,
                         "har_window_segments_static: %lld samples at window_len %d need %zu bytes of scratch but the "
                         "built-in array holds %zu; the limit at this window_len is %lld samples - use "
                         "har_window_segments, or split the stream at a bout boundary",
                         (long long)n, window_len, need, SEG_STATIC_BYTES, max_n);
    }
    unsigned char* base = nullptr;
    if (int rc = seg_static_base(&base)) return rc;
    if (n == 0) return HAR_OK;
    cudaStream_t s = (cudaStream_t)stream;
    int64_t* d_total = nullptr;
    if (int rc = seg_launch(d_user, d_activity, d_ts, n, window_len, gap_ns, base, d_win_start, d_win_user, d_win_activity,
                            &d_total, s)) return rc;
    int64_t h_total = 0;
    HAR_CUDA(cudaMemcpyAsync(&h_total, d_total, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    HAR_CUDA(cudaStreamSynchronize(s));
    *num_windows = h_total;
    return HAR_OK;
}

HAR_API int har_window_features(const double* d_x, const double* d_y, const double* d_z, const int64_t* d_ts,
                                int64_t n, const int64_t* d_win_start, int64_t num_windows, int32_t window_len,
                                double* d_out, int64_t ld_out, har_stream_t stream) {
    if (window_len < 3 || window_len > 256)
        return har::fail(HAR_E_UNSUPPORTED, "har_window_features: window_len must be in [3, 256], got %d", window_len);
    if (ld_out < HAR_NUM_WINDOW_FEATURES) return har::fail(HAR_E_INVALID, "har_window_features: ld_out < 43");
    if (num_windows <= 0) return HAR_OK;
    return features_launch(d_x, d_y, d_z, d_ts, n, d_win_start, num_windows, nullptr, window_len, d_out, ld_out,
                           (cudaStream_t)stream);
}

HAR_API int har_window_transform(const int32_t* d_user, const int32_t* d_activity, const int64_t* d_ts,
                                 const double* d_x, const double* d_y, const double* d_z, int64_t n, int32_t window_len,
                                 int64_t gap_ns, int64_t* d_win_start, int32_t* d_win_user, int32_t* d_win_activity,
                                 double* d_out, int64_t ld_out, int64_t* num_windows, har_stream_t stream) {
    if (int rc = seg_check_args("har_window_transform", n, window_len, gap_ns, d_ts, num_windows)) return rc;
    if (window_len < 3 || window_len > 256)
        return har::fail(HAR_E_UNSUPPORTED, "har_window_transform: window_len must be in [3, 256], got %d", window_len);
    if (ld_out < HAR_NUM_WINDOW_FEATURES) return har::fail(HAR_E_INVALID, "har_window_transform: ld_out < 43");
    if (!d_ts) return har::fail(HAR_E_INVALID, "har_window_transform: d_ts is NULL (the PEAK columns are time differences)");
    if (n == 0) return HAR_OK;
    cudaStream_t s = (cudaStream_t)stream;
    har::DevBuf<unsigned char> scratch;
    HAR_CUDA(scratch.alloc(seg_scratch_bytes(har::ceil_div(n, SEG_TILE), SEG_TILE / window_len + 1), s));
    int64_t* d_total = nullptr;
    // with window_len 200 the feature kernel also fills win_user / win_activity (it has every window's start at hand):
    // the emit pass then writes the window starts only
    const bool fused = window_len == W4_L;
    int64_t* h_total = fused ? pinned_total() : nullptr;
    CountLane& lane = count_lane();
    const bool early = fused && h_total && lane.ensure() == cudaSuccess;
    har::prof_span_begin(s);
    if (int rc = seg_launch(d_user, d_activity, d_ts, n, window_len, gap_ns, scratch.p, d_win_start, fused ? nullptr : d_win_user,
                            fused ? nullptr : d_win_activity, &d_total, s, early ? lane.scanned : nullptr)) return rc;
    if (fused) {
        // the feature kernel takes the window count from device memory: the kernels are queued back to back
        const int64_t cap = n / window_len;
        if (cap > 0)
            if (int rc = features_launch(d_x, d_y, d_z, d_ts, n, d_win_start, cap, d_total, window_len, d_out, ld_out, s,
                                         d_user, d_activity, d_win_user, d_win_activity)) return rc;
        har::prof_span_end(har::PROF_WINDOW_SPAN, s, (double)n);
        if (!early) return read_total(d_total, num_windows, s);
        // the count is ready as soon as the scan is: copy it on the side stream and wait for that copy only (the
        // feature kernel keeps running; `stream` must not free the scratch before it ends, hence the join below)
        HAR_CUDA(cudaStreamWaitEvent(lane.side, lane.scanned, 0));
        HAR_CUDA(cudaMemcpyAsync(h_total, d_total, sizeof(int64_t), cudaMemcpyDeviceToHost, lane.side));
        HAR_CUDA(cudaEventRecord(lane.copied, lane.side));
        HAR_CUDA(cudaStreamWaitEvent(s, lane.copied, 0));
        HAR_CUDA(cudaEventSynchronize(lane.copied));
        *num_windows = *h_total;
        return HAR_OK;
    }
    if (int rc = read_total(d_total, num_windows, s)) return rc;
    if (*num_windows > 0)
        return features_launch(d_x, d_y, d_z, d_ts, n, d_win_start, *num_windows, nullptr, window_len, d_out, ld_out, s);
    return HAR_OK;
}
