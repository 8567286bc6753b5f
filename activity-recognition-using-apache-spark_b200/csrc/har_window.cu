// WISDM window transform on sm_100a: segmentation (K1) and per-window features (K2/K3).
//
// No reference implementation exists for this stage: the reference ships only
// the transformed table (Main/wisdm_main_ver_0.0/data/wisdm_data.csv, schema line 1) and reads it at
// Main/main.py:18-20.  Definitions follow SURVEY Appendix B as fixed in DESIGN.md; oracle/har_oracle.c
// (orc_window_segments / orc_window_features) is the arbiter.
//
// K1  har_window_segments: three launches, the stream is read ONCE (8 B/sample, +8 with gap_ns):
//       k_seg_tiles  per 2048-sample tile: boundary flags, first/last boundary, the windows whose
//                    bout starts inside the tile (block scans: prefix-max + prefix-sum)
//       k_seg_scan   one CTA: prefix-max of last boundaries and prefix-sum of window counts over tiles
//       k_seg_emit   one thread per tile writes win_start / win_user / win_activity
// K2/K3 k_window_features: one warp per window; x/y/z slices (3 x 1.6 KB) are staged into shared
//       memory with cp.async.bulk (TMA 1-D) behind an mbarrier, double-buffered, so each sample is read
//       from HBM exactly once; reductions are warp shuffles in fp64 with a fixed butterfly order.
#include "har_common.cuh"

#include <algorithm>
#include <cstdlib>

namespace {

constexpr int SEG_THREADS = 256;
constexpr int SEG_PER_THREAD = 8;
constexpr int SEG_TILE = SEG_THREADS * SEG_PER_THREAD;

__device__ __forceinline__ int64_t warp_incl_max(int64_t v) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int64_t o = __shfl_up_sync(0xffffffffu, v, d);
        if ((threadIdx.x & 31) >= d) v = v > o ? v : o;
    }
    return v;
}
__device__ __forceinline__ int64_t warp_incl_sum(int64_t v) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int64_t o = __shfl_up_sync(0xffffffffu, v, d);
        if ((threadIdx.x & 31) >= d) v += o;
    }
    return v;
}

// Block-wide exclusive scans (blockDim.x <= 1024).  `total` receives the block aggregate.
template <bool IS_MAX>
__device__ __forceinline__ int64_t block_excl_scan(int64_t v, int64_t identity, int64_t* smem /*[33]*/, int64_t* total) {
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    int64_t incl = IS_MAX ? warp_incl_max(v) : warp_incl_sum(v);
    __syncthreads();
    if (lane == 31) smem[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int64_t w = lane < nw ? smem[lane] : identity;
        int64_t wi = IS_MAX ? warp_incl_max(w) : warp_incl_sum(w);
        smem[lane] = wi;
    }
    __syncthreads();
    int64_t excl = __shfl_up_sync(0xffffffffu, incl, 1);
    if (lane == 0) excl = identity;
    int64_t base = wid ? smem[wid - 1] : identity;
    if (total) *total = smem[nw - 1];
    if (IS_MAX) return excl > base ? excl : base;
    return excl + base;
}

__global__ void __launch_bounds__(SEG_THREADS)
k_seg_tiles(const int32_t* __restrict__ user, const int32_t* __restrict__ act, const int64_t* __restrict__ ts,
            int64_t n, int32_t L, int64_t gap_ns, int vec_ok,
            int64_t* __restrict__ first_b, int64_t* __restrict__ last_b, int32_t* __restrict__ nlocal,
            int64_t* __restrict__ tile_wins, int32_t cap) {
    __shared__ int64_t sm[33];
    __shared__ unsigned long long s_first;
    const int64_t tile = blockIdx.x;
    const int64_t base = tile * SEG_TILE + (int64_t)threadIdx.x * SEG_PER_THREAD;
    if (threadIdx.x == 0) s_first = ~0ULL;

    int32_t u[SEG_PER_THREAD], a[SEG_PER_THREAD];
    int32_t pu = 0, pa = 0;
    int64_t pt = 0;
    if (vec_ok && base + SEG_PER_THREAD <= n) {
        const int4* u4 = reinterpret_cast<const int4*>(user + base);
        const int4* a4 = reinterpret_cast<const int4*>(act + base);
        int4 v0 = __ldg(u4), v1 = __ldg(u4 + 1), w0 = __ldg(a4), w1 = __ldg(a4 + 1);
        u[0] = v0.x; u[1] = v0.y; u[2] = v0.z; u[3] = v0.w; u[4] = v1.x; u[5] = v1.y; u[6] = v1.z; u[7] = v1.w;
        a[0] = w0.x; a[1] = w0.y; a[2] = w0.z; a[3] = w0.w; a[4] = w1.x; a[5] = w1.y; a[6] = w1.z; a[7] = w1.w;
    } else {
#pragma unroll
        for (int j = 0; j < SEG_PER_THREAD; j++) {
            int64_t i = base + j;
            u[j] = i < n ? user[i] : 0;
            a[j] = i < n ? act[i] : 0;
        }
    }
    if (base > 0 && base < n) { pu = user[base - 1]; pa = act[base - 1]; if (gap_ns > 0) pt = ts[base - 1]; }

    unsigned flags = 0;
    int64_t my_last = -1, my_first = -1;
#pragma unroll
    for (int j = 0; j < SEG_PER_THREAD; j++) {
        int64_t i = base + j;
        if (i < n) {
            bool b = (i == 0) || u[j] != pu || a[j] != pa;
            if (gap_ns > 0) {
                int64_t t = ts[i];
                b = b || (i > 0 && t - pt > gap_ns);
                pt = t;
            }
            if (b) { flags |= 1u << j; my_last = i; if (my_first < 0) my_first = i; }
            pu = u[j]; pa = a[j];
        }
    }
    __syncthreads();
    if (my_first >= 0) atomicMin(&s_first, (unsigned long long)my_first);
    int64_t tot_last;
    int64_t carry = block_excl_scan<true>(my_last, -1, sm, &tot_last);
    // windows whose bout starts in this tile: window END at i with (i - bstart + 1) % L == 0
    int cnt = 0;
    int64_t cur = carry;
    int64_t starts[SEG_PER_THREAD];
#pragma unroll
    for (int j = 0; j < SEG_PER_THREAD; j++) {
        int64_t i = base + j;
        starts[j] = -1;
        if (i < n) {
            if (flags & (1u << j)) cur = i;
            if (cur >= 0 && (i - cur + 1) % L == 0) { starts[j] = i - L + 1; cnt++; }
        }
    }
    int64_t tot_cnt;
    int64_t ord = block_excl_scan<false>(cnt, 0, sm, &tot_cnt);
#pragma unroll
    for (int j = 0; j < SEG_PER_THREAD; j++)
        if (starts[j] >= 0) tile_wins[tile * cap + (ord++)] = starts[j];
    if (threadIdx.x == 0) {
        first_b[tile] = s_first == ~0ULL ? -1 : (int64_t)s_first;
        last_b[tile] = tot_last;
        nlocal[tile] = (int32_t)tot_cnt;
    }
}

__global__ void __launch_bounds__(1024)
k_seg_scan(const int64_t* __restrict__ first_b, const int64_t* __restrict__ last_b,
           const int32_t* __restrict__ nlocal, int64_t num_tiles, int64_t n, int32_t L,
           int64_t* __restrict__ bstart_in, int32_t* __restrict__ npre, int64_t* __restrict__ win_base,
           int64_t* __restrict__ total_out) {
    __shared__ int64_t sm[33];
    int64_t carry_b = -1, carry_c = 0;
    for (int64_t t0 = 0; t0 < num_tiles; t0 += blockDim.x) {
        int64_t t = t0 + threadIdx.x;
        bool ok = t < num_tiles;
        int64_t lb = ok ? last_b[t] : -1, tot_b;
        int64_t bin = block_excl_scan<true>(lb, -1, sm, &tot_b);
        if (carry_b > bin) bin = carry_b;
        int64_t c = 0;
        if (ok) {
            int64_t tb = t * SEG_TILE, fb = first_b[t];
            int64_t end_pre = fb >= 0 ? fb : (tb + SEG_TILE < n ? tb + SEG_TILE : n);
            int64_t pre = 0;
            if (bin >= 0) pre = (end_pre - bin) / L - (tb - bin) / L;
            bstart_in[t] = bin;
            npre[t] = (int32_t)pre;
            c = pre + nlocal[t];
        }
        int64_t tot_c;
        int64_t ex = block_excl_scan<false>(c, 0, sm, &tot_c);
        if (ok) win_base[t] = carry_c + ex;
        carry_c += tot_c;
        if (tot_b > carry_b) carry_b = tot_b;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total_out = carry_c;
}

__global__ void k_seg_emit(const int64_t* __restrict__ bstart_in, const int32_t* __restrict__ npre,
                           const int32_t* __restrict__ nlocal, const int64_t* __restrict__ win_base,
                           const int64_t* __restrict__ tile_wins, int32_t cap, int64_t num_tiles, int32_t L,
                           const int32_t* __restrict__ user, const int32_t* __restrict__ act,
                           int64_t* __restrict__ win_start, int32_t* __restrict__ win_user,
                           int32_t* __restrict__ win_act) {
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= num_tiles) return;
    int64_t o = win_base[t];
    int pre = npre[t];
    if (pre > 0) {
        int64_t b = bstart_in[t], tb = t * SEG_TILE;
        int64_t m0 = (tb - b) / L;
        for (int k = 0; k < pre; k++) {
            int64_t s = b + (m0 + k) * L;
            win_start[o] = s;
            if (win_user) win_user[o] = user[s];
            if (win_act) win_act[o] = act[s];
            o++;
        }
    }
    int nl = nlocal[t];
    for (int k = 0; k < nl; k++) {
        int64_t s = tile_wins[t * cap + k];
        win_start[o] = s;
        if (win_user) win_user[o] = user[s];
        if (win_act) win_act[o] = act[s];
        o++;
    }
}

// ---------------------------------------------------------------------------------------------
// K2/K3
// ---------------------------------------------------------------------------------------------
constexpr int WF_WARPS = 8;
constexpr int WF_STAGES = 2;

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

// insert h into a descending triple (multiset semantics)
__device__ __forceinline__ void top3_insert(double h, double& p1, double& p2, double& p3) {
    if (h > p1) { p3 = p2; p2 = p1; p1 = h; }
    else if (h > p2) { p3 = p2; p2 = h; }
    else if (h > p3) { p3 = h; }
}

struct AxisOut {
    double mean, absdev, stddev, peak;
    unsigned long long lo5, hi5;   // bin counts, 12-bit fields: bins 0-4 / 5-9
};

// One axis of one window.  v points at the window's first sample in shared memory.
__device__ __forceinline__ AxisOut axis_features(const double* v, const int64_t* __restrict__ ts, int L, int lane) {
    AxisOut o;
    double sum = 0.0, lo = INFINITY, hi = -INFINITY;
    for (int i = lane; i < L; i += 32) {
        double a = v[i];
        sum += a; lo = fmin(lo, a); hi = fmax(hi, a);
    }
    sum = warp_sum_d(sum); lo = warp_min_d(lo); hi = warp_max_d(hi);
    const double mean = sum / (double)L;
    const double scale = hi > lo ? 10.0 / (hi - lo) : 0.0;
    double sq = 0.0, ab = 0.0;
    unsigned long long lo5 = 0, hi5 = 0;
    int npk = 0;
    double p1 = -INFINITY, p2 = -INFINITY, p3 = -INFINITY;
    for (int i = lane; i < L; i += 32) {
        double a = v[i];
        double d = a - mean;
        sq = fma(d, d, sq);
        ab += fabs(d);
        double q = floor((a - lo) * scale);
        int b = q >= 9.0 ? 9 : (int)q;
        if (b < 5) lo5 += 1ULL << (12 * b); else hi5 += 1ULL << (12 * (b - 5));
        if (i >= 1 && i + 1 < L && a > v[i - 1] && a >= v[i + 1]) { npk++; top3_insert(a, p1, p2, p3); }
    }
    sq = warp_sum_d(sq); ab = warp_sum_d(ab);
    o.lo5 = warp_sum_u64(lo5); o.hi5 = warp_sum_u64(hi5);
    o.mean = mean;
    o.absdev = ab / (double)L;
    o.stddev = sqrt(sq / (double)L);
    npk = __reduce_add_sync(0xffffffffu, npk);
    // merge the per-lane top-3 peak heights (disjoint multisets) with an xor butterfly
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1) {
        double q1 = shfl_xor_d(p1, m), q2 = shfl_xor_d(p2, m), q3 = shfl_xor_d(p3, m);
        top3_insert(q1, p1, p2, p3); top3_insert(q2, p1, p2, p3); top3_insert(q3, p1, p2, p3);
    }
    o.peak = __longlong_as_double(0x7ff8000000000000LL);
    if (npk >= 3) {
        // threshold ladder tau_k = P - ((k+1)*0.1)*(P - lo), k = 0..8, then lo: first one admitting 3 peaks
        const double P = p1, R = __dsub_rn(P, lo);
        double tau = lo;
#pragma unroll 1
        for (int k = 0; k < 9; k++) {
            double t = __dsub_rn(P, __dmul_rn(__dmul_rn((double)(k + 1), 0.1), R));
            if (t <= p3) { tau = t; break; }
        }
        int first = 0x7fffffff, last = -1, cnt = 0;
        for (int i = lane; i < L; i += 32) {
            double a = v[i];
            if (i >= 1 && i + 1 < L && a > v[i - 1] && a >= v[i + 1] && a >= tau) { first = min(first, i); last = max(last, i); cnt++; }
        }
        first = __reduce_min_sync(0xffffffffu, first);
        last = __reduce_max_sync(0xffffffffu, last);
        cnt = __reduce_add_sync(0xffffffffu, cnt);
        o.peak = ((double)(ts[last] - ts[first]) / (double)(cnt - 1)) / 1.0e6;
    }
    return o;
}

__global__ void __launch_bounds__(WF_WARPS * 32, 2)
k_window_features(const double* __restrict__ x, const double* __restrict__ y, const double* __restrict__ z,
                  const int64_t* __restrict__ ts, const int64_t* __restrict__ win_start, int64_t nw, int64_t n,
                  int L, int stride /* doubles per staged axis slice, even */, double* __restrict__ out, int64_t ld_out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // per warp: [STAGES][3][stride] doubles, [43+1] doubles of output staging, STAGES mbarriers
    const size_t per_warp = ((size_t)WF_STAGES * 3 * stride + 44) * sizeof(double) + WF_STAGES * sizeof(uint64_t);
    unsigned char* mine = smem_raw + per_warp * warp;
    double* buf = reinterpret_cast<double*>(mine);
    double* ostage = buf + (size_t)WF_STAGES * 3 * stride;
    uint64_t* bars = reinterpret_cast<uint64_t*>(ostage + 44);

    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < WF_STAGES; s++) mbar_init(&bars[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();

    const int64_t gw = (int64_t)blockIdx.x * WF_WARPS + warp;
    const int64_t nwarps = (int64_t)gridDim.x * WF_WARPS;
    const double* axes[3] = {x, y, z};

    // issue the loads of window w into stage s; returns nothing (all lanes call, lane 0 issues)
    auto issue = [&](int64_t w, int s) {
        int64_t st = win_start[w];
        int64_t s0 = st & ~(int64_t)1;
        int64_t e1 = (st + L + 1) & ~(int64_t)1;          // exclusive, rounded up to even
        double* dst = buf + (size_t)s * 3 * stride;
        bool tma_ok = e1 <= n && ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) |
                                   reinterpret_cast<uintptr_t>(z)) & 15) == 0;
        if (tma_ok) {
            if (lane == 0) {
                uint32_t bytes = (uint32_t)((e1 - s0) * sizeof(double));
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_expect_tx(&bars[s], 3 * bytes);
#pragma unroll
                for (int a = 0; a < 3; a++) bulk_g2s(dst + a * stride, axes[a] + s0, bytes, &bars[s]);
            }
        } else {
            // tail window (or unaligned base pointers): plain loads, then a plain arrive
            for (int a = 0; a < 3; a++)
                for (int i = lane; i < L; i += 32) dst[a * stride + (st - s0) + i] = axes[a][st + i];
            __syncwarp();
            if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&bars[s])) : "memory");
        }
    };

    uint32_t phase_bits = 0;      // bit s = parity to wait for on stage s
    int s = 0;
    if (gw < nw) issue(gw, 0);
    for (int64_t w = gw; w < nw; w += nwarps) {
        int64_t wn = w + nwarps;
        if (wn < nw) issue(wn, s ^ 1);
        mbar_wait(&bars[s], (phase_bits >> s) & 1u);
        phase_bits ^= 1u << s;
        const int64_t st = win_start[w];
        const int off = (int)(st & 1);
        const double* base = buf + (size_t)s * 3 * stride + off;
        const int64_t* tsw = ts + st;

        double rs = 0.0;
        for (int i = lane; i < L; i += 32) {
            double a = base[i], b = base[stride + i], c = base[2 * stride + i];
            rs += sqrt(fma(c, c, fma(b, b, a * a)));
        }
        rs = warp_sum_d(rs);
#pragma unroll 1
        for (int a = 0; a < 3; a++) {
            AxisOut o = axis_features(base + a * stride, tsw, L, lane);
            if (lane == 0) {
#pragma unroll
                for (int b = 0; b < 5; b++) {
                    ostage[a * 10 + b] = (double)((o.lo5 >> (12 * b)) & 0xfff) / (double)L;
                    ostage[a * 10 + 5 + b] = (double)((o.hi5 >> (12 * b)) & 0xfff) / (double)L;
                }
                ostage[30 + a] = o.mean;
                ostage[33 + a] = o.peak;
                ostage[36 + a] = o.absdev;
                ostage[39 + a] = o.stddev;
            }
        }
        if (lane == 0) ostage[42] = rs / (double)L;
        __syncwarp();
        double* dst = out + w * ld_out;
        dst[lane] = ostage[lane];
        if (lane < HAR_NUM_WINDOW_FEATURES - 32) dst[32 + lane] = ostage[32 + lane];
        __syncwarp();
        s ^= 1;
    }
}

#include "har_window_v2.cuh"
#include "har_window_v3.cuh"

}  // namespace

HAR_API int har_window_segments(const int32_t* d_user, const int32_t* d_activity, const int64_t* d_ts,
                                int64_t n, int32_t window_len, int64_t gap_ns, int64_t* d_win_start,
                                int32_t* d_win_user, int32_t* d_win_activity, int64_t* num_windows,
                                har_stream_t stream) {
    if (!num_windows) return har::fail(HAR_E_INVALID, "har_window_segments: num_windows is NULL");
    *num_windows = 0;
    if (n < 0 || window_len < 2) return har::fail(HAR_E_INVALID, "har_window_segments: need n >= 0 and window_len >= 2");
    if (n == 0) return HAR_OK;
    if (gap_ns > 0 && !d_ts) return har::fail(HAR_E_INVALID, "har_window_segments: gap_ns > 0 needs d_ts");
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t num_tiles = har::ceil_div(n, SEG_TILE);
    const int32_t cap = SEG_TILE / window_len + 1;
    har::DevBuf<int64_t> first_b, last_b, bstart_in, win_base, tile_wins, total;
    har::DevBuf<int32_t> nlocal, npre;
    HAR_CUDA(first_b.alloc(num_tiles, s)); HAR_CUDA(last_b.alloc(num_tiles, s));
    HAR_CUDA(bstart_in.alloc(num_tiles, s)); HAR_CUDA(win_base.alloc(num_tiles, s));
    HAR_CUDA(tile_wins.alloc((size_t)num_tiles * cap, s)); HAR_CUDA(total.alloc(1, s));
    HAR_CUDA(nlocal.alloc(num_tiles, s)); HAR_CUDA(npre.alloc(num_tiles, s));
    int vec_ok = ((reinterpret_cast<uintptr_t>(d_user) | reinterpret_cast<uintptr_t>(d_activity)) & 15) == 0;
    har::prof_begin(s);
    k_seg_tiles<<<(unsigned)num_tiles, SEG_THREADS, 0, s>>>(d_user, d_activity, d_ts, n, window_len, gap_ns, vec_ok,
                                                            first_b.p, last_b.p, nlocal.p, tile_wins.p, cap);
    HAR_LAUNCH_CHECK();
    har::prof_end(har::PROF_SEG_TILES, s, (double)n);
    k_seg_scan<<<1, 1024, 0, s>>>(first_b.p, last_b.p, nlocal.p, num_tiles, n, window_len, bstart_in.p, npre.p,
                                  win_base.p, total.p);
    HAR_LAUNCH_CHECK();
    k_seg_emit<<<(unsigned)har::ceil_div(num_tiles, 128), 128, 0, s>>>(bstart_in.p, npre.p, nlocal.p, win_base.p,
                                                                       tile_wins.p, cap, num_tiles, window_len,
                                                                       d_user, d_activity, d_win_start, d_win_user,
                                                                       d_win_activity);
    HAR_LAUNCH_CHECK();
    int64_t h_total = 0;
    HAR_CUDA(cudaMemcpyAsync(&h_total, total.p, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    HAR_CUDA(cudaStreamSynchronize(s));
    *num_windows = h_total;
    return HAR_OK;
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
// cached; the kernels get the looked-up pointer.  That look-up happens inside the first call made on
// each device, so make that first call during start-up, before any stream capture begins.
//
// All calls on a device share the one array.  Two calls whose kernels can overlap on the GPU (different
// streams, or different host threads) would overwrite each other's tile summaries: callers must
// serialise them.  The array keeps whatever the previous call left in it; that is harmless here because
// every entry a call reads (per-tile first/last boundary, counts, prefix sums, the window list of each
// tile up to its own count, and the total) is written earlier in the same call, so nothing has to be
// cleared to zero beforehand.
// ---------------------------------------------------------------------------------------------
constexpr size_t SEG_STATIC_BYTES = 4u << 20;      // HAR_SEGMENTS_STATIC_BYTES
static __device__ __align__(256) unsigned char g_seg_static[SEG_STATIC_BYTES];

static int seg_static_base(unsigned char** base) {
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
    if (!num_windows) return har::fail(HAR_E_INVALID, "har_window_segments_static: num_windows is NULL");
    *num_windows = 0;
    if (n < 0 || window_len < 2) return har::fail(HAR_E_INVALID, "har_window_segments_static: need n >= 0 and window_len >= 2");
    if (gap_ns > 0 && !d_ts) return har::fail(HAR_E_INVALID, "har_window_segments_static: gap_ns > 0 needs d_ts");
    // carve the fixed array: 4 x int64 + 2 x int32 per tile, `cap` window slots per tile, one int64 total
    const int64_t num_tiles = har::ceil_div(n, SEG_TILE);
    const int32_t cap = SEG_TILE / window_len + 1;
    const size_t per_tile = 4 * sizeof(int64_t) + 2 * sizeof(int32_t) + (size_t)cap * sizeof(int64_t);
    const size_t need = 256 + (size_t)num_tiles * per_tile;
    if (need > SEG_STATIC_BYTES) {
        const long long max_n = (long long)((SEG_STATIC_BYTES - 256) / per_tile) * SEG_TILE;
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
    int64_t* total = reinterpret_cast<int64_t*>(base);
    int64_t* first_b = reinterpret_cast<int64_t*>(base + 256);
    int64_t* last_b = first_b + num_tiles;
    int64_t* bstart_in = last_b + num_tiles;
    int64_t* win_base = bstart_in + num_tiles;
    int64_t* tile_wins = win_base + num_tiles;
    int32_t* nlocal = reinterpret_cast<int32_t*>(tile_wins + num_tiles * cap);
    int32_t* npre = nlocal + num_tiles;
    int vec_ok = ((reinterpret_cast<uintptr_t>(d_user) | reinterpret_cast<uintptr_t>(d_activity)) & 15) == 0;
    k_seg_tiles<<<(unsigned)num_tiles, SEG_THREADS, 0, s>>>(d_user, d_activity, d_ts, n, window_len, gap_ns, vec_ok,
                                                            first_b, last_b, nlocal, tile_wins, cap);
    HAR_LAUNCH_CHECK();
    k_seg_scan<<<1, 1024, 0, s>>>(first_b, last_b, nlocal, num_tiles, n, window_len, bstart_in, npre, win_base, total);
    HAR_LAUNCH_CHECK();
    k_seg_emit<<<(unsigned)har::ceil_div(num_tiles, 128), 128, 0, s>>>(bstart_in, npre, nlocal, win_base, tile_wins, cap,
                                                                       num_tiles, window_len, d_user, d_activity,
                                                                       d_win_start, d_win_user, d_win_activity);
    HAR_LAUNCH_CHECK();
    int64_t h_total = 0;
    HAR_CUDA(cudaMemcpyAsync(&h_total, total, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
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
    cudaStream_t s = (cudaStream_t)stream;
    // HAR_WINDOW_KERNEL = v3 (default; 8 lanes/window, register resident, compact code) | v2 (16 lanes/window,
    // unrolled over axes) | tma (first version: one warp per window, cp.async.bulk staging) - kept for A/B timing
    const char* sel = std::getenv("HAR_WINDOW_KERNEL");
    const std::string which = sel ? sel : "v3";
    if (which == "v3") {
        const int64_t quads = har::ceil_div(num_windows, 4);
        const unsigned blocks = (unsigned)std::min<int64_t>(har::ceil_div(quads, 4), (int64_t)har::kNumSMs * 3 * 8);
        har::prof_begin(s);
        if (window_len == 200)
            k_window_features_v3<25, true><<<blocks, 128, 0, s>>>(d_x, d_y, d_z, d_ts, d_win_start, num_windows, window_len, d_out, ld_out);
        else if (window_len <= 128)
            k_window_features_v3<16, false><<<blocks, 128, 0, s>>>(d_x, d_y, d_z, d_ts, d_win_start, num_windows, window_len, d_out, ld_out);
        else
            k_window_features_v3<32, false><<<blocks, 128, 0, s>>>(d_x, d_y, d_z, d_ts, d_win_start, num_windows, window_len, d_out, ld_out);
        HAR_LAUNCH_CHECK();
        har::prof_end(har::PROF_WINDOW_FEATURES, s, (double)num_windows * window_len);
        return HAR_OK;
    }
    if (which == "v2") {
        const int64_t pairs = har::ceil_div(num_windows, 2);
        int64_t blocks = std::min<int64_t>(har::ceil_div(pairs, 4), (int64_t)har::kNumSMs * 3 * 8);
        const int kiter = (window_len + 15) / 16;
        har::prof_begin(s);
        if (kiter <= 13)
            k_window_features_v2<13><<<(unsigned)blocks, 128, 0, s>>>(d_x, d_y, d_z, d_ts, d_win_start, num_windows, window_len, d_out, ld_out);
        else
            k_window_features_v2<16><<<(unsigned)blocks, 128, 0, s>>>(d_x, d_y, d_z, d_ts, d_win_start, num_windows, window_len, d_out, ld_out);
        HAR_LAUNCH_CHECK();
        har::prof_end(har::PROF_WINDOW_FEATURES, s, (double)num_windows * window_len);
        return HAR_OK;
    }
    const int stride = (window_len + 3) & ~1;
    const size_t per_warp = ((size_t)WF_STAGES * 3 * stride + 44) * sizeof(double) + WF_STAGES * sizeof(uint64_t);
    const size_t smem = per_warp * WF_WARPS;
    HAR_CUDA(cudaFuncSetAttribute(k_window_features, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t blocks = har::ceil_div(num_windows, WF_WARPS);
    int64_t max_blocks = (int64_t)har::kNumSMs * 2 * 4;       // 2 resident CTAs/SM, a few waves of work each
    if (blocks > max_blocks) blocks = max_blocks;
    har::prof_begin(s);
    k_window_features<<<(unsigned)blocks, WF_WARPS * 32, smem, s>>>(d_x, d_y, d_z, d_ts, d_win_start, num_windows, n,
                                                                   window_len, stride, d_out, ld_out);
    HAR_LAUNCH_CHECK();
    har::prof_end(har::PROF_WINDOW_FEATURES, s, (double)num_windows * window_len);
    return HAR_OK;
}
