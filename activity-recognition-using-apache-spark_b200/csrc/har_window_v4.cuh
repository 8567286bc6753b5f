// Window transform, second generation (included by har_window.cu inside its anonymous namespace).
//
// k_seg_tiles/scan/emit segmentation: one streaming pass over (user, activity[, ts]) that leaves a 64-bit summary and the
//                       local window starts per 2048-sample tile, a one-CTA scan of the summaries, and the emit pass.
// k_window_features_v4  window length 200: a warp owns four windows at a time, eight lanes per window, every lane 25
//                       CONSECUTIVE samples.  The four 1.6 KB slices of one axis are brought into shared memory by
//                       cp.async.bulk (TMA 1-D, SASS UBLKCP) behind an mbarrier, two items ahead of the arithmetic, so
//                       each sample crosses HBM->SM once and costs one LDS; neighbours for the local-maximum test are
//                       registers of the same lane (one shuffle per lane edge), the 10-bin counts are packed 5-bit
//                       fields accumulated by one fp64 add per sample (no atomics), x^2+y^2+z^2 stays in registers
//                       across the three axes.  All discrete decisions use the oracle's exact fp64 operations.
#pragma once

// ---------------------------------------------------------------------------------------------------------------------
// segmentation: three launches, the stream is read once
//   k_seg_tiles  one CTA per 2048-sample tile: bout boundaries, the windows of bouts that START inside the tile (as
//                16-bit tile-relative starts) and a 64-bit summary {first boundary, last boundary, local windows}
//   k_seg_scan   one CTA over the summaries: the bout that is open when each tile begins (prefix maximum of the last
//                boundaries), how many of its windows end inside the tile (closed form), and the tile's first output
//                position (prefix sum).  Every thread walks a contiguous run of tiles, so the block-wide scans run once.
//   k_seg_emit   eight lanes per tile write win_start / win_user / win_activity
// A single-pass version (tiles exchanging their running totals by decoupled look-back) was measured slower on B200
// (0.73 ms per 1e8 samples against 0.29 + 0.16 + 0.03 ms for the round-1 kernels): with ~10 windows per tile there is
// too little work per tile to hide the look-back round trips through L2, and seven of eight warps wait at the barrier.
// ---------------------------------------------------------------------------------------------------------------------
struct SegScratch {                // carved from one allocation (or from the built-in array of the static entry point)
    unsigned long long* summary;   // [tiles]  bits 0-15 first boundary + 1 (0 = none), 16-31 last boundary + 1, 32-47 local windows
    int64_t* open_bout;            // [tiles]  start of the first window of the open bout that ends inside the tile
    int64_t* out_base;             // [tiles]  output position of the tile's first window
    int32_t* npre;                 // [tiles]  windows of the open bout that end inside the tile
    uint16_t* local;               // [tiles][cap] tile-relative starts of the local windows
    int64_t* total;
    unsigned long long* chain;     // [slabs of 8192 tiles][2] hand-over words of k_seg_scan; must start at zero
};
__host__ __device__ inline size_t seg_scratch_bytes_per_tile(int cap) { return 8 + 8 + 8 + 4 + (((size_t)cap * 2 + 3) & ~(size_t)3); }

// x / L for 0 <= x < 2^21 / L by a multiply-high with magic = floor(2^32 / L) + 1 (exact while x * L < 2^32)
__device__ __forceinline__ unsigned umod_small(unsigned x, unsigned L, unsigned magic) { return x - __umulhi(x, magic) * L; }

__global__ void __launch_bounds__(SEG_THREADS)
k_seg_tiles(const int32_t* __restrict__ user, const int32_t* __restrict__ act, const int64_t* __restrict__ ts,
            int64_t n, int32_t L, unsigned magic, int64_t gap_ns, int vec_ok, SegScratch sc, int32_t cap) {
    __shared__ __align__(16) int sm_last[SEG_THREADS / 32], sm_first[SEG_THREADS / 32], sm_cnt[SEG_THREADS / 32];
    const int64_t tile = blockIdx.x;
    const int64_t tile0 = tile * SEG_TILE;
    const int rel0 = (int)threadIdx.x * SEG_PER_THREAD;
    const int64_t base = tile0 + rel0;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    // whole tile inside the stream, aligned columns, no timestamp gaps: no per-sample bounds or gap tests
    const bool fast = vec_ok && gap_ns <= 0 && tile0 + SEG_TILE <= n;

    int32_t u[SEG_PER_THREAD], a[SEG_PER_THREAD];
    unsigned flags = 0, valid = 0xffu;
    if (fast) {
        const int4* u4 = reinterpret_cast<const int4*>(user + base);
        const int4* a4 = reinterpret_cast<const int4*>(act + base);
        int4 v0 = __ldg(u4), v1 = __ldg(u4 + 1), w0 = __ldg(a4), w1 = __ldg(a4 + 1);
        u[0] = v0.x; u[1] = v0.y; u[2] = v0.z; u[3] = v0.w; u[4] = v1.x; u[5] = v1.y; u[6] = v1.z; u[7] = v1.w;
        a[0] = w0.x; a[1] = w0.y; a[2] = w0.z; a[3] = w0.w; a[4] = w1.x; a[5] = w1.y; a[6] = w1.z; a[7] = w1.w;
        // the sample before this thread's first one: the neighbour lane has it, except for lane 0 of each warp
        int32_t pu = __shfl_up_sync(0xffffffffu, u[SEG_PER_THREAD - 1], 1);
        int32_t pa = __shfl_up_sync(0xffffffffu, a[SEG_PER_THREAD - 1], 1);
        if (lane == 0 && base > 0) { pu = __ldg(user + base - 1); pa = __ldg(act + base - 1); }
#pragma unroll
        for (int j = 0; j < SEG_PER_THREAD; j++) {
            flags |= ((u[j] != pu) | (a[j] != pa) ? 1u : 0u) << j;
            pu = u[j]; pa = a[j];
        }
        if (base == 0) flags |= 1u;
    } else {
        int32_t pu = 0, pa = 0;
        int64_t pt = 0;
        valid = 0;
#pragma unroll
        for (int j = 0; j < SEG_PER_THREAD; j++) {
            const int64_t i = base + j;
            u[j] = i < n ? user[i] : 0;
            a[j] = i < n ? act[i] : 0;
            if (i < n) valid |= 1u << j;
        }
        if (base > 0 && base < n) { pu = user[base - 1]; pa = act[base - 1]; if (gap_ns > 0) pt = ts[base - 1]; }
#pragma unroll
        for (int j = 0; j < SEG_PER_THREAD; j++) {
            const int64_t i = base + j;
            if (i < n) {
                bool b = (i == 0) || u[j] != pu || a[j] != pa;
                if (gap_ns > 0) {
                    const int64_t t = ts[i];
                    b = b || (i > 0 && t - pt > gap_ns);
                    pt = t;
                }
                if (b) flags |= 1u << j;
                pu = u[j]; pa = a[j];
            }
        }
    }
    const int my_last = flags ? rel0 + 31 - __clz(flags) : -1;
    const int my_first = flags ? rel0 + __ffs(flags) - 1 : SEG_TILE;
    // latest boundary before this thread (exclusive prefix maximum), first boundary of the tile
    int incl = my_last, wfirst = my_first;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        incl = max(incl, __shfl_up_sync(0xffffffffu, incl, d));       // lanes below d get their own value back: no-op
        wfirst = min(wfirst, __shfl_xor_sync(0xffffffffu, wfirst, d));
    }
    if (lane == 31) sm_last[wid] = incl;
    if (lane == 0) sm_first[wid] = wfirst;
    __syncthreads();
    int carry = __shfl_up_sync(0xffffffffu, incl, 1);
    if (lane == 0) carry = -1;
    int tile_last = -1, tile_first = SEG_TILE;
    {
        const int4 l0 = *reinterpret_cast<const int4*>(sm_last), l1 = *reinterpret_cast<const int4*>(sm_last + 4);
        const int4 f0 = *reinterpret_cast<const int4*>(sm_first), f1 = *reinterpret_cast<const int4*>(sm_first + 4);
        const int lv[8] = {l0.x, l0.y, l0.z, l0.w, l1.x, l1.y, l1.z, l1.w};
#pragma unroll
        for (int w = 0; w < SEG_THREADS / 32; w++) {
            if (w < wid) carry = max(carry, lv[w]);
            tile_last = max(tile_last, lv[w]);
        }
        tile_first = min(min(min(f0.x, f0.y), min(f0.z, f0.w)), min(min(f1.x, f1.y), min(f1.z, f1.w)));
    }
    // windows of bouts that START in this tile: a window ends where the running length of the bout hits a multiple of L
    int cnt = 0;
    unsigned ends = 0;
    if (L > SEG_PER_THREAD) {
        // at most one window can end inside these 8 samples, and only one of the bout that is open at the thread's start
        if (carry >= 0) {
            const int run = (int)umod_small((unsigned)(rel0 - carry), (unsigned)L, magic);   // samples of the bout so far, mod L
            const int pend = L - run - 1;                                                    // it would end at this offset
            const int stop = flags ? __ffs(flags) - 1 : SEG_PER_THREAD;                      // the bout ends before this offset
            if (pend < stop && ((valid >> pend) & 1u)) { ends = 1u << pend; cnt = 1; }
        }
    } else {
        int run = carry >= 0 ? (int)umod_small((unsigned)(rel0 - carry), (unsigned)L, magic) : -1;
#pragma unroll
        for (int j = 0; j < SEG_PER_THREAD; j++) {
            if ((valid >> j) & 1u) {
                if (flags & (1u << j)) run = 0;
                if (run >= 0 && ++run == L) { run = 0; ends |= 1u << j; cnt++; }
            }
        }
    }
    int cincl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int o = __shfl_up_sync(0xffffffffu, cincl, d);
        if (lane >= d) cincl += o;
    }
    if (lane == 31) sm_cnt[wid] = cincl;
    __syncthreads();
    int ord = cincl - cnt, nlocal = 0;
    {
        const int4 c0 = *reinterpret_cast<const int4*>(sm_cnt), c1 = *reinterpret_cast<const int4*>(sm_cnt + 4);
        const int cv[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
#pragma unroll
        for (int w = 0; w < SEG_THREADS / 32; w++) {
            if (w < wid) ord += cv[w];
            nlocal += cv[w];
        }
    }
    if (cnt) {
        uint16_t* dst = sc.local + (size_t)tile * cap;
#pragma unroll
        for (int j = 0; j < SEG_PER_THREAD; j++)
            if (ends & (1u << j)) dst[ord++] = (uint16_t)(rel0 + j - L + 1);
    }
    if (threadIdx.x == 0)
        sc.summary[tile] = (unsigned long long)(tile_first == SEG_TILE ? 0 : tile_first + 1) |
                           ((unsigned long long)(tile_last + 1) << 16) | ((unsigned long long)nlocal << 32);
}

// One CTA per slab of 8192 tile summaries, all slabs at once.  A slab is brought into shared memory with coalesced
// loads; every thread owns eight consecutive tiles of it, so the two block-wide scans (prefix maximum of the last
// boundaries, prefix sum of the window counts) run once, and the three per-tile results leave through shared memory
// too.  What a slab needs from the slabs before it - the bout that is open at its start, and the number of windows
// before it - travels along a chain of two 64-bit words per slab in global memory (0 = not published yet, the host
// clears them): the first is published as soon as the slab knows its own last boundary, the second after its count.
// CTAs are dispatched in index order, so a CTA only ever waits for CTAs that are running or done.
constexpr int SEG_SCAN_THREADS = 1024;
constexpr int SEG_SCAN_PER = 8;
constexpr int SEG_SCAN_SLAB = SEG_SCAN_THREADS * SEG_SCAN_PER;
constexpr int SEG_SCAN_PAD = SEG_SCAN_SLAB + SEG_SCAN_SLAB / 8;          // one pad element per thread's run: 72-byte lane stride
__device__ __forceinline__ int seg_pad(int k) { return k + (k >> 3); }
__device__ __forceinline__ unsigned long long seg_chain_wait(const unsigned long long* p) {
    unsigned long long v;
    do { asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); } while (v == 0);
    return v;
}
__device__ __forceinline__ void seg_chain_post(unsigned long long* p, unsigned long long v) {
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__global__ void __launch_bounds__(SEG_SCAN_THREADS)
k_seg_scan(SegScratch sc, int64_t num_tiles, int64_t n, int32_t L, unsigned magic) {
    extern __shared__ __align__(16) unsigned char seg_scan_smem[];
    unsigned long long* s_sum = reinterpret_cast<unsigned long long*>(seg_scan_smem);           // [PAD]
    int64_t* s_out = reinterpret_cast<int64_t*>(seg_scan_smem + (size_t)SEG_SCAN_PAD * 8);      // [PAD] first open-bout window, then output base
    int32_t* s_pre = reinterpret_cast<int32_t*>(seg_scan_smem + (size_t)SEG_SCAN_PAD * 16);     // [PAD]
    __shared__ int s_wlast[32], s_wcnt[32];
    __shared__ int64_t s_carry_b, s_carry_n;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const unsigned uL = (unsigned)L;
    // (x / L, x % L) for x < L + 2048 by multiply-high while (L + 2048) * L < 2^32; plain division for longer windows
    const bool small_L = (unsigned long long)(uL + SEG_TILE) * uL < (1ULL << 32);
    auto udiv = [&](unsigned x) { return small_L ? __umulhi(x, magic) : x / uL; };
    const int64_t slab = blockIdx.x;
    const int64_t slab0 = slab * SEG_SCAN_SLAB;
    const int64_t in_slab = min((int64_t)SEG_SCAN_SLAB, num_tiles - slab0);
    const int64_t slab_pos = slab0 * SEG_TILE;            // positions inside a slab fit 32 bits (8192 tiles x 2048 samples)
    unsigned long long* chain_b = sc.chain + 2 * slab;    // open bout after this slab, + 2 (1 = none)
    unsigned long long* chain_n = chain_b + 1;            // windows up to the end of this slab, + 1
#pragma unroll
    for (int i = 0; i < SEG_SCAN_PER; i++) {              // eight independent loads in flight per thread
        const int k = threadIdx.x + i * SEG_SCAN_THREADS;
        s_sum[seg_pad(k)] = k < in_slab ? sc.summary[slab0 + k] : 0ULL;
    }
    __syncthreads();
    const int k0 = threadIdx.x * SEG_SCAN_PER;
    const int kp = seg_pad(k0);                           // this thread's eight entries are contiguous behind the pad
    unsigned long long sv[SEG_SCAN_PER];
#pragma unroll
    for (int j = 0; j < SEG_SCAN_PER; j++) sv[j] = s_sum[kp + j];
    // 1: last boundary inside this thread's tiles (slab-relative) -> exclusive prefix maximum over the threads
    int mine = -1;
#pragma unroll
    for (int j = 0; j < SEG_SCAN_PER; j++) {
        const int lb = (int)((unsigned)(sv[j] >> 16) & 0xffffu);
        if (lb) mine = (k0 + j) * SEG_TILE + lb - 1;
    }
    int incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) incl = max(incl, __shfl_up_sync(0xffffffffu, incl, d));
    if (lane == 31) s_wlast[wid] = incl;
    __syncthreads();
    int wbase = -1;                                       // latest boundary in the warps before this one
    {
        int wi = s_wlast[lane];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) wi = max(wi, __shfl_up_sync(0xffffffffu, wi, d));
        const int excl = __shfl_up_sync(0xffffffffu, wi, 1);
        wbase = __shfl_sync(0xffffffffu, lane == 0 ? -1 : excl, wid);
    }
    if (wid == 0) {
        int wi = s_wlast[lane];
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) wi = max(wi, __shfl_xor_sync(0xffffffffu, wi, d));
        if (lane == 0) {
            // what is open after this slab is its own last boundary (published at once), else whatever was open before it
            if (wi >= 0) seg_chain_post(chain_b, (unsigned long long)(slab_pos + wi + 2));
            const int64_t before = slab > 0 ? (int64_t)seg_chain_wait(chain_b - 2) - 2 : -1;
            if (wi < 0) seg_chain_post(chain_b, (unsigned long long)(before + 2));
            s_carry_b = before;
        }
    }
    __syncthreads();
    const int64_t carry_b = s_carry_b;
    int open_rel = __shfl_up_sync(0xffffffffu, incl, 1);
    if (lane == 0) open_rel = -1;
    open_rel = max(open_rel, wbase);
    const int64_t open = open_rel >= 0 ? slab_pos + open_rel : carry_b;
    // 2: per tile, the first window of the open bout that ends inside the tile and how many of them do; count this
    // thread's windows.  r = (tile start - bout start) mod L is carried from tile to tile, so the only 64-bit
    // division is the one that starts the thread's run.
    int total = 0;
    {
        const int64_t tb0 = slab_pos + (int64_t)k0 * SEG_TILE;
        bool have = open >= 0;
        unsigned r = have ? (unsigned)((tb0 - open) % L) : 0u;
#pragma unroll
        for (int j = 0; j < SEG_SCAN_PER; j++) {
            const unsigned long long s = sv[j];
            const unsigned fb = (unsigned)s & 0xffffu, lb = (unsigned)(s >> 16) & 0xffffu, nl = (unsigned)(s >> 32) & 0xffffu;
            const int64_t tb = tb0 + (int64_t)j * SEG_TILE;
            const unsigned span = fb ? fb - 1 : (unsigned)min((int64_t)SEG_TILE, n - tb);     // samples before the first boundary
            const unsigned pre = (have && k0 + j < in_slab) ? udiv(r + span) : 0u;
            s_out[kp + j] = tb - r;                        // start of the first such window
            s_pre[kp + j] = (int)pre;
            total += (int)(pre + nl);
            const unsigned x = lb ? (unsigned)(SEG_TILE - (lb - 1)) : r + SEG_TILE;
            have = have || lb != 0;
            r = x - udiv(x) * uL;
        }
    }
    int cincl = total;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, cincl, d);
        if (lane >= d) cincl += o;
    }
    if (lane == 31) s_wcnt[wid] = cincl;
    __syncthreads();
    int woff = 0;
    {
        const int x = s_wcnt[lane];
        int wi = x;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int o = __shfl_up_sync(0xffffffffu, wi, d);
            if (lane >= d) wi += o;
        }
        woff = __shfl_sync(0xffffffffu, wi - x, wid);
        const int slab_total = __shfl_sync(0xffffffffu, wi, 31);
        if (threadIdx.x == 0) {
            const int64_t before = slab > 0 ? (int64_t)seg_chain_wait(chain_n - 2) - 1 : 0;
            s_carry_n = before;
            seg_chain_post(chain_n, (unsigned long long)(before + slab_total + 1));
            if (slab == gridDim.x - 1) *sc.total = before + slab_total;
        }
    }
    // first open-bout windows out (coalesced) while thread 0 may still be waiting for the count before this slab
#pragma unroll
    for (int i = 0; i < SEG_SCAN_PER; i++) {
        const int k = threadIdx.x + i * SEG_SCAN_THREADS;
        if (k < in_slab) { sc.open_bout[slab0 + k] = s_out[seg_pad(k)]; sc.npre[slab0 + k] = s_pre[seg_pad(k)]; }
    }
    __syncthreads();
    int64_t off = s_carry_n + woff + cincl - total;
#pragma unroll
    for (int j = 0; j < SEG_SCAN_PER; j++) {
        const unsigned nl = (unsigned)(sv[j] >> 32) & 0xffffu;
        s_out[kp + j] = off;
        off += s_pre[kp + j] + nl;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < SEG_SCAN_PER; i++) {
        const int k = threadIdx.x + i * SEG_SCAN_THREADS;
        if (k < in_slab) sc.out_base[slab0 + k] = s_out[seg_pad(k)];
    }
}

__global__ void __launch_bounds__(256)
k_seg_emit(SegScratch sc, int32_t cap, int64_t num_tiles, int32_t L, const int32_t* __restrict__ user,
           const int32_t* __restrict__ act, int64_t* __restrict__ win_start, int32_t* __restrict__ win_user,
           int32_t* __restrict__ win_act) {
    const int64_t t = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    const int sub = threadIdx.x & 7;
    if (t >= num_tiles) return;
    const int64_t o = sc.out_base[t];
    const int pre = sc.npre[t];
    const int nl = (int)((sc.summary[t] >> 32) & 0xffffu);
    const int64_t tb = t * SEG_TILE;
    if (pre > 0) {
        const int64_t b = sc.open_bout[t];
        for (int k = sub; k < pre; k += 8) {
            const int64_t s = b + (int64_t)k * L;
            win_start[o + k] = s;
            if (win_user) win_user[o + k] = user[s];
            if (win_act) win_act[o + k] = act[s];
        }
    }
    const uint16_t* loc = sc.local + (size_t)t * cap;
    for (int k = sub; k < nl; k += 8) {
        const int64_t s = tb + loc[k];
        win_start[o + pre + k] = s;
        if (win_user) win_user[o + pre + k] = user[s];
        if (win_act) win_act[o + pre + k] = act[s];
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// features, window length 200
// ---------------------------------------------------------------------------------------------------------------------
constexpr int W4_L = 200;
constexpr int W4_PER_LANE = 25;            // consecutive samples per lane, 8 lanes per window
constexpr int W4_SLOT = 216;               // doubles per staged window slice: >= 202, == 8 (mod 16) so that the four
                                           // windows of a warp fall on disjoint bank groups when every lane reads 8 bytes
constexpr int W4_STAGES = 2;
constexpr int W4_WARPS = 4;
constexpr int W4_STAGE_DOUBLES = 4 * W4_SLOT;
constexpr size_t W4_SMEM_PER_WARP = (size_t)W4_STAGES * W4_STAGE_DOUBLES * sizeof(double) + W4_STAGES * sizeof(uint64_t);

__device__ __forceinline__ double rsqrt_approx_f64(double s) {
    double r;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(s));
    return r;
}
__device__ __noinline__ double precise_sqrt_v4(double s) { return sqrt(s); }

// the 25-bit mask of (v[j] OP t); written so that the compares land in predicates and are packed a few at a time
#define W4_MASK_GE(dst, t)                                                       \
    {                                                                            \
        unsigned _m = 0;                                                         \
        _Pragma("unroll") for (int j = 0; j < W4_PER_LANE; j++) _m |= (v[j] >= (t) ? 1u : 0u) << j; \
        dst = _m;                                                                \
    }

// (128 registers / 4 CTAs per SM was measured: the spills cost more than the extra warps bring, 1.30 vs 0.82 ms)
__global__ void __launch_bounds__(W4_WARPS * 32, 3)
k_window_features_v4(const double* __restrict__ x, const double* __restrict__ y, const double* __restrict__ z,
                     const int64_t* __restrict__ ts, const int64_t* __restrict__ win_start,
                     int64_t nw_host, const int64_t* __restrict__ nw_dev, int64_t n, int tma_ok,
                     double* __restrict__ out, int64_t ld_out,
                     const int32_t* __restrict__ user, const int32_t* __restrict__ act,
                     int32_t* __restrict__ win_user, int32_t* __restrict__ win_act) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, sub = lane & 7, grp = lane >> 3;
    unsigned char* mine = smem_raw + W4_SMEM_PER_WARP * warp;
    double* buf = reinterpret_cast<double*>(mine);
    uint64_t* bars = reinterpret_cast<uint64_t*>(buf + (size_t)W4_STAGES * W4_STAGE_DOUBLES);
    const int64_t nw = nw_dev ? *nw_dev : nw_host;
    const double qnan = __longlong_as_double(0x7ff8000000000000LL);

    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < W4_STAGES; s++) mbar_init(&bars[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();

    const int64_t nquads = (nw + 3) >> 2;
    const int64_t q0 = (int64_t)blockIdx.x * W4_WARPS + warp, qstep = (int64_t)gridDim.x * W4_WARPS;
    if (q0 >= nquads) return;
    auto start_of = [&](int64_t q) {                     // first sample of this lane group's window in quad q
        int64_t w = q * 4 + grp;
        if (w >= nw) w = nw - 1;
        return win_start[w];
    };
    // stage the slices of axis a (window starts stw) into stage sg: all lanes call
    auto issue = [&](int64_t stw, int a, int sg) {
        const double* src = a == 0 ? x : (a == 1 ? y : z);
        const int64_t s0 = stw & ~(int64_t)1;
        const bool bulk = tma_ok && s0 + 202 <= n;
        double* dst = buf + (size_t)sg * W4_STAGE_DOUBLES + grp * W4_SLOT;
        if (!bulk) {                                   // last window of the stream or unaligned columns: plain loads
            for (int i = sub; i < W4_L; i += 8) dst[(stw - s0) + i] = src[stw + i];
        }
        const unsigned bulk_mask = __ballot_sync(0xffffffffu, bulk);
        if (lane == 0) mbar_expect_tx(&bars[sg], 1616u * __popc(bulk_mask & 0x01010101u));
        __syncwarp();                                  // the plain stores and the expected byte count are in place
        if (sub == 0 && bulk) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            bulk_g2s(dst, src + s0, 1616u, &bars[sg]);
        }
    };
    // items are (quad, axis) pairs in order; the producer side runs W4_STAGES items ahead of the consumer side
    // The window starts are fetched one quad ahead on both sides, so that nobody waits for that load.
    int64_t pq = q0, pst = start_of(q0);
    int pa = 0, ps = 0;
    auto issue_next = [&]() {
        if (pq < nquads) {
            issue(pst, pa, ps);
            ps ^= 1;
            if (++pa == 3) { pa = 0; pq += qstep; if (pq < nquads) pst = start_of(pq); }
        }
    };
    static_assert(W4_STAGES == 2, "stage index toggles");
    issue_next(); issue_next();

    uint32_t phase_bits = 0;
    double r2[W4_PER_LANE];
    int64_t st = 0;
    double* dstrow = nullptr;
    bool live = false;
    // PEAK of the previous item: its two timestamps are requested at the end of that item and used in the middle of
    // the next one, so that the round trip to memory is not waited for
    int64_t pk_ta = 0, pk_tb = 0;
    int pk_cnt = 2;
    double* pk_ptr = nullptr;
    auto flush_peak = [&]() {
        if (pk_ptr) {
            *pk_ptr = ((double)(pk_tb - pk_ta) / (double)(pk_cnt - 1)) / 1.0e6;
            pk_ptr = nullptr;
        }
    };

    int s = 0;
    int64_t st_next = start_of(q0);
    for (int64_t q = q0; q < nquads; q += qstep) {
      {
        int64_t w = q * 4 + grp;
        live = w < nw;
        if (!live) w = nw - 1;
        st = st_next;
        if (q + qstep < nquads) st_next = start_of(q + qstep);
        dstrow = out + w * ld_out;
        // the window's user / activity, when the caller left them to this kernel (har_window_transform): two gathers
        // per window that would otherwise be a pass of their own
        if (live && win_user && sub == 6) win_user[w] = __ldg(user + st);
        if (live && win_act && sub == 7) win_act[w] = __ldg(act + st);
      }
#pragma unroll 1
      for (int a = 0; a < 3; a++, s ^= 1) {
        mbar_wait(&bars[s], (phase_bits >> s) & 1u);
        phase_bits ^= 1u << s;
        const double* p = buf + (size_t)s * W4_STAGE_DOUBLES + grp * W4_SLOT + (st & 1) + sub * W4_PER_LANE;
        double v[W4_PER_LANE];
#pragma unroll
        for (int j = 0; j < W4_PER_LANE; j++) v[j] = p[j];
        __syncwarp();
        issue_next();                                           // the slot is free: its samples are in registers
        if (a == 0) {
#pragma unroll
            for (int j = 0; j < W4_PER_LANE; j++) r2[j] = 0.0;
        }

        // ---- local maxima: up[j] = v[j+1] > v[j]; a peak is up[j-1] && !up[j] (no NaN in the stream) ----
        const double nxt = __shfl_down_sync(0xffffffffu, v[0], 1);       // first sample of the next lane
        unsigned up = 0;
        double sum0 = 0.0, sum1 = 0.0, hi = v[24], lo = v[24];
        int hpair = 12;                                            // pair that holds this lane's maximum (12: the last sample)
#pragma unroll
        for (int j = 0; j < W4_PER_LANE; j++) {
            const double r = j + 1 < W4_PER_LANE ? v[j + 1] : nxt;
            const bool g = r > v[j];
            up |= (g ? 1u : 0u) << j;
            if ((j & 1) == 0 && j + 1 < W4_PER_LANE) {           // one compare orders the pair: larger to hi, smaller to lo
                const double pmx = g ? r : v[j], pmn = g ? v[j] : r;
                const bool nh = pmx > hi;
                hi = nh ? pmx : hi;
                hpair = nh ? (j >> 1) : hpair;
                lo = pmn < lo ? pmn : lo;
            }
            if (j & 1) sum1 += v[j]; else sum0 += v[j];
        }
        double sum = sum0 + sum1;
        const double lane_hi = hi;
        const unsigned up_prev = __shfl_up_sync(0xffffffffu, up, 1);
        unsigned pk = ((up << 1) | (sub ? (up_prev >> 24) & 1u : 0u)) & ~up & 0x1ffffffu;
        if (sub == 0) pk &= ~1u;                                   // first sample of the window
        if (sub == 7) pk &= ~(1u << 24);                           // last sample of the window
#pragma unroll
        for (int m = 4; m >= 1; m >>= 1) {
            sum += shfl_xor_d(sum, m);
            const double ol = shfl_xor_d(lo, m), oh = shfl_xor_d(hi, m);
            lo = ol < lo ? ol : lo;
            hi = oh > hi ? oh : hi;
        }
        // work that does not need the window-wide sum / minimum / maximum sits here, behind the shuffles above and
        // under the division below: the squares for RESULTANT and the test "is the maximum one of the peaks"
#pragma unroll
        for (int j = 0; j < W4_PER_LANE; j++) r2[j] = fma(v[j], v[j], r2[j]);
        const double mean = sum * 0.005;                           // within an ulp of sum / 200 (AVG is compared to 1e-6)
        const double scale = hi > lo ? 10.0 / (hi - lo) : 0.0;     // exact: decides the bins
        // "the window maximum is one of the peaks": looked up at one place per lane, the sample that gave the lane its
        // maximum.  A yes is always right; a no can be wrong when the maximum occurs twice in a lane (the exact search
        // below then finds it anyway).
        const int hpos = hpair < 12 ? 2 * hpair + (int)((up >> (2 * hpair)) & 1u) : 24;
        const bool lane_top = lane_hi == hi && ((pk >> hpos) & 1u);

        // ---- deviations and the bin counts.  Eleven 4-bit fields per accumulator (bin 10 = the samples equal to the
        // maximum, folded into bin 9 at the end): an accumulator takes 13 samples at most, so no field overflows and the
        // sum stays an exact integer below 2^44; no clamp per sample. ----
        double sq = 0.0, sq1 = 0.0, ab = 0.0, ab1 = 0.0, packed = 0.0, packed1 = 0.0;
#pragma unroll
        for (int j = 0; j < W4_PER_LANE; j++) {
            const double d = v[j] - mean;
            const double q = __dmul_rn(v[j] - lo, scale);
            const int bq = __double2loint(__dadd_rd(q, 4503599627370496.0));               // floor(q), 0 <= q <= 10
            const double one = __hiloint2double((1023 << 20) + (bq << 22), 0);             // 2^(4 * bin)
            if (j & 1) { sq1 = fma(d, d, sq1); ab1 += fabs(d); packed1 += one; }
            else { sq = fma(d, d, sq); ab += fabs(d); packed += one; }
        }
        sq += sq1; ab += ab1;
        int npk = __popc(pk);
        // 4-bit fields -> 8-bit fields (a window has 200 samples): even bins 0,2,4,6 | 8,10 and odd bins 1,3,5,7 | 9
        const unsigned long long pa = (unsigned long long)__double2ll_rn(packed), pb = (unsigned long long)__double2ll_rn(packed1);
        const unsigned long long M4 = 0x0f0f0f0f0f0f0f0full;
        const unsigned long long ev = (pa & M4) + (pb & M4), od = ((pa >> 4) & M4) + ((pb >> 4) & M4);
        unsigned e_lo = (unsigned)ev, e_hi = (unsigned)(ev >> 32), o_lo = (unsigned)od, o_hi = (unsigned)(od >> 32);
        int top_is_peak = lane_top ? 1 : 0;
#pragma unroll
        for (int m = 4; m >= 1; m >>= 1) {
            sq += shfl_xor_d(sq, m);
            ab += shfl_xor_d(ab, m);
        }
        // (integer warp reductions with a different member mask per window group run group by group: measured slower
        // than three shuffle steps)
        unsigned npk_top = (unsigned)npk | ((unsigned)top_is_peak << 16);
#pragma unroll
        for (int m = 4; m >= 1; m >>= 1) {
            npk_top += __shfl_xor_sync(0xffffffffu, npk_top, m);
            e_lo += __shfl_xor_sync(0xffffffffu, e_lo, m);
            e_hi += __shfl_xor_sync(0xffffffffu, e_hi, m);
            o_lo += __shfl_xor_sync(0xffffffffu, o_lo, m);
            o_hi += __shfl_xor_sync(0xffffffffu, o_hi, m);
        }
        npk = (int)(npk_top & 0xffffu);
        top_is_peak = (npk_top >> 16) != 0;
        // ---- highest peak: the window maximum when it sits on a peak (almost always), else a max over the peaks ----
        double pm = hi;
        const bool has = npk >= 3;
        if (__any_sync(0xffffffffu, has && !top_is_peak)) {
            double m2 = -INFINITY;
#pragma unroll
            for (int j = 0; j < W4_PER_LANE; j++) m2 = ((pk >> j) & 1u) && v[j] > m2 ? v[j] : m2;
#pragma unroll
            for (int m = 4; m >= 1; m >>= 1) { const double o = shfl_xor_d(m2, m); m2 = o > m2 ? o : m2; }
            if (!top_is_peak) pm = m2;
        }
        // ---- threshold ladder tau_k = P - ((k+1)*0.1)*(P - lo), k = 0..8, then lo: first rung admitting >= 3 peaks ----
        unsigned adm = pk;                                         // rung 9 (tau = lo) admits every peak
        {
            const double R = __dsub_rn(pm, lo);
            bool open = has;
#pragma unroll 1
            for (int kk = 0; kk < 9; kk++) {
                if (!__any_sync(0xffffffffu, open)) break;
                const double t = __dsub_rn(pm, __dmul_rn(__dmul_rn((double)(kk + 1), 0.1), R));
                unsigned mt;
                W4_MASK_GE(mt, t);
                mt &= pk;
                int c3 = __popc(mt);
#pragma unroll
                for (int m = 4; m >= 1; m >>= 1) c3 += __shfl_xor_sync(0xffffffffu, c3, m);
                if (open && c3 >= 3) { adm = mt; open = false; }
            }
        }
        // first / last admitted sample and their number: (199 - first) and last share one word (both maxima of 8-bit values)
        int fl = adm ? ((199 - (sub * W4_PER_LANE + (__ffs(adm) - 1))) << 16) | (sub * W4_PER_LANE + (31 - __clz(adm))) : 0;
        int cnt = __popc(adm);
#pragma unroll
        for (int m = 4; m >= 1; m >>= 1) {
            const int o = __shfl_xor_sync(0xffffffffu, fl, m);
            fl = max(fl & 0xffff0000, o & 0xffff0000) | max(fl & 0xffff, o & 0xffff);
            cnt += __shfl_xor_sync(0xffffffffu, cnt, m);
        }
        const int first = 199 - (fl >> 16), last = fl & 0xffff;
        // ---- the 14 outputs of this axis: every lane one bin, then bins 8-9 / AVG / ABSDEV / STDDEV; PEAK is deferred ----
        flush_peak();                                              // the previous item's PEAK: its timestamps have arrived
        if (live) {
            // bin fraction k / 200 exactly: q = k * 0.005 corrected by one fused step (checked for k = 0..200 against the
            // correctly rounded quotient; the plain product is off by an ulp for 24 of them)
            auto frac200 = [](unsigned k) {
                const double a = (double)k, q = a * 0.005;
                return fma(fma(-q, 200.0, a), 0.005, q);
            };
            // lane `sub` writes bin `sub` (0..7): even bins sit in e_lo, odd ones in o_lo, byte sub / 2
            dstrow[a * 10 + sub] = frac200((((sub & 1) ? o_lo : e_lo) >> (8 * (sub >> 1))) & 0xffu);
            if (sub < 2) {
                // bin 8, and bin 9 with the samples equal to the maximum (field 10)
                dstrow[a * 10 + 8 + sub] = frac200(sub ? (o_hi & 0xffu) + ((e_hi >> 8) & 0xffu) : (e_hi & 0xffu));
            } else if (sub == 2) {
                dstrow[30 + a] = mean;
            } else if (sub == 3) {
                if (has) {                                         // admitted peaks: first / last timestamp, used later
                    pk_ta = ts[st + first]; pk_tb = ts[st + last]; pk_cnt = cnt; pk_ptr = dstrow + 33 + a;
                } else {
                    dstrow[33 + a] = qnan;
                }
            } else if (sub == 4) {
                dstrow[36 + a] = ab * 0.005;
            } else if (sub == 5) {
                const double var = sq * 0.005;
                const double r = rsqrt_approx_f64(var);
                double sd = var * r;
                sd = fma(0.5 * r, fma(-sd, sd, var), sd);
                if (!(var > 1e-280 && var < 1e280)) sd = precise_sqrt_v4(var);
                dstrow[39 + a] = sd;
            }
        }
        // ---- RESULTANT after the third axis: sqrt by the fp64 rsqrt seed + one Newton step (rel. error < 1e-13) ----
        if (a == 2) {
            double rs = 0.0, rs1 = 0.0;
#pragma unroll
            for (int j = 0; j < W4_PER_LANE; j++) {
                const double s2 = r2[j];
                const double r = rsqrt_approx_f64(s2);
                double yv = s2 * r;
                yv = fma(0.5 * r, fma(-yv, yv, s2), yv);
                if (j & 1) rs1 += yv; else rs += yv;
            }
            rs += rs1;
            // a zero or denormal square (seed +inf) or an infinite one (seed 0) turns its term into NaN, huge ones overflow:
            // one test on the sum catches them all
            const bool odd = !(rs < 1e300);
            if (__any_sync(0xffffffffu, odd)) {                    // the library sqrt
                rs = 0.0;
#pragma unroll
                for (int j = 0; j < W4_PER_LANE; j++) rs += precise_sqrt_v4(r2[j]);
            }
#pragma unroll
            for (int m = 4; m >= 1; m >>= 1) rs += shfl_xor_d(rs, m);
            if (live && sub == 0) dstrow[42] = rs * 0.005;
        }
      }
    }
    flush_peak();
}
