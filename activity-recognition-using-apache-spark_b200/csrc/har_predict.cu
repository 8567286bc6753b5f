// K10: batched forest predict, one row per thread.
//
// Stands in for {DecisionTree,RandomForest}ClassificationModel.transform reached from
// Main/main.py:305 and :486: predictRaw (walk every tree: continuous `x[f] <= threshold`, categorical
// `leftCategories.contains(x[f])`), raw2probabilityInPlace (raw / sum(raw)), raw2prediction (first
// argmax).  RF votes are accumulated in tree order with plain fp64 adds of the pre-divided leaf
// vectors counts/total, which is the same sequence of IEEE operations MLlib performs.
//
// Device layout: 16-byte nodes {feature|flags, left child (right = left + 1) | leaf index, threshold or
// category mask}, children adjacent so one 128-bit load serves a visit; leaf vectors f64 [leaves][C].
#include <algorithm>
#include <cstdlib>
#include <deque>

#include "har_common.cuh"
#include "har_forest.h"
#include "har_pnode.cuh"

namespace {

using har::PNode; using har::P_CAT; using har::P_DEFLEFT; using har::P_FMASK;
constexpr int PRED_THREADS = 128;

template <bool STAGE>
__global__ void __launch_bounds__(PRED_THREADS)
k_predict(const PNode* __restrict__ nodes, const int64_t* __restrict__ roots, int T, const double* __restrict__ leaf,
          int C, const double* __restrict__ X, int64_t n, int64_t ld, int F, int fstride,
          double* __restrict__ raw, double* __restrict__ prob, double* __restrict__ pred) {
    extern __shared__ __align__(16) double psm[];
    double* acc = psm;                                   // [C][PRED_THREADS]
    double* xs = psm + (size_t)C * PRED_THREADS;         // [PRED_THREADS][fstride] when STAGE
    const int64_t row0 = (int64_t)blockIdx.x * PRED_THREADS;
    const int64_t i = row0 + threadIdx.x;
    if (STAGE) {
        const int64_t rows_here = min((int64_t)PRED_THREADS, n - row0);
        for (int64_t t = threadIdx.x; t < rows_here * F; t += PRED_THREADS) {
            int64_t r = t / F;
            int f = (int)(t - r * F);
            xs[r * fstride + f] = X[(row0 + r) * ld + f];
        }
        __syncthreads();
    }
    if (i >= n) return;
    const double* x = STAGE ? xs + (size_t)threadIdx.x * fstride : X + i * ld;
    for (int c = 0; c < C; c++) acc[c * PRED_THREADS + threadIdx.x] = 0.0;
    // Four trees are walked at once: a walk is a chain of dependent node loads (L2 latency each), so interleaving
    // independent walks multiplies the loads in flight per thread.  The leaf vectors are then added in tree order,
    // which keeps MLlib's sequence of fp64 additions.
    auto step = [&](PNode& nd) {
        const double v = x[nd.feature & P_FMASK];
        bool go_left;
        if (nd.feature & P_CAT) {
            const bool in_dom = v >= 0.0 && v < 64.0 && v == floor(v);
            go_left = in_dom ? (bool)((nd.mask >> (int)v) & 1ULL) : (bool)(nd.feature & P_DEFLEFT);
        } else {
            go_left = v <= nd.thr;
        }
        nd = nodes[nd.left + (go_left ? 0 : 1)];
    };
    int t = 0;
    for (; t + 4 <= T; t += 4) {
        PNode n0 = nodes[roots[t]], n1 = nodes[roots[t + 1]], n2 = nodes[roots[t + 2]], n3 = nodes[roots[t + 3]];
        while ((n0.feature >= 0) | (n1.feature >= 0) | (n2.feature >= 0) | (n3.feature >= 0)) {
            if (n0.feature >= 0) step(n0);
            if (n1.feature >= 0) step(n1);
            if (n2.feature >= 0) step(n2);
            if (n3.feature >= 0) step(n3);
        }
        const double* l0 = leaf + (int64_t)n0.left * C; const double* l1 = leaf + (int64_t)n1.left * C;
        const double* l2 = leaf + (int64_t)n2.left * C; const double* l3 = leaf + (int64_t)n3.left * C;
        for (int c = 0; c < C; c++) {
            double a = acc[c * PRED_THREADS + threadIdx.x];
            a += l0[c]; a += l1[c]; a += l2[c]; a += l3[c];
            acc[c * PRED_THREADS + threadIdx.x] = a;
        }
    }
    for (; t < T; t++) {
        PNode nd = nodes[roots[t]];
        while (nd.feature >= 0) step(nd);
        const double* lv = leaf + (int64_t)nd.left * C;
        for (int c = 0; c < C; c++) acc[c * PRED_THREADS + threadIdx.x] += lv[c];
    }
    double sum = 0.0;
    for (int c = 0; c < C; c++) sum += acc[c * PRED_THREADS + threadIdx.x];
    int best = 0;
    double mv = acc[threadIdx.x];
    for (int c = 0; c < C; c++) {
        const double a = acc[c * PRED_THREADS + threadIdx.x];
        if (raw) raw[i * C + c] = a;
        if (prob) prob[i * C + c] = sum != 0.0 ? a / sum : a;
        if (a > mv) { mv = a; best = c; }
    }
    if (pred) pred[i] = (double)best;
}

// ---- binned predict ------------------------------------------------------------------------------------------------
using har::CNode; using har::C_LEAF; using har::C_CAT; using har::C_DEFLEFT; using har::BIN_OUTSIDE;

// PNode -> CNode.  A continuous node's threshold must be one of its feature's candidates (exact, in Double.compare
// order) and must not be -0.0 (TreePoint.findBin orders -0.0 before +0.0, `x <= -0.0` does not); otherwise *bad is set
// and the forest keeps the f64 walk.
__global__ void k_compact_nodes(const PNode* __restrict__ in, int64_t np, const double* __restrict__ thr,
                                const int32_t* __restrict__ nsplits, int max_bins, CNode* __restrict__ out,
                                unsigned long long* __restrict__ cmask, uint2* __restrict__ walk, uint32_t* __restrict__ leaf_of,
                                int* __restrict__ bad /* [0] not binnable, [1] categorical nodes */) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= np) return;
    const PNode p = in[i];
    CNode c;
    c.feature = 0; c.split = 0; c.flags = 0; c.left = (uint32_t)p.left;
    unsigned long long m = 0;
    if (p.feature < 0) {
        c.flags = C_LEAF;
    } else if (p.feature & P_CAT) {
        c.feature = (uint16_t)(p.feature & P_FMASK);
        c.flags = C_CAT | ((p.feature & P_DEFLEFT) ? C_DEFLEFT : 0);
        m = p.mask;
    } else {
        const int f = p.feature & P_FMASK;
        c.feature = (uint16_t)f;
        const int ns = nsplits[f];
        const int b = har::java_find_bin(thr + (int64_t)f * max_bins, ns, p.thr);
        const bool exact = b < ns && har::jd_key(thr[(int64_t)f * max_bins + b]) == har::jd_key(p.thr);
        const bool neg_zero = __double_as_longlong(p.thr) == (long long)0x8000000000000000ULL;
        if (!exact || neg_zero || b > 255) atomicExch(bad, 1);
        c.split = (uint8_t)b;
    }
    out[i] = c;
    cmask[i] = m;
    // walk table: {feature | split << 24, left child}; a leaf tests "bin of feature 0 <= 255" and is its own left child,
    // so a walk needs no leaf test: it simply stops moving
    uint2 w;
    if (c.flags & C_LEAF) { w.x = 255u << 24; w.y = (uint32_t)i; leaf_of[i] = c.left; }
    else { w.x = (uint32_t)c.feature | ((uint32_t)c.split << 24); w.y = c.left; leaf_of[i] = 0; }
    walk[i] = w;
    if (c.flags & C_CAT) atomicAdd(bad + 1, 1);
}

constexpr int PB_CMAX = 8;              // class votes live in registers up to this many classes

// R rows per thread, NT threads.  The CTA first bins its NT * R rows (coalesced f64 reads, binary search among <= 255 candidates
// read through L1, one byte per feature into shared memory at an odd word stride), then every thread walks all trees
// over its rows' bytes, K trees at a time: R * K independent pointer chases per thread hide the gather latency, and
// because every CTA of an SM is inside the same K trees at about the same time, those trees' nodes (32 KB each at
// depth 12) stay in L1.  Leaf vectors are added in tree order (MLlib's sequence of fp64 additions).
template <int NT, int R, int K, bool SYNC>
__global__ void __launch_bounds__(NT)
k_predict_bins(const CNode* __restrict__ nodes, const unsigned long long* __restrict__ cmask, const int64_t* __restrict__ roots,
               int T, const double* __restrict__ leaf, int C, const double* __restrict__ X, int64_t n, int64_t ld, int F,
               const double* __restrict__ thr, const int32_t* __restrict__ nsplits, int max_bins, int bstride,
               double* __restrict__ raw, double* __restrict__ prob, double* __restrict__ pred) {
    extern __shared__ __align__(16) unsigned char pbs[];
    uint8_t* xs = pbs;                                                   // [NT * R][bstride]
    constexpr int ROWS = NT * R;
    const int64_t row0 = (int64_t)blockIdx.x * ROWS;
    const int rows_here = (int)min((int64_t)ROWS, n - row0);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int r = warp; r < rows_here; r += NT / 32) {
        const double* xr = X + (row0 + r) * ld;
        uint8_t* br = xs + (size_t)r * bstride;
        for (int f = lane; f < F; f += 32) {
            const double x = xr[f];
            const int ns = nsplits[f];
            int b;
            if (ns > 0) b = har::java_find_bin(thr + (int64_t)f * max_bins, ns, x);
            else b = (x >= 0.0 && x < 64.0 && x == floor(x)) ? (int)x : BIN_OUTSIDE;      // category (features never split on: unused)
            br[f] = (uint8_t)b;
        }
    }
    __syncthreads();
    // thread t owns rows t, t + 128, ... of the tile (rows past the end walk row 0 of the tile and are not written)
    const uint8_t* x[R];
    bool live[R];
#pragma unroll
    for (int r = 0; r < R; r++) {
        const int lr = threadIdx.x + r * NT;
        live[r] = lr < rows_here;
        x[r] = xs + (size_t)(live[r] ? lr : 0) * bstride;
    }
    double acc[R][PB_CMAX];
#pragma unroll
    for (int r = 0; r < R; r++)
#pragma unroll
        for (int c = 0; c < PB_CMAX; c++) acc[r][c] = 0.0;
    auto step = [&](const uint8_t* xr, CNode& nd, uint32_t& at) {
        const unsigned b = xr[nd.feature];
        bool go_left;
        if (nd.flags & C_CAT) go_left = b != BIN_OUTSIDE ? (bool)((cmask[at] >> b) & 1ULL) : (bool)(nd.flags & C_DEFLEFT);
        else go_left = b <= nd.split;
        at = nd.left + (go_left ? 0u : 1u);
        nd = nodes[at];
    };
    int t = 0;
    for (; t + K <= T; t += K) {
        // one CTA per SM: a barrier per tree group keeps every warp of the SM inside the same K trees, whose nodes
        // then stay in L1 (without it warps drift apart by many trees and the walks go to L2)
        if (SYNC) __syncthreads();
        CNode nd[R][K];
        uint32_t at[R][K];
#pragma unroll
        for (int k = 0; k < K; k++) {
            const uint32_t rt = (uint32_t)roots[t + k];
            const CNode rn = nodes[rt];
#pragma unroll
            for (int r = 0; r < R; r++) { at[r][k] = rt; nd[r][k] = rn; }
        }
        for (;;) {
            unsigned all_leaf = C_LEAF;
#pragma unroll
            for (int r = 0; r < R; r++)
#pragma unroll
                for (int k = 0; k < K; k++) all_leaf &= nd[r][k].flags;
            if (all_leaf) break;
#pragma unroll
            for (int r = 0; r < R; r++)
#pragma unroll
                for (int k = 0; k < K; k++)
                    if (!(nd[r][k].flags & C_LEAF)) step(x[r], nd[r][k], at[r][k]);
        }
#pragma unroll
        for (int k = 0; k < K; k++)                                       // tree order
#pragma unroll
            for (int r = 0; r < R; r++) {
                const double* lv = leaf + (int64_t)nd[r][k].left * C;
#pragma unroll
                for (int c = 0; c < PB_CMAX; c++) if (c < C) acc[r][c] += lv[c];
            }
    }
    for (; t < T; t++) {
        const uint32_t rt = (uint32_t)roots[t];
#pragma unroll
        for (int r = 0; r < R; r++) {
            uint32_t a0 = rt;
            CNode n0 = nodes[a0];
            while (!(n0.flags & C_LEAF)) step(x[r], n0, a0);
            const double* lv = leaf + (int64_t)n0.left * C;
#pragma unroll
            for (int c = 0; c < PB_CMAX; c++) if (c < C) acc[r][c] += lv[c];
        }
    }
#pragma unroll
    for (int r = 0; r < R; r++) {
        if (!live[r]) continue;
        const int64_t i = row0 + threadIdx.x + r * NT;
        double sum = 0.0;
#pragma unroll
        for (int c = 0; c < PB_CMAX; c++) if (c < C) sum += acc[r][c];
        int best = 0;
        double mv = acc[r][0];
#pragma unroll
        for (int c = 0; c < PB_CMAX; c++) {
            if (c < C) {
                const double a = acc[r][c];
                if (raw) raw[i * C + c] = a;
                if (prob) prob[i * C + c] = sum != 0.0 ? a / sum : a;
                if (a > mv) { mv = a; best = c; }
            }
        }
        if (pred) pred[i] = (double)best;
    }
}

// Forests without categorical splits (the window-feature table, configs 3-5): the same tiling as k_predict_bins, but
// the walk is branch-free - 8-byte nodes {feature | split << 24, left child} whose leaves loop onto themselves, so a
// visit is: mask, add, LDS.U8, shift, compare, add, address, LDG.64.  R rows per thread advance together through one
// tree; every fourth level the warp votes whether all its walks have stopped.  CM bounds the classes held in registers.
template <int NT, int R, int CM>
__global__ void __launch_bounds__(NT)
k_predict_walk(const uint2* __restrict__ walk, const uint32_t* __restrict__ leaf_of, const int64_t* __restrict__ roots, int T,
               const double* __restrict__ leaf, int C, const double* __restrict__ X, int64_t n, int64_t ld, int F,
               const double* __restrict__ thr, const int32_t* __restrict__ nsplits, int max_bins, int bstride,
               double* __restrict__ raw, double* __restrict__ prob, double* __restrict__ pred) {
    extern __shared__ __align__(16) unsigned char pbs[];
    uint8_t* xs = pbs;                                                   // [NT * R][bstride]
    constexpr int ROWS = NT * R;
    const int64_t row0 = (int64_t)blockIdx.x * ROWS;
    const int rows_here = (int)min((int64_t)ROWS, n - row0);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int r = warp; r < rows_here; r += NT / 32) {
        const double* xr = X + (row0 + r) * ld;
        uint8_t* br = xs + (size_t)r * bstride;
        for (int f = lane; f < F; f += 32) {
            const int ns = nsplits[f];
            br[f] = ns > 0 ? (uint8_t)har::java_find_bin(thr + (int64_t)f * max_bins, ns, xr[f]) : (uint8_t)0;
        }
    }
    __syncthreads();
    const uint8_t* x[R];
    bool live[R];
#pragma unroll
    for (int r = 0; r < R; r++) {
        const int lr = threadIdx.x + r * NT;
        live[r] = lr < rows_here;
        x[r] = xs + (size_t)(live[r] ? lr : 0) * bstride;
    }
    double acc[R][CM];
#pragma unroll
    for (int r = 0; r < R; r++)
#pragma unroll
        for (int c = 0; c < CM; c++) acc[r][c] = 0.0;
    for (int t = 0; t < T; t++) {
        const uint32_t rt = (uint32_t)roots[t];
        const uint2 rn = __ldg(walk + rt);
        uint2 nd[R];
        uint32_t at[R];
#pragma unroll
        for (int r = 0; r < R; r++) { nd[r] = rn; at[r] = rt; }
        for (;;) {
#pragma unroll
            for (int lv = 0; lv < 4; lv++) {
#pragma unroll
                for (int r = 0; r < R; r++) {
                    const unsigned b = x[r][nd[r].x & 0xffffu];
                    at[r] = nd[r].y + (b > (nd[r].x >> 24) ? 1u : 0u);
                    nd[r] = __ldg(walk + at[r]);
                }
            }
            bool done = true;
#pragma unroll
            for (int r = 0; r < R; r++) done = done && nd[r].y == at[r];
            if (__all_sync(0xffffffffu, done)) break;
        }
#pragma unroll
        for (int r = 0; r < R; r++) {
            const double* lv = leaf + (int64_t)leaf_of[at[r]] * C;
#pragma unroll
            for (int c = 0; c < CM; c++) if (c < C) acc[r][c] += lv[c];
        }
    }
#pragma unroll
    for (int r = 0; r < R; r++) {
        if (!live[r]) continue;
        const int64_t i = row0 + threadIdx.x + r * NT;
        double sum = 0.0;
#pragma unroll
        for (int c = 0; c < CM; c++) if (c < C) sum += acc[r][c];
        int best = 0;
        double mv = acc[r][0];
#pragma unroll
        for (int c = 0; c < CM; c++) {
            if (c < C) {
                const double a = acc[r][c];
                if (raw) raw[i * C + c] = a;
                if (prob) prob[i * C + c] = sum != 0.0 ? a / sum : a;
                if (a > mv) { mv = a; best = c; }
            }
        }
        if (pred) pred[i] = (double)best;
    }
}

__global__ void k_csr_to_dense(const int64_t* __restrict__ indptr, const int32_t* __restrict__ indices,
                               const double* __restrict__ values, int64_t n, double* __restrict__ out, int64_t ld) {
    // one warp per row
    int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (r >= n) return;
    for (int64_t k = indptr[r] + lane; k < indptr[r + 1]; k += 32) out[r * ld + indices[k]] = values[k];
}

void free_device_tables(har_forest* f) {
    if (f->d_nodes) cudaFree(f->d_nodes);
    if (f->d_leaf) cudaFree(f->d_leaf);
    if (f->d_roots) cudaFree(f->d_roots);
    if (f->d_slot_stats) cudaFree(f->d_slot_stats);
    f->d_nodes = nullptr; f->d_leaf = nullptr; f->d_roots = nullptr; f->d_slot_stats = nullptr; f->num_slots = 0;
    f->dev = -1; f->leaf_mode = -1;
    if (f->d_cnodes) cudaFree(f->d_cnodes);
    if (f->d_cmask) cudaFree(f->d_cmask);
    if (f->d_bin_thr) cudaFree(f->d_bin_thr);
    if (f->d_bin_ns) cudaFree(f->d_bin_ns);
    if (f->d_walk) cudaFree(f->d_walk);
    if (f->d_leaf_of) cudaFree(f->d_leaf_of);
    f->d_walk = nullptr; f->d_leaf_of = nullptr;
    f->d_cnodes = nullptr; f->d_cmask = nullptr; f->d_bin_thr = nullptr; f->d_bin_ns = nullptr; f->num_cnodes = 0; f->binnable = -1;
}

// Compact tables of the binned path, derived once from the PNode tables (after build_device_tables).  Sets
// f->binnable; 0 means "use the f64 walk" and is not an error.
int build_compact_tables(har_forest* f, cudaStream_t s) {
    if (f->binnable >= 0) return HAR_OK;
    // HAR_PREDICT_BINS=0 (read when a forest's tables are built, so tests can flip it per forest) keeps the f64 walk
    const char* env_bins = std::getenv("HAR_PREDICT_BINS");
    const bool off = env_bins && std::atoi(env_bins) == 0;
    const int64_t np = f->d_slot_stats ? f->num_slots : (int64_t)f->feature.size();
    const int bstride = 4 * (((f->F + 3) >> 2) | 1);           // bytes per staged row: a multiple of 4 with an odd word count
    if (off || f->nsplits.empty() || f->max_bins < 2 || f->max_bins > 256 || f->F >= 65535 || np <= 0 || np >= (1LL << 32) ||
        (size_t)128 * bstride > 96 * 1024) {
        f->binnable = 0;
        return HAR_OK;
    }
    har::pool_setup();
    HAR_CUDA(cudaMallocAsync(&f->d_cnodes, sizeof(CNode) * np, s));
    HAR_CUDA(cudaMallocAsync((void**)&f->d_cmask, sizeof(unsigned long long) * np, s));
    HAR_CUDA(cudaMallocAsync(&f->d_walk, sizeof(uint2) * np, s));
    HAR_CUDA(cudaMallocAsync((void**)&f->d_leaf_of, sizeof(uint32_t) * np, s));
    HAR_CUDA(cudaMallocAsync((void**)&f->d_bin_thr, sizeof(double) * f->thresholds.size(), s));
    HAR_CUDA(cudaMallocAsync((void**)&f->d_bin_ns, sizeof(int32_t) * f->F, s));
    har::DevBuf<int> bad;
    HAR_CUDA(bad.alloc(2, s));
    HAR_CUDA(cudaMemsetAsync(bad.p, 0, 2 * sizeof(int), s));
    HAR_CUDA(cudaMemcpyAsync(f->d_bin_thr, f->thresholds.data(), sizeof(double) * f->thresholds.size(), cudaMemcpyHostToDevice, s));
    HAR_CUDA(cudaMemcpyAsync(f->d_bin_ns, f->nsplits.data(), sizeof(int32_t) * f->F, cudaMemcpyHostToDevice, s));
    k_compact_nodes<<<(unsigned)har::ceil_div(np, 256), 256, 0, s>>>((const PNode*)f->d_nodes, np, f->d_bin_thr, f->d_bin_ns, f->max_bins,
                                                                     (CNode*)f->d_cnodes, f->d_cmask, (uint2*)f->d_walk, f->d_leaf_of, bad.p);
    HAR_LAUNCH_CHECK();
    int h_bad[2] = {0, 0};
    HAR_CUDA(cudaMemcpyAsync(h_bad, bad.p, 2 * sizeof(int), cudaMemcpyDeviceToHost, s));
    HAR_CUDA(cudaStreamSynchronize(s));      // thresholds / nsplits are host vectors of the handle; once per forest
    f->num_cnodes = np;
    f->binnable = h_bad[0] ? 0 : 1;
    f->has_cat_nodes = h_bad[1] != 0;
    return HAR_OK;
}

int build_device_tables(har_forest* f, int is_forest, cudaStream_t s) {
    int dev = 0;
    HAR_CUDA(cudaGetDevice(&dev));
    if (f->d_nodes && f->dev == dev && f->leaf_mode == is_forest) return HAR_OK;
    // the leaf vectors are about to be rewritten (or freed): a predict of the other mode may still be reading them on
    // another stream
    if (f->ev_pred) HAR_CUDA(cudaStreamWaitEvent(s, (cudaEvent_t)f->ev_pred, 0));
    if (f->d_nodes && f->dev == dev && f->d_slot_stats) {
        // tables written by the fit on this device: only the leaf vectors depend on the predict mode
        har::pool_setup();
        if (!f->d_leaf) HAR_CUDA(cudaMallocAsync((void**)&f->d_leaf, sizeof(double) * std::max<int64_t>(f->num_slots * f->C, 1), s));
        har::k_fin_leaf_table<<<(unsigned)har::ceil_div(std::max<int64_t>(f->num_slots, 1), 256), 256, 0, s>>>(
            f->d_slot_stats, f->num_slots, f->C, is_forest, f->d_leaf);
        HAR_LAUNCH_CHECK();
        f->leaf_mode = is_forest;
        return HAR_OK;
    }
    if (int rc = har_forest_ensure_host(f)) return rc;
    free_device_tables(f);
    const int T = f->T, C = f->C;
    const int64_t N = (int64_t)f->feature.size();
    std::vector<PNode> pn((size_t)N);
    std::vector<int64_t> roots(T), leaf_base(T + 1, 0);
    // trees are independent: count leaves per tree, then lay every tree out (BFS, children adjacent) in parallel
    har::parallel_for(T, N > 4096, [&](int t) {
        int64_t c = 0;
        for (int64_t k = f->tree_off[t]; k < f->tree_off[t + 1]; k++) c += f->feature[k] < 0;
        leaf_base[t + 1] = c;
    });
    for (int t = 0; t < T; t++) leaf_base[t + 1] += leaf_base[t];
    std::vector<double> leaf((size_t)leaf_base[T] * C);
    har::parallel_for(T, N > 4096, [&](int t) {
        const int64_t base = f->tree_off[t];
        int64_t next = base, nleaf = leaf_base[t];
        roots[t] = next++;
        std::vector<std::pair<int64_t, int64_t>> q;      // BFS queue of (flat index, device index)
        q.reserve((size_t)(f->tree_off[t + 1] - base));
        q.emplace_back(base, roots[t]);
        for (size_t qi = 0; qi < q.size(); qi++) {
            const int64_t src = q[qi].first, dst = q[qi].second;
            PNode& d = pn[(size_t)dst];
            if (f->feature[src] < 0) {
                d.feature = -1; d.left = (int32_t)nleaf; d.thr = 0.0;
                const double* st = &f->stats[(size_t)src * C];
                double* lv = &leaf[(size_t)nleaf * C];
                nleaf++;
                double total = 0.0;
                for (int c = 0; c < C; c++) total += st[c];
                for (int c = 0; c < C; c++) lv[c] = is_forest ? (total != 0.0 ? st[c] / total : 0.0) : st[c];
            } else {
                const int ar = f->arity[src];
                d.feature = f->feature[src];
                if (ar > 0) {
                    har::pnode_set_categorical(d, f->feature[src], ar, f->left_mask[src]);
                } else {
                    d.thr = f->threshold[src];
                }
                d.left = (int32_t)next;
                q.emplace_back(base + f->left[src], next);
                q.emplace_back(base + f->right[src], next + 1);
                next += 2;
            }
        }
    });
    har::pool_setup();
    HAR_CUDA(cudaMallocAsync(&f->d_nodes, sizeof(PNode) * std::max<int64_t>(N, 1), s));
    HAR_CUDA(cudaMallocAsync((void**)&f->d_leaf, sizeof(double) * std::max<size_t>(leaf.size(), 1), s));
    HAR_CUDA(cudaMallocAsync((void**)&f->d_roots, sizeof(int64_t) * std::max(T, 1), s));
    HAR_CUDA(cudaMemcpyAsync(f->d_nodes, pn.data(), sizeof(PNode) * N, cudaMemcpyHostToDevice, s));
    HAR_CUDA(cudaMemcpyAsync(f->d_leaf, leaf.data(), sizeof(double) * leaf.size(), cudaMemcpyHostToDevice, s));
    HAR_CUDA(cudaMemcpyAsync(f->d_roots, roots.data(), sizeof(int64_t) * T, cudaMemcpyHostToDevice, s));
    HAR_CUDA(cudaStreamSynchronize(s));
    f->dev = dev; f->leaf_mode = is_forest;
    return HAR_OK;
}

}  // namespace

HAR_API int har_forest_predict(const har_forest_t* fc, const double* d_X, int64_t n, int64_t ld, int32_t is_forest,
                               double* d_raw, double* d_prob, double* d_pred, har_stream_t stream) {
    if (!fc) return har::fail(HAR_E_INVALID, "har_forest_predict: NULL forest");
    har_forest* f = const_cast<har_forest*>(fc);
    if (n <= 0) return HAR_OK;
    if (ld < f->F) return har::fail(HAR_E_INVALID, "har_forest_predict: ld (%lld) < numFeatures (%d)", (long long)ld, f->F);
    cudaStream_t s = (cudaStream_t)stream;
    if (int rc = build_device_tables(f, is_forest ? 1 : 0, s)) return rc;
    const int C = f->C, F = f->F;
    if (int rc = build_compact_tables(f, s)) return rc;
    if (f->binnable == 1 && C <= PB_CMAX) {
        const int bstride = 4 * (((F + 3) >> 2) | 1);
        // large batches: one 1024-thread CTA per SM, two rows per thread, one tree at a time behind a barrier;
        // small batches: 128-thread CTAs, one row per thread, four trees in flight (fills the SMs, lowest latency)
        static const int shape = [] { const char* e = std::getenv("HAR_PREDICT_SHAPE"); return e ? std::atoi(e) : 0; }();
        har::prof_begin(s);
        if (!f->has_cat_nodes && shape < 10) {
#define HAR_PW_LAUNCH(NT_, R_, CM_)                                                                                             \
            {                                                                                                                   \
                const size_t smem = ((size_t)NT_ * R_ * bstride + 15) & ~(size_t)15;                                            \
                HAR_CUDA(cudaFuncSetAttribute(k_predict_walk<NT_, R_, CM_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
                k_predict_walk<NT_, R_, CM_><<<(unsigned)har::ceil_div(n, NT_ * R_), NT_, smem, s>>>(                           \
                    (const uint2*)f->d_walk, f->d_leaf_of, f->d_roots, f->T, f->d_leaf, C, d_X, n, ld, F, f->d_bin_thr,         \
                    f->d_bin_ns, f->max_bins, bstride, d_raw, d_prob, d_pred);                                                  \
            }
            const bool small = n < (int64_t)har::kNumSMs * 2048;
            // measured on 500 trees x depth 12 (M rows/s at 2e7 rows): 128 x 2 rows 102, 256 x 2 94, 256 x 4 84, 128 x 1 73
            if (C <= 6) {
                if (shape == 2 || (shape == 0 && small)) HAR_PW_LAUNCH(128, 1, 6)
                else if (shape == 4) HAR_PW_LAUNCH(256, 4, 6)
                else if (shape == 5) HAR_PW_LAUNCH(256, 2, 6)
                else HAR_PW_LAUNCH(128, 2, 6)
            } else {
                if (shape == 2 || (shape == 0 && small)) HAR_PW_LAUNCH(128, 1, 8)
                else HAR_PW_LAUNCH(256, 2, 8)
            }
#undef HAR_PW_LAUNCH
            HAR_LAUNCH_CHECK();
            har::prof_end(har::PROF_PREDICT, s, (double)n);
            return HAR_OK;
        }
#define HAR_PB_LAUNCH(NT_, R_, K_, SYNC_)                                                                                        \
        {                                                                                                                       \
            const size_t smem = ((size_t)NT_ * R_ * bstride + 15) & ~(size_t)15;                                                \
            HAR_CUDA(cudaFuncSetAttribute(k_predict_bins<NT_, R_, K_, SYNC_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
            k_predict_bins<NT_, R_, K_, SYNC_><<<(unsigned)har::ceil_div(n, NT_ * R_), NT_, smem, s>>>(                         \
                (const CNode*)f->d_cnodes, f->d_cmask, f->d_roots, f->T, f->d_leaf, C, d_X, n, ld, F, f->d_bin_thr, f->d_bin_ns, \
                f->max_bins, bstride, d_raw, d_prob, d_pred);                                                                   \
        }
        // (one 1024-thread CTA per SM with a barrier per tree was measured: 69 M rows/s against 88 for these)
        if (n >= (int64_t)har::kNumSMs * 2048) HAR_PB_LAUNCH(256, 2, 1, false)
        else HAR_PB_LAUNCH(128, 1, 4, false)
#undef HAR_PB_LAUNCH
        HAR_LAUNCH_CHECK();
        har::prof_end(har::PROF_PREDICT, s, (double)n);
        return HAR_OK;
    }
    const int fstride = F | 1;
    const size_t acc_bytes = (size_t)C * PRED_THREADS * sizeof(double);
    const size_t stage_bytes = (size_t)PRED_THREADS * fstride * sizeof(double);
    const unsigned blocks = (unsigned)har::ceil_div(n, PRED_THREADS);
    har::prof_begin(s);
    if (acc_bytes + stage_bytes <= 96 * 1024) {
        HAR_CUDA(cudaFuncSetAttribute(k_predict<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(acc_bytes + stage_bytes)));
        k_predict<true><<<blocks, PRED_THREADS, acc_bytes + stage_bytes, s>>>((const PNode*)f->d_nodes, f->d_roots, f->T, f->d_leaf, C,
                                                                              d_X, n, ld, F, fstride, d_raw, d_prob, d_pred);
    } else {
        k_predict<false><<<blocks, PRED_THREADS, acc_bytes, s>>>((const PNode*)f->d_nodes, f->d_roots, f->T, f->d_leaf, C, d_X, n, ld, F,
                                                                 fstride, d_raw, d_prob, d_pred);
    }
    HAR_LAUNCH_CHECK();
    har::prof_end(har::PROF_PREDICT, s, (double)n);
    if (!f->ev_pred) {
        cudaEvent_t e = nullptr;
        HAR_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        f->ev_pred = e;
    }
    HAR_CUDA(cudaEventRecord((cudaEvent_t)f->ev_pred, s));
    return HAR_OK;
}

HAR_API int har_forest_import(int32_t T, int32_t C, int32_t F, const int64_t* tree_off, const int32_t* feature,
                              const int32_t* left, const int32_t* right, const int32_t* arity, const double* threshold,
                              const uint64_t* left_mask, const double* stats, har_forest_t** out) {
    if (!out || T < 1 || C < 1 || C > HAR_MAX_CLASSES || F < 1) return har::fail(HAR_E_INVALID, "har_forest_import: bad shape");
    if (!tree_off || !feature || !left || !right || !arity || !threshold || !left_mask || !stats)
        return har::fail(HAR_E_INVALID, "har_forest_import: NULL array");
    if (tree_off[0] != 0) return har::fail(HAR_E_INVALID, "har_forest_import: tree_off[0] must be 0");
    for (int t = 0; t < T; t++)
        if (tree_off[t + 1] <= tree_off[t]) return har::fail(HAR_E_INVALID, "har_forest_import: tree %d is empty or tree_off decreases", t);
    const int64_t N = tree_off[T];
    // every tree must be a proper binary tree in pre-order: children inside the tree and after their parent, every
    // node reached exactly once (a malformed model would otherwise send the predict kernel out of bounds or in circles)
    {
        std::vector<uint8_t> seen((size_t)N, 0);
        for (int t = 0; t < T; t++) {
            const int64_t base = tree_off[t], size = tree_off[t + 1] - base;
            seen[(size_t)base] = 1;
            for (int64_t k = 0; k < size; k++) {
                const int64_t g = base + k;
                if (!seen[(size_t)g]) return har::fail(HAR_E_INVALID, "har_forest_import: node %lld of tree %d is not reachable from the root", (long long)k, t);
                if (feature[g] >= F) return har::fail(HAR_E_INVALID, "har_forest_import: node %lld tests feature %d >= %d", (long long)g, feature[g], F);
                if (feature[g] < 0) continue;
                if (!(arity[g] == 0 || (arity[g] >= 2 && arity[g] <= 64)))
                    return har::fail(HAR_E_INVALID, "har_forest_import: node %lld has arity %d (0 or 2..64 expected)", (long long)g, arity[g]);
                for (const int32_t ch : {left[g], right[g]}) {
                    if (ch <= k || ch >= size) return har::fail(HAR_E_INVALID, "har_forest_import: node %lld of tree %d has a child outside (%lld, %lld)", (long long)k, t, (long long)k, (long long)size);
                    if (seen[(size_t)(base + ch)]++) return har::fail(HAR_E_INVALID, "har_forest_import: node %lld of tree %d has two parents", (long long)ch, t);
                }
            }
        }
    }
    har_forest* f = new har_forest();
    f->T = T; f->C = C; f->F = F;
    f->tree_off.assign(tree_off, tree_off + T + 1);
    f->feature.assign(feature, feature + N); f->left.assign(left, left + N); f->right.assign(right, right + N);
    f->arity.assign(arity, arity + N); f->threshold.assign(threshold, threshold + N);
    f->left_mask.assign(left_mask, left_mask + N); f->stats.assign(stats, stats + N * C);
    *out = f;
    return HAR_OK;
}

HAR_API void har_forest_free(har_forest_t* f) {
    if (!f) return;
    if (f->ev_pred) { cudaEventSynchronize((cudaEvent_t)f->ev_pred); cudaEventDestroy((cudaEvent_t)f->ev_pred); f->ev_pred = nullptr; }
    free_device_tables(f);
    delete f;
}

HAR_API int har_csr_to_dense(const int64_t* d_indptr, const int32_t* d_indices, const double* d_values, int64_t n,
                             int32_t F, double* d_out, int64_t ld, har_stream_t stream) {
    if (n <= 0) return HAR_OK;
    if (ld < F) return har::fail(HAR_E_INVALID, "har_csr_to_dense: ld < num_features");
    cudaStream_t s = (cudaStream_t)stream;
    HAR_CUDA(cudaMemsetAsync(d_out, 0, sizeof(double) * (size_t)n * ld, s));
    k_csr_to_dense<<<(unsigned)har::ceil_div(n * 32, 256), 256, 0, s>>>(d_indptr, d_indices, d_values, n, d_out, ld);
    HAR_LAUNCH_CHECK();
    return HAR_OK;
}
