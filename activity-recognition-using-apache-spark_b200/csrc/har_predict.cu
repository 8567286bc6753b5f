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
}

int build_device_tables(har_forest* f, int is_forest, cudaStream_t s) {
    int dev = 0;
    HAR_CUDA(cudaGetDevice(&dev));
    if (f->d_nodes && f->dev == dev && f->leaf_mode == is_forest) return HAR_OK;
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
    return HAR_OK;
}

HAR_API int har_forest_import(int32_t T, int32_t C, int32_t F, const int64_t* tree_off, const int32_t* feature,
                              const int32_t* left, const int32_t* right, const int32_t* arity, const double* threshold,
                              const uint64_t* left_mask, const double* stats, har_forest_t** out) {
    if (!out || T < 1 || C < 1 || C > HAR_MAX_CLASSES || F < 1) return har::fail(HAR_E_INVALID, "har_forest_import: bad shape");
    const int64_t N = tree_off[T];
    for (int64_t k = 0; k < N; k++)
        if (feature[k] >= F) return har::fail(HAR_E_INVALID, "har_forest_import: node %lld tests feature %d >= %d", (long long)k, feature[k], F);
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
