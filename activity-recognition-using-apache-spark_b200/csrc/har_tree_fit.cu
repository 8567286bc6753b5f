// Host driver of the MLlib-compatible tree trainer + the stage-level entry points of
// include/har_b200.h (har_find_splits, har_bin_features, har_bag_weights, har_node_histogram,
// har_forest_fit, har_forest_export ...).
//
// Stands in for ml.tree.impl.RandomForest.run, reached from Main/main.py:300 (`dt.fit(train)`) and
// :481 (`rf.fit(train)`).  The driver-side control flow (node stack, selectNodesToSplit memory-bounded
// groups, feature-subset draws, child creation, pruning) follows SURVEY.md A.5-A.9 on the host, as it
// does on Spark's driver; everything that touches rows runs in the kernels of har_tree_kernels.cuh.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <deque>
#include <memory>
#include <thread>

#include "har_forest.h"
#include "har_tree_kernels.cuh"

using namespace har;

namespace {

// ---- DecisionTreeMetadata.buildMetadata -------------------------------------------
struct Meta {
    int F = 0, C = 0, mpb = 0;
    std::vector<int32_t> arity, num_bins, num_splits, hist_bins;
    std::vector<uint8_t> kind;
    std::vector<int32_t> loff;     // [F+1] word offsets of the full-width node histogram
};

int build_meta(Meta& m, int F, int C, int64_t n_global, int max_bins, const int32_t* arity) {
    m.F = F; m.C = C;
    m.mpb = (int)std::min<int64_t>(max_bins, n_global);
    m.arity.assign(F, 0); m.num_bins.assign(F, 0); m.num_splits.assign(F, 0); m.hist_bins.assign(F, 0);
    m.kind.assign(F, KIND_CONT);
    const int max_unordered = (int)std::floor(std::log((double)(m.mpb / 2 + 1)) / std::log(2.0) + 1.0);
    for (int f = 0; f < F; f++) {
        int k = arity ? arity[f] : 0;
        if (k <= 1) k = 0;
        m.arity[f] = k;
        if (k > m.mpb)
            return fail(HAR_E_INVALID,
                        "DecisionTree requires maxBins (= %d) to be at least as large as the number of values in each "
                        "categorical feature, but categorical feature %d has %d values. Considering remove this and "
                        "other categorical features with a large number of values, or add more training examples.",
                        m.mpb, f, k);
        if (k > MAX_CAT)
            return fail(HAR_E_UNSUPPORTED, "categorical feature %d has %d values; this library supports at most %d", f, k, MAX_CAT);
        if (k == 0) { m.num_bins[f] = m.mpb; m.num_splits[f] = m.mpb - 1; }
        else if (C > 2 && k <= max_unordered) { m.kind[f] = KIND_UNORDERED; m.num_bins[f] = (1 << (k - 1)) - 1; m.num_splits[f] = m.num_bins[f]; }
        else { m.kind[f] = KIND_ORDERED; m.num_bins[f] = k; m.num_splits[f] = k - 1; }
    }
    return HAR_OK;
}

int finish_meta(Meta& m) {
    m.loff.assign(m.F + 1, 0);
    int64_t off = 0;
    for (int f = 0; f < m.F; f++) {
        m.hist_bins[f] = m.arity[f] ? m.arity[f] : m.num_bins[f];
        m.loff[f] = (int32_t)off;
        off += (int64_t)m.hist_bins[f] * m.C;
        if (off > 0x7fffffff) return fail(HAR_E_UNSUPPORTED, "node histogram exceeds 2^31 counters");
    }
    m.loff[m.F] = (int32_t)off;
    return HAR_OK;
}

// ---- RandomForest.findSplits row sample (RDD.sample -> BernoulliSampler / GapSampling) -----------
void sample_rows(int64_t n_global, int mpb, int64_t seed, const int64_t* part_off, int num_parts,
                 int partition_offset, std::vector<int64_t>& out) {
    int64_t required = std::max<int64_t>((int64_t)mpb * mpb, 10000);
    out.clear();
    if (required >= n_global) {
        for (int64_t i = part_off[0]; i < part_off[num_parts]; i++) out.push_back(i);
        return;
    }
    const double fraction = (double)required / (double)n_global;
    XorShift sx(seed);
    JavaRand jr((int64_t)sx.next_int());
    for (int p = 0; p < partition_offset; p++) jr.next_long();
    for (int p = 0; p < num_parts; p++) {
        XorShift rng(jr.next_long());
        if (fraction <= 0.4) {
            const double lnq = std::log1p(-fraction);
            auto advance = [&]() { double u = std::max(rng.next_double(), 5e-11); return (int)(std::log(u) / lnq); };
            int drop = advance();
            for (int64_t i = part_off[p]; i < part_off[p + 1]; i++) {
                if (drop > 0) { drop--; continue; }
                drop = advance();
                out.push_back(i);
            }
        } else {
            for (int64_t i = part_off[p]; i < part_off[p + 1]; i++)
                if (rng.next_double() <= fraction) out.push_back(i);
        }
    }
}

int find_splits_impl(const double* d_X, int64_t n, int F, int64_t ld, const Meta& m, int max_bins, int64_t seed,
                     const int64_t* part_off, int num_parts, const har_dist* dist, double* thresholds,
                     int32_t* nsplits, cudaStream_t s) {
    std::fill(thresholds, thresholds + (size_t)F * max_bins, 0.0);
    std::fill(nsplits, nsplits + F, 0);
    std::vector<int32_t> cont;
    for (int f = 0; f < F; f++) if (m.arity[f] == 0) cont.push_back(f);
    const int nfc = (int)cont.size();
    if (nfc == 0) return HAR_OK;
    const int64_t n_global = dist ? dist->global_rows : n;
    const int world = dist ? dist->world_size : 1, rank = dist ? dist->rank : 0;
    std::vector<int64_t> rows;
    sample_rows(n_global, m.mpb, seed, part_off, num_parts, dist ? dist->partition_offset : 0, rows);
    const int64_t ns_local = (int64_t)rows.size();
    int64_t col0 = 0, ns_total = ns_local;
    if (world > 1) {
        DevBuf<int32_t> cnt;
        HAR_CUDA(cnt.alloc(world, s));
        HAR_CUDA(cudaMemsetAsync(cnt.p, 0, sizeof(int32_t) * world, s));
        int32_t mine = (int32_t)ns_local;
        HAR_CUDA(cudaMemcpyAsync(cnt.p + rank, &mine, sizeof(int32_t), cudaMemcpyHostToDevice, s));
        if (int rc = dist_allreduce(dist, cnt.p, world, s)) return rc;
        std::vector<int32_t> all(world);
        HAR_CUDA(cudaMemcpyAsync(all.data(), cnt.p, sizeof(int32_t) * world, cudaMemcpyDeviceToHost, s));
        HAR_CUDA(cudaStreamSynchronize(s));
        ns_total = 0;
        for (int r = 0; r < world; r++) { if (r == rank) col0 = ns_total; ns_total += all[r]; }
    }
    if (ns_total == 0) return HAR_OK;
    DevBuf<double> S, thr;
    DevBuf<int64_t> d_rows;
    DevBuf<int32_t> d_cont, d_ns;
    DevBuf<uint64_t> scratch;
    HAR_CUDA(S.alloc((size_t)nfc * ns_total, s));
    HAR_CUDA(d_rows.alloc(std::max<int64_t>(ns_local, 1), s));
    HAR_CUDA(d_cont.alloc(nfc, s)); HAR_CUDA(d_ns.alloc(nfc, s)); HAR_CUDA(thr.alloc((size_t)nfc * max_bins, s));
    if (world > 1) HAR_CUDA(cudaMemsetAsync(S.p, 0, sizeof(double) * nfc * ns_total, s));
    HAR_CUDA(cudaMemcpyAsync(d_cont.p, cont.data(), sizeof(int32_t) * nfc, cudaMemcpyHostToDevice, s));
    if (ns_local) {
        HAR_CUDA(cudaMemcpyAsync(d_rows.p, rows.data(), sizeof(int64_t) * ns_local, cudaMemcpyHostToDevice, s));
        k_gather_sample<<<(unsigned)ceil_div(ns_local * nfc, 256), 256, 0, s>>>(d_X, ld, d_rows.p, ns_local, d_cont.p, nfc,
                                                                                S.p, ns_total, col0);
        HAR_LAUNCH_CHECK();
    }
    // all-gather of the sample as an integer all-reduce: every slot is non-zero on exactly one rank
    if (int rc = dist_allreduce(dist, S.p, (int64_t)nfc * ns_total * 2, s)) return rc;
    int64_t npad = 2;
    while (npad < ns_total) npad <<= 1;
    size_t smem = 0;
    if (npad <= 16384) smem = (size_t)npad * sizeof(uint64_t);
    else HAR_CUDA(scratch.alloc((size_t)nfc * npad, s));
    HAR_CUDA(cudaMemsetAsync(thr.p, 0, sizeof(double) * nfc * max_bins, s));
    HAR_CUDA(cudaFuncSetAttribute(k_find_splits, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 * 8));
    prof_begin(s);
    k_find_splits<<<nfc, 512, smem, s>>>(S.p, ns_total, npad, m.mpb - 1, max_bins, scratch.p, thr.p, d_ns.p);
    HAR_LAUNCH_CHECK();
    prof_end(PROF_FIND_SPLITS, s, (double)ns_total);
    std::vector<double> h_thr((size_t)nfc * max_bins);
    std::vector<int32_t> h_ns(nfc);
    HAR_CUDA(cudaMemcpyAsync(h_thr.data(), thr.p, sizeof(double) * nfc * max_bins, cudaMemcpyDeviceToHost, s));
    HAR_CUDA(cudaMemcpyAsync(h_ns.data(), d_ns.p, sizeof(int32_t) * nfc, cudaMemcpyDeviceToHost, s));
    HAR_CUDA(cudaStreamSynchronize(s));
    for (int j = 0; j < nfc; j++) {
        nsplits[cont[j]] = h_ns[j];
        std::copy(h_thr.begin() + (size_t)j * max_bins, h_thr.begin() + (size_t)(j + 1) * max_bins,
                  thresholds + (size_t)cont[j] * max_bins);
    }
    return HAR_OK;
}

int bin_features_impl(const double* d_X, int64_t n, int F, int64_t ld, const int32_t* arity, const double* thresholds,
                      const int32_t* nsplits, int max_bins, uint8_t* d_bins, int64_t ldb, cudaStream_t s) {
    if (n == 0) return HAR_OK;
    DevBuf<int32_t> d_ar, d_ns;
    DevBuf<double> d_thr;
    DevBuf<unsigned long long> bad;
    HAR_CUDA(d_ar.alloc(F, s)); HAR_CUDA(d_ns.alloc(F, s)); HAR_CUDA(d_thr.alloc((size_t)F * max_bins, s));
    HAR_CUDA(bad.alloc(1, s));
    std::vector<int32_t> ar(F, 0);
    if (arity) std::copy(arity, arity + F, ar.begin());
    HAR_CUDA(cudaMemcpyAsync(d_ar.p, ar.data(), sizeof(int32_t) * F, cudaMemcpyHostToDevice, s));
    HAR_CUDA(cudaMemcpyAsync(d_ns.p, nsplits, sizeof(int32_t) * F, cudaMemcpyHostToDevice, s));
    HAR_CUDA(cudaMemcpyAsync(d_thr.p, thresholds, sizeof(double) * F * max_bins, cudaMemcpyHostToDevice, s));
    HAR_CUDA(cudaMemsetAsync(bad.p, 0xff, sizeof(unsigned long long), s));
    int64_t blocks = std::min<int64_t>(ceil_div(n * F, 256), (int64_t)kNumSMs * 32);
    prof_begin(s);
    k_bin_features<<<(unsigned)blocks, 256, 0, s>>>(d_X, n, F, ld, d_ar.p, d_thr.p, d_ns.p, max_bins, d_bins, ldb, bad.p);
    HAR_LAUNCH_CHECK();
    prof_end(PROF_BIN, s, (double)n);
    unsigned long long h_bad = 0;
    HAR_CUDA(cudaMemcpyAsync(&h_bad, bad.p, sizeof h_bad, cudaMemcpyDeviceToHost, s));
    HAR_CUDA(cudaStreamSynchronize(s));
    if (h_bad != ~0ULL) {
        int64_t i = (int64_t)(h_bad / F);
        int f = (int)(h_bad % F);
        double v = 0;
        cudaMemcpy(&v, d_X + i * ld + f, sizeof(double), cudaMemcpyDeviceToHost);
        return fail(HAR_E_INVALID,
                    "DecisionTree given invalid data: Feature %d is categorical with values in {0,...,%d}, but a data "
                    "point gives it value %g.\n  Bad data point: row %lld", f, ar[f] - 1, v, (long long)i);
    }
    return HAR_OK;
}

// BaggedPoint.convertToBaggedRDD: Poisson(rate) with replacement when numTrees > 1, else Bernoulli(rate)
int launch_bag(int64_t n, int T, double rate, int64_t seed, const int64_t* d_part_off, int num_parts, int partition_offset,
               uint8_t* d_w, cudaStream_t s) {
    const unsigned grid = (unsigned)ceil_div(num_parts, 32);
    if (T > 1) {
        const double lim = std::ceil(1000.0 * rate);
        const int ilimit = lim >= 2147483647.0 ? 2147483647 : (int)lim;
        const size_t smem = BAG_SMEM_BYTES;
        HAR_CUDA(cudaFuncSetAttribute(k_bag_poisson, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_bag_poisson<<<grid, BAG_THREADS, smem, s>>>(n, T, std::exp(-rate), ilimit, seed, d_part_off, num_parts, partition_offset, d_w);
    } else {
        k_bag_weights<<<grid, 32, 0, s>>>(n, T, rate, seed, d_part_off, num_parts, partition_offset, d_w);
    }
    HAR_LAUNCH_CHECK();
    return HAR_OK;
}

int bag_weights_impl(int64_t n, int T, double rate, int64_t seed, const int64_t* part_off, int num_parts,
                     int partition_offset, uint8_t* d_w, cudaStream_t s) {
    if (n == 0) return HAR_OK;
    if (T == 1 && rate == 1.0) { HAR_CUDA(cudaMemsetAsync(d_w, 1, (size_t)n, s)); return HAR_OK; }
    DevBuf<int64_t> d_po;
    HAR_CUDA(d_po.alloc(num_parts + 1, s));
    HAR_CUDA(cudaMemcpyAsync(d_po.p, part_off, sizeof(int64_t) * (num_parts + 1), cudaMemcpyHostToDevice, s));
    prof_begin(s);
    if (int rc = launch_bag(n, T, rate, seed, d_po.p, num_parts, partition_offset, d_w, s)) return rc;
    prof_end(PROF_BAG, s, (double)n * T);
    HAR_CUDA(cudaStreamSynchronize(s));       // part_off is a caller buffer
    return HAR_OK;
}

// ---- Scala 2.11 mutable.HashMap[Int, _] iteration order (findBestSplits' nodesForGroup.foreach) ----
uint32_t bswap32(uint32_t x) { return (x >> 24) | ((x >> 8) & 0xff00u) | ((x << 8) & 0xff0000u) | (x << 24); }
int scala_index(uint32_t hcode, int table_len) {
    uint32_t i = bswap32(hcode * 0x9e3775cdu) * 0x9e3775cdu;
    uint32_t rotated = (i >> 4) | (i << 28);                 // seedvalue = bitCount(initial table length - 1) = 4
    int ones = table_len - 1;
    return (int)((rotated >> __builtin_clz((unsigned)ones)) & (uint32_t)ones);
}
std::vector<int> scala_hashmap_order(const std::vector<int>& keys) {
    int tlen = 16, thr = 12, size = 0;
    std::vector<int> head(tlen, -1), next(keys.size(), -1);
    for (int k = 0; k < (int)keys.size(); k++) {
        int h = scala_index((uint32_t)keys[k], tlen);
        next[k] = head[h]; head[h] = k;
        if (++size > thr) {
            int nlen = tlen * 2;
            std::vector<int> nh(nlen, -1);
            for (int i = tlen - 1; i >= 0; i--)
                for (int e = head[i]; e >= 0;) {
                    int e1 = next[e];
                    int hh = scala_index((uint32_t)keys[e], nlen);
                    next[e] = nh[hh]; nh[hh] = e;
                    e = e1;
                }
            head.swap(nh); tlen = nlen; thr = (int)((int64_t)nlen * 750 / 1000);
        }
    }
    std::vector<int> out;
    for (int i = tlen - 1; i >= 0; i--) for (int e = head[i]; e >= 0; e = next[e]) out.push_back(e);
    return out;
}

// SamplingUtils.reservoirSampleAndCount(0 until F, k, seed)  
void reservoir_features(int F, int k, int64_t seed, int32_t* out) {
    int i = 0;
    for (; i < k && i < F; i++) out[i] = i;
    if (i < k) return;
    int64_t l = i;
    XorShift r(seed);
    for (int item = i; item < F; item++) {
        l++;
        int64_t j = (int64_t)(r.next_double() * (double)l);
        if (j < k) out[j] = item;
    }
}

// ---- LearningNode ---------------------------------------------------------------------------------
struct LNode {
    int32_t id = 1;
    bool is_leaf = false, valid = true;
    int feature = -1, kind = 0, split = 0, arity = 0;
    double threshold = 0, gain = 0, impurity = 0;
    uint64_t mask = 0;
    int left = -1, right = -1;
    // rows of this rank at the node
    int64_t start = 0;
    int32_t count = 0;
    int buf = 0;
    // sibling subtraction: the node's parent, its position in the batch being assembled, and (for a split node whose
    // two children both go on) the kept all-reduced counters until both children have been processed
    int parent = -1;
    int level = 0;                  // depth of the node (root 0)
    bool pushed = true;             // went onto the node stack when it was created (not a leaf by depth or purity)
    bool derive_pref = false;       // of a sibling pair, the child with more weighted instances: derived when both are in a batch
    int batch_pos = -1;
    int64_t batch_stamp = -1;
    uint32_t* kept = nullptr;
    int kids_pending = 0;
    std::vector<int32_t> subset;
};

// Device slots for kept parent counters ([hist_len] + [C] words each), carved from slabs that live until the fit ends.
struct HistStore {
    size_t slot_words = 0;
    std::deque<DevBuf<uint32_t>> slabs;
    std::vector<uint32_t*> free_slots;
    cudaError_t get(uint32_t** out, cudaStream_t s) {
        if (free_slots.empty()) {
            const size_t per = std::max<size_t>(64, (size_t(64) << 20) / (slot_words * 4));   // 64 MB slabs
            slabs.emplace_back();
            cudaError_t e = slabs.back().alloc(per * slot_words, s);
            if (e != cudaSuccess) { slabs.pop_back(); return e; }
            for (size_t i = per; i-- > 0;) free_slots.push_back(slabs.back().p + i * slot_words);
        }
        *out = free_slots.back();
        free_slots.pop_back();
        return cudaSuccess;
    }
    void put(uint32_t* p) { free_slots.push_back(p); }
};
int level_of(int id) { int l = 0; while (id > 1) { id >>= 1; l++; } return l; }
int first_argmax(const uint32_t* s, int C) {
    int b = -1; uint32_t mv = 0;
    for (int c = 0; c < C; c++) if (b < 0 || s[c] > mv) { b = c; mv = s[c]; }
    return b;
}

struct GroupItem { int tree; int node; };

// Number of node groups RandomForest.run would have formed for finished trees whose nodes all cost the same
// (all features per node): replays the LIFO node stack, the memory-bounded selectNodesToSplit and the
// per-group push order (Scala mutable.HashMap order of the tree indices, pop order inside a tree).
int64_t simulate_groups(const std::vector<std::vector<LNode>>& trees, int64_t cost, int64_t max_memory) {
    const int T = (int)trees.size();
    std::vector<GroupItem> stack, group;
    for (int t = 0; t < T; t++) stack.push_back({t, 0});
    std::vector<std::vector<int>> pushes(T);
    std::vector<char> seen(T, 0);
    std::vector<int> tkeys;
    int64_t groups = 0;
    while (!stack.empty()) {
        group.clear();
        int64_t mem = 0;
        while (!stack.empty()) {
            if (mem + cost <= max_memory || mem == 0) { group.push_back(stack.back()); stack.pop_back(); mem += cost; }
            else break;
        }
        groups++;
        tkeys.clear();
        for (auto& gi : group) if (!seen[gi.tree]) { seen[gi.tree] = 1; tkeys.push_back(gi.tree); }
        for (auto& gi : group) {
            const LNode& nd = trees[gi.tree][gi.node];
            if (nd.left < 0) continue;
            if (trees[gi.tree][nd.left].pushed) pushes[gi.tree].push_back(nd.left);
            if (trees[gi.tree][nd.right].pushed) pushes[gi.tree].push_back(nd.right);
        }
        const std::vector<int> torder = scala_hashmap_order(tkeys);
        for (int oi : torder) {
            const int t = tkeys[oi];
            for (int ni : pushes[t]) stack.push_back({t, ni});
            pushes[t].clear();
            seen[t] = 0;
        }
    }
    return groups;
}

// LearningNode.toNode(prune = true) into the flat pre-order layout
int64_t emit_node(har_forest& fo, const std::vector<LNode>& tree, const std::vector<uint32_t>& tstats, int idx, int64_t base,
                  bool* is_leaf_out) {
    const LNode& n = tree[idx];
    const int C = fo.C;
    const int64_t me = (int64_t)fo.feature.size();
    fo.feature.push_back(-1); fo.left.push_back(-1); fo.right.push_back(-1); fo.node_id.push_back(n.id);
    fo.arity.push_back(0); fo.threshold.push_back(0.0); fo.left_mask.push_back(0);
    const uint32_t* nstats = tstats.data() + (size_t)idx * C;
    fo.prediction.push_back((double)first_argmax(nstats, C));
    fo.impurity.push_back(0.0); fo.gain.push_back(0.0);
    for (int c = 0; c < C; c++) fo.stats.push_back((double)nstats[c]);
    if (n.left >= 0) {
        bool ll = false, rl = false;
        int64_t li = emit_node(fo, tree, tstats, n.left, base, &ll);
        int64_t ri = emit_node(fo, tree, tstats, n.right, base, &rl);
        if (ll && rl && fo.prediction[base + li] == fo.prediction[base + ri]) {
            const size_t keep = (size_t)me + 1;
            fo.feature.resize(keep); fo.left.resize(keep); fo.right.resize(keep); fo.node_id.resize(keep);
            fo.arity.resize(keep); fo.threshold.resize(keep); fo.left_mask.resize(keep); fo.prediction.resize(keep);
            fo.impurity.resize(keep); fo.gain.resize(keep); fo.stats.resize(keep * C);
            fo.impurity[me] = n.impurity;
            *is_leaf_out = true;
            return me - base;
        }
        fo.feature[me] = n.feature; fo.left[me] = (int32_t)li; fo.right[me] = (int32_t)ri;
        fo.arity[me] = n.kind ? n.arity : 0;
        fo.threshold[me] = n.threshold; fo.left_mask[me] = n.mask; fo.gain[me] = n.gain; fo.impurity[me] = n.impurity;
        *is_leaf_out = false;
        return me - base;
    }
    fo.impurity[me] = n.valid ? n.impurity : -1.0;
    *is_leaf_out = true;
    return me - base;
}

// LearningNode.toNode(prune = true) for every tree (in parallel), concatenated in tree order into the flat arrays
void emit_forest(har_forest* fo, int T, int C, const std::vector<std::vector<LNode>>& trees,
                 const std::vector<std::vector<uint32_t>>& tree_stats, bool trace) {
    const auto t_emit = std::chrono::steady_clock::now();
    fo->tree_off.assign(T + 1, 0);
    {
        std::vector<har_forest> parts(T);
        parallel_for(T, true, [&](int t) {
            har_forest& pf = parts[t];
            pf.C = C;
            const size_t nn = trees[t].size();
            pf.feature.reserve(nn); pf.left.reserve(nn); pf.right.reserve(nn); pf.node_id.reserve(nn); pf.arity.reserve(nn);
            pf.threshold.reserve(nn); pf.left_mask.reserve(nn); pf.prediction.reserve(nn); pf.impurity.reserve(nn);
            pf.gain.reserve(nn); pf.stats.reserve(nn * C);
            bool leaf = false;
            emit_node(pf, trees[t], tree_stats[t], 0, 0, &leaf);
        });
        if (trace) fprintf(stderr, "[har_forest_fit]   toNode: parallel emit %.2f ms\n", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_emit).count());
        size_t total_nodes = 0;
        for (int t = 0; t < T; t++) { fo->tree_off[t] = (int64_t)total_nodes; total_nodes += parts[t].feature.size(); }
        // one host thread per output array: the first touch of the fresh allocations (page faults) and the copies
        // run in parallel instead of serialising ~12 MB of inserts
        har_forest* fp = fo;
        auto cat = [&](auto har_forest::*member, size_t per_node) {
            auto& dst = fp->*member;
            dst.reserve(total_nodes * per_node);
            for (int t = 0; t < T; t++) { const auto& src = parts[t].*member; dst.insert(dst.end(), src.begin(), src.end()); }
        };
        parallel_for(11, true, [&](int k) {
            switch (k) {
                case 0: cat(&har_forest::stats, (size_t)C); break;
                case 1: cat(&har_forest::feature, 1); break;
                case 2: cat(&har_forest::left, 1); break;
                case 3: cat(&har_forest::right, 1); break;
                case 4: cat(&har_forest::node_id, 1); break;
                case 5: cat(&har_forest::arity, 1); break;
                case 6: cat(&har_forest::threshold, 1); break;
                case 7: cat(&har_forest::left_mask, 1); break;
                case 8: cat(&har_forest::prediction, 1); break;
                case 9: cat(&har_forest::impurity, 1); break;
                default: cat(&har_forest::gain, 1); break;
            }
        });
    }
    fo->tree_off[T] = (int64_t)fo->feature.size();
    fo->total_nodes = (int64_t)fo->feature.size();
}

// ---- read-back of trees grown on the device ---------------------------------------------------------------
// The device-driven fit leaves the trained nodes in per-set pools on the GPU (level order; see k_grow_* / k_fin_*)
// and writes the predict tables from them.  Everything the *host* may still want — the flat model arrays of
// har_forest_export, Spark's node-group count — is rebuilt from the pools on first use.
void build_tree_from_pool(std::vector<LNode>& tr, std::vector<uint32_t>& ts, int root_pool_idx, const NodeResult* R,
                          const uint32_t* RS, const PoolMeta* M, int C, int max_depth, const std::vector<int32_t>& arity,
                          const std::vector<double>& thresholds, int MB) {
    tr.clear(); tr.emplace_back();                                     // root: id 1, level 0
    ts.assign(C, 0);
    std::vector<std::pair<int, int>> todo;                             // (pool index, LNode index), parents first
    todo.emplace_back(root_pool_idx, 0);
    for (size_t k = 0; k < todo.size(); k++) {
        const int pi = todo[k].first, ni = todo[k].second;
        const NodeResult& r = R[pi];
        const uint32_t* rs = RS + (size_t)pi * 3 * C;
        const int level = tr[ni].level;
        const bool is_leaf = !(r.gain > 0.0) || level == max_depth;
        {
            LNode& nd = tr[ni];
            nd.is_leaf = is_leaf; nd.valid = r.valid != 0; nd.gain = r.gain; nd.impurity = r.impurity;
            for (int c = 0; c < C; c++) ts[(size_t)ni * C + c] = rs[2 * C + c];
        }
        if (is_leaf) continue;
        const bool child_leaf = level + 1 == max_depth;
        const bool ll = child_leaf || r.left_imp == 0.0, rl = child_leaf || r.right_imp == 0.0;
        const int li = (int)tr.size();
        tr.emplace_back(); tr.emplace_back();
        LNode& nd = tr[ni];
        LNode& Lc = tr[li]; LNode& Rc = tr[li + 1];
        Lc.id = nd.id * 2; Rc.id = Lc.id + 1;
        Lc.parent = Rc.parent = ni;
        Lc.level = Rc.level = level + 1;
        Lc.is_leaf = ll; Rc.is_leaf = rl; Lc.impurity = r.left_imp; Rc.impurity = r.right_imp;
        Lc.pushed = !ll; Rc.pushed = !rl;
        ts.resize((size_t)(li + 2) * C);
        for (int c = 0; c < C; c++) { ts[(size_t)li * C + c] = rs[c]; ts[(size_t)(li + 1) * C + c] = rs[C + c]; }
        nd.left = li; nd.right = li + 1;
        nd.feature = r.feature; nd.kind = r.kind; nd.split = r.split; nd.mask = r.left_mask; nd.arity = arity[r.feature];
        nd.threshold = r.kind == KIND_CONT ? thresholds[(size_t)r.feature * MB + r.split] : 0.0;
        if (M[pi].left >= 0) todo.emplace_back(M[pi].left, li);
        if (M[pi].right >= 0) todo.emplace_back(M[pi].right, li + 1);
    }
}

struct LazyPools {
    int T = 0, C = 0, MB = 0, max_depth = 0, num_sets = 0, dev = 0;
    std::vector<int32_t> arity;
    std::vector<double> thresholds;
    struct SetPool {
        std::vector<int> my_trees;
        int64_t N = 0;
        NodeResult* res = nullptr; uint32_t* stats = nullptr; PoolMeta* meta = nullptr;     // device, owned
    } set[2];
    std::mutex mu;
    bool built = false;
    std::vector<std::vector<LNode>> trees;
    std::vector<std::vector<uint32_t>> tree_stats;
    ~LazyPools() {
        for (auto& sp : set) { if (sp.res) cudaFree(sp.res); if (sp.stats) cudaFree(sp.stats); if (sp.meta) cudaFree(sp.meta); }
    }
    // pools -> host -> per-tree LearningNode vectors (once); the device copies are released afterwards
    int build() {
        std::lock_guard<std::mutex> lk(mu);
        if (built) return HAR_OK;
        int cur = 0;
        HAR_CUDA(cudaGetDevice(&cur));
        if (cur != dev) HAR_CUDA(cudaSetDevice(dev));
        trees.assign(T, {}); tree_stats.assign(T, {});
        for (int k = 0; k < num_sets; k++) {
            SetPool& sp = set[k];
            if (sp.N == 0) continue;
            std::vector<NodeResult> R((size_t)sp.N);
            std::vector<uint32_t> RS((size_t)sp.N * 3 * C);
            std::vector<PoolMeta> M((size_t)sp.N);
            HAR_CUDA(cudaMemcpy(R.data(), sp.res, sizeof(NodeResult) * sp.N, cudaMemcpyDeviceToHost));
            HAR_CUDA(cudaMemcpy(RS.data(), sp.stats, sizeof(uint32_t) * (size_t)sp.N * 3 * C, cudaMemcpyDeviceToHost));
            HAR_CUDA(cudaMemcpy(M.data(), sp.meta, sizeof(PoolMeta) * sp.N, cudaMemcpyDeviceToHost));
            parallel_for((int)sp.my_trees.size(), sp.N >= 512, [&](int q) {
                const int t = sp.my_trees[q];
                build_tree_from_pool(trees[t], tree_stats[t], q, R.data(), RS.data(), M.data(), C, max_depth, arity, thresholds, MB);
            });
            cudaFree(sp.res); cudaFree(sp.stats); cudaFree(sp.meta);
            sp.res = nullptr; sp.stats = nullptr; sp.meta = nullptr;
        }
        if (cur != dev) cudaSetDevice(cur);
        built = true;
        return HAR_OK;
    }
};

template <typename T>
struct Pinned {
    T* p = nullptr;
    size_t cap = 0;
    ~Pinned() { if (p) cudaFreeHost(p); }
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = std::max(n, cap * 2);
        return cudaMallocHost((void**)&p, cap * sizeof(T));
    }
};


// A second stream forked from / joined into the caller's stream with events (created on first use).
struct SideStream {
    cudaStream_t s = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    bool active = false;
    ~SideStream() {
        if (active) cudaStreamSynchronize(s);
        if (ev_fork) cudaEventDestroy(ev_fork);
        if (ev_join) cudaEventDestroy(ev_join);
        if (s) cudaStreamDestroy(s);
    }
    cudaError_t start() {                               // independent of any other stream
        cudaError_t e;
        if (!s) {
            if ((e = cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking)) != cudaSuccess) return e;
            if ((e = cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming)) != cudaSuccess) return e;
            if ((e = cudaEventCreateWithFlags(&ev_join, cudaEventDisableTiming)) != cudaSuccess) return e;
        }
        active = true;
        return cudaSuccess;
    }
    cudaError_t fork(cudaStream_t main) {
        cudaError_t e;
        if (!s) {
            if ((e = cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking)) != cudaSuccess) return e;
            if ((e = cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming)) != cudaSuccess) return e;
            if ((e = cudaEventCreateWithFlags(&ev_join, cudaEventDisableTiming)) != cudaSuccess) return e;
        }
        if ((e = cudaEventRecord(ev_fork, main)) != cudaSuccess) return e;
        active = true;
        return cudaStreamWaitEvent(s, ev_fork, 0);
    }
    cudaError_t join(cudaStream_t main) {
        if (!active) return cudaSuccess;
        cudaError_t e = cudaEventRecord(ev_join, s);
        if (e != cudaSuccess) return e;
        active = false;
        return cudaStreamWaitEvent(main, ev_join, 0);
    }
};

}  // namespace

// =====================================================================================================
HAR_API int har_find_splits(const double* d_X, int64_t n, int32_t F, int64_t ld, const int32_t* arity, int32_t max_bins,
                            int64_t seed, const int64_t* part_off, int32_t num_parts, const har_dist* dist,
                            double* thresholds, int32_t* nsplits, har_stream_t stream) {
    if (F <= 0 || max_bins < 2 || max_bins > HAR_MAX_BINS) return fail(HAR_E_INVALID, "har_find_splits: bad F / max_bins");
    Meta m;
    if (int rc = build_meta(m, F, 3, dist ? dist->global_rows : n, max_bins, arity)) return rc;
    int64_t po[2] = {0, n};
    if (!part_off) { part_off = po; num_parts = 1; }
    return find_splits_impl(d_X, n, F, ld, m, max_bins, seed, part_off, num_parts, dist, thresholds, nsplits, (cudaStream_t)stream);
}

HAR_API int har_bin_features(const double* d_X, int64_t n, int32_t F, int64_t ld, const int32_t* arity,
                             const double* thresholds, const int32_t* nsplits, int32_t max_bins, uint8_t* d_bins,
                             int64_t ldb, har_stream_t stream) {
    if (F <= 0 || ldb < F) return fail(HAR_E_INVALID, "har_bin_features: need ldb >= num_features");
    return bin_features_impl(d_X, n, F, ld, arity, thresholds, nsplits, max_bins, d_bins, ldb, (cudaStream_t)stream);
}

HAR_API int har_bag_weights(int64_t n, int32_t T, double rate, int64_t seed, const int64_t* part_off, int32_t num_parts,
                            int32_t partition_offset, uint8_t* d_weights, har_stream_t stream) {
    if (T < 1 || !(rate > 0.0) || rate > 1.0) return fail(HAR_E_INVALID, "har_bag_weights: need num_trees >= 1 and rate in (0, 1]");
    int64_t po[2] = {0, n};
    if (!part_off) { part_off = po; num_parts = 1; }
    return bag_weights_impl(n, T, rate, seed, part_off, num_parts, partition_offset, d_weights, (cudaStream_t)stream);
}

HAR_API int har_node_histogram(const uint8_t* d_bins, int64_t ldb, const uint8_t* d_labels, const uint8_t* d_weights,
                               const int32_t* d_rows, int64_t nrows, int32_t F, int32_t C, const int64_t* hist_off,
                               uint32_t* d_hist, uint32_t* d_parent, har_stream_t stream) {
    if (C < 1 || C > HAR_MAX_CLASSES) return fail(HAR_E_UNSUPPORTED, "num_classes must be in [1, %d]", HAR_MAX_CLASSES);
    if (nrows <= 0) return HAR_OK;
    if (nrows > 0x7fffffff || hist_off[F] > 0x7fffffff) return fail(HAR_E_UNSUPPORTED, "har_node_histogram: sizes exceed 2^31");
    cudaStream_t s = (cudaStream_t)stream;
    std::vector<int32_t> loff(F + 1);
    for (int f = 0; f <= F; f++) loff[f] = (int32_t)hist_off[f];
    for (int f = 0; f < F; f++)
        if ((int64_t)(loff[f + 1] - loff[f]) > HIST_TILE_WORDS)
            return fail(HAR_E_UNSUPPORTED, "feature %d needs %d counters; tile limit is %d", f, loff[f + 1] - loff[f], HIST_TILE_WORDS);
    const int chunk_rows = 2048;
    NodeDesc nd{};
    nd.start = 0; nd.hist_base = 0; nd.count = (int32_t)nrows; nd.tree = 0; nd.buf = 0; nd.level = 0; nd.sub_off = 0; nd.sib = -1;
    nd.nf = F; nd.hist_len = loff[F]; nd.id = 1; nd.wtotal = 0xffffffffu;
    int32_t item_off[2] = {0, (int32_t)ceil_div(nrows, chunk_rows)};
    DevBuf<NodeDesc> d_nd; DevBuf<int32_t> d_io, d_loff; DevBuf<uint8_t> d_kind;
    HAR_CUDA(d_nd.alloc(1, s)); HAR_CUDA(d_io.alloc(2, s)); HAR_CUDA(d_loff.alloc(F + 1, s)); HAR_CUDA(d_kind.alloc(F, s));
    HAR_CUDA(cudaMemcpyAsync(d_nd.p, &nd, sizeof nd, cudaMemcpyHostToDevice, s));
    HAR_CUDA(cudaMemcpyAsync(d_io.p, item_off, sizeof item_off, cudaMemcpyHostToDevice, s));
    HAR_CUDA(cudaMemcpyAsync(d_loff.p, loff.data(), sizeof(int32_t) * (F + 1), cudaMemcpyHostToDevice, s));
    HAR_CUDA(cudaMemsetAsync(d_kind.p, KIND_CONT, F, s));
    const size_t smem = (HIST_TILE_WORDS + 32 + HIST_FTAB + HIST_FTAB / 16) * sizeof(uint32_t);
    const bool vec = (ldb % 16 == 0) && ((reinterpret_cast<uintptr_t>(d_bins) & 15) == 0);
    int max_width = 0;
    for (int f = 0; f < F; f++) max_width = std::max(max_width, loff[f + 1] - loff[f]);
    const int keys = (max_width + 15) & ~15;
    bool lanes = C <= 8 && F <= HISTL_MAX_FEATURES && keys <= HISTL_MAX_KEYS && nrows * ldb < (1LL << 32);
    for (int f = 0; f < F && lanes; f++) lanes = (loff[f + 1] - loff[f]) % C == 0;
    if (const char* e = std::getenv("HAR_HIST_LANES")) lanes = lanes && std::atoi(e) != 0;
    if (lanes) {
        const size_t sm = histl_smem_bytes(keys, F);
        HAR_CUDA(cudaFuncSetAttribute(k_node_hist_lanes<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
        k_node_hist_lanes<false><<<item_off[1], HIST_THREADS, sm, s>>>(d_nd.p, d_io.p, 1, chunk_rows, d_bins, (uint32_t)ldb, d_labels, d_weights,
                                                                nrows, d_rows, nullptr, d_loff.p, d_kind.p, C, F, keys, 0, d_hist, d_parent);
    } else if (vec) {
        HAR_CUDA(cudaFuncSetAttribute(k_node_hist<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_node_hist<false><<<item_off[1], HIST_THREADS, smem, s>>>(d_nd.p, d_io.p, 1, chunk_rows, d_bins, ldb, d_labels, d_weights,
                                                                  nrows, d_rows, nullptr, nullptr, d_loff.p, d_kind.p, C, d_hist, d_parent);
    } else {
        HAR_CUDA(cudaFuncSetAttribute(k_node_hist<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_node_hist<true><<<item_off[1], HIST_THREADS, smem, s>>>(d_nd.p, d_io.p, 1, chunk_rows, d_bins, ldb, d_labels, d_weights,
                                                                 nrows, d_rows, nullptr, nullptr, d_loff.p, d_kind.p, C, d_hist, d_parent);
    }
    HAR_LAUNCH_CHECK();
    HAR_CUDA(cudaStreamSynchronize(s));
    return HAR_OK;
}

// =====================================================================================================
HAR_API int har_forest_fit(const double* d_X, int64_t n, int32_t F, int64_t ld, const double* d_y, const int32_t* arity,
                           const int64_t* part_off_in, int32_t num_parts, const har_fit_params* p, const har_dist* dist,
                           const uint8_t* d_weights_in, har_stream_t stream, har_forest_t** out) {
    ArenaScope scratch_scope;                            // device scratch of this call comes from the thread's arena
    // host threads for the driver-side work: the ranks of a node share its cores
    host_threads_set_ranks(dist ? dist->world_size : 1);
    if (!out || !p) return fail(HAR_E_INVALID, "har_forest_fit: NULL params / out");
    *out = nullptr;
    cudaStream_t s = (cudaStream_t)stream;
    const int T = p->num_trees, C = p->num_classes, MB = p->max_bins;
    if (p->max_depth < 0 || p->max_depth > 30)
        return fail(HAR_E_INVALID, "DecisionTree currently only supports maxDepth <= 30, but was given maxDepth = %d.", p->max_depth);
    if (MB < 2) return fail(HAR_E_INVALID, "maxBins must be >= 2, got %d", MB);
    if (MB > HAR_MAX_BINS) return fail(HAR_E_UNSUPPORTED, "maxBins = %d exceeds this library's limit of %d", MB, HAR_MAX_BINS);
    if (C < 2 || C > HAR_MAX_CLASSES) return fail(HAR_E_UNSUPPORTED, "numClasses = %d outside [2, %d]", C, HAR_MAX_CLASSES);
    if (T < 1 || F < 1 || n < 0 || n > 0x7fffffff) return fail(HAR_E_INVALID, "har_forest_fit: bad num_trees / num_features / n");
    if (p->min_instances_per_node < 1) return fail(HAR_E_INVALID, "minInstancesPerNode must be >= 1");
    if (!(p->subsampling_rate > 0.0) || p->subsampling_rate > 1.0) return fail(HAR_E_INVALID, "subsamplingRate must be in (0, 1]");
    if (p->features_per_node < 1 || p->features_per_node > F) return fail(HAR_E_INVALID, "features_per_node must be in [1, num_features]");
    const int64_t n_global = dist ? dist->global_rows : n;
    const int world = dist ? dist->world_size : 1;
    if (n_global < 1) return fail(HAR_E_INVALID, "DecisionTree requires a non-empty training set");
    int64_t po_default[2] = {0, n};
    const int64_t* part_off = part_off_in;
    if (!part_off) { part_off = po_default; num_parts = 1; }
    const int part_offset = dist ? dist->partition_offset : 0;
    const long long launches0 = launches_since_reset();
    // HAR_FIT_TRACE=1: wall-clock split of the fit (stderr), for tuning only
    const bool trace = std::getenv("HAR_FIT_TRACE") != nullptr;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto ms_since = [&](std::chrono::steady_clock::time_point a) { return std::chrono::duration<double, std::milli>(now() - a).count(); };
    double t_prep = 0, t_wait = 0, t_post = 0;
    auto t_begin = now();
    std::string setup_marks;                             // HAR_FIT_TRACE: host time at the setup milestones
    auto mark = [&](const char* what) { if (trace) { char b[64]; snprintf(b, sizeof b, " %s %.2f", what, ms_since(t_begin)); setup_marks += b; } };

    Meta m;
    if (int rc = build_meta(m, F, C, n_global, MB, arity)) return rc;

    // labels -> u8 (Classifier.extractLabeledPoints validation)
    // Bagging first, on a side stream forked from the caller's: the serial-RNG kernel runs under split finding and
    // binning below.  (Letting it start before the caller's pending work — the window-feature kernel — was tried:
    // its 111 KB CTAs then wait for free SMs and it finishes later.)
    const bool need_w = d_weights_in || !(T == 1 && p->subsampling_rate == 1.0);
    SideStream side;                                     // declared before the buffers that are freed on it
    DevBuf<uint8_t> d_wbuf;
    DevBuf<int64_t> d_po;
    const uint8_t* d_w = d_weights_in;
    if (need_w && !d_weights_in && n > 0) {
        HAR_CUDA(d_wbuf.alloc((size_t)T * n, s));
        d_w = d_wbuf.p;
        if (T == 1 && p->subsampling_rate == 1.0) {
            HAR_CUDA(cudaMemsetAsync(d_wbuf.p, 1, (size_t)n, s));
        } else {
            HAR_CUDA(d_po.alloc(num_parts + 1, s));
            HAR_CUDA(cudaMemcpyAsync(d_po.p, part_off, sizeof(int64_t) * (num_parts + 1), cudaMemcpyHostToDevice, s));
            HAR_CUDA(side.fork(s));                      // side stream waits for the allocations / copy above
            prof_begin(side.s);
            if (int rc = launch_bag(n, T, p->subsampling_rate, p->seed, d_po.p, num_parts, part_offset, d_wbuf.p, side.s)) return rc;
            prof_end(PROF_BAG, side.s, (double)n * T);
        }
    }

    // labels -> u8 (validated like Classifier.extractLabeledPoints); the verdict is read at the first sync below
    DevBuf<uint8_t> d_lab; DevBuf<unsigned long long> d_bad;
    static thread_local Pinned<unsigned long long> h_bad_pin;
    HAR_CUDA(h_bad_pin.reserve(1));
    HAR_CUDA(d_lab.alloc(std::max<int64_t>(n, 1), s)); HAR_CUDA(d_bad.alloc(1, s));
    HAR_CUDA(cudaMemsetAsync(d_bad.p, 0xff, sizeof(unsigned long long), s));
    if (n) { k_labels_to_u8<<<(unsigned)ceil_div(n, 256), 256, 0, s>>>(d_y, n, C, d_lab.p, d_bad.p); HAR_LAUNCH_CHECK(); }
    h_bad_pin.p[0] = ~0ULL;
    HAR_CUDA(cudaMemcpyAsync(h_bad_pin.p, d_bad.p, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));

    mark("bag+labels launched");
    std::unique_ptr<har_forest> fo(new har_forest());
    fo->T = T; fo->C = C; fo->F = F; fo->max_bins = MB; fo->max_depth = p->max_depth;
    fo->thresholds.assign((size_t)F * MB, 0.0);
    fo->nsplits.assign(F, 0);
    if (int rc = find_splits_impl(d_X, n, F, ld, m, MB, p->seed, part_off, num_parts, dist, fo->thresholds.data(),
                                  fo->nsplits.data(), s)) return rc;
    mark("find_splits");
    HAR_CUDA(cudaStreamSynchronize(s));
    if (h_bad_pin.p[0] != ~0ULL) {
        const unsigned long long h_bad = h_bad_pin.p[0];
        double v = 0;
        cudaMemcpy(&v, d_y + h_bad, sizeof(double), cudaMemcpyDeviceToHost);
        return fail(HAR_E_INVALID, "Classifier given dataset with invalid label %g.  Labels must be integers in range [0, %d).", v, C);
    }
    for (int f = 0; f < F; f++) if (m.arity[f] == 0) { m.num_splits[f] = fo->nsplits[f]; m.num_bins[f] = fo->nsplits[f] + 1; }
    if (int rc = finish_meta(m)) return rc;
    for (int f = 0; f < F; f++)
        if (m.hist_bins[f] * C > HIST_TILE_WORDS) return fail(HAR_E_UNSUPPORTED, "feature %d needs more than %d counters", f, HIST_TILE_WORDS);

    const int64_t ldb = (F + 15) & ~15;
    DevBuf<uint8_t> d_bins;
    HAR_CUDA(d_bins.alloc((size_t)std::max<int64_t>(n, 1) * ldb, s));
    if (ldb != F) HAR_CUDA(cudaMemsetAsync(d_bins.p, 0, (size_t)std::max<int64_t>(n, 1) * ldb, s));   // zero padding bytes
    if (int rc = bin_features_impl(d_X, n, F, ld, arity, fo->thresholds.data(), fo->nsplits.data(), MB, d_bins.p, ldb, s)) return rc;

    // feature-major copy for the row partition (HAR_PART_COLUMNS=0: gather from the row-major table as before)
    DevBuf<uint8_t> d_binsT;
    const int64_t nT = (n + 31) & ~(int64_t)31;
    {
        static const bool off = [] { const char* e = std::getenv("HAR_PART_COLUMNS"); return e && std::atoi(e) == 0; }();
        if (!off && n > 0 && p->max_depth > 0) {
            HAR_CUDA(d_binsT.alloc((size_t)F * nT, s));
            k_bins_transpose<<<dim3((unsigned)ceil_div(n, 32), (unsigned)ceil_div(F, 32)), 256, 0, s>>>(d_bins.p, n, F, ldb, d_binsT.p, nT);
            HAR_LAUNCH_CHECK();
        }
    }
    mark("binned");
    HAR_CUDA(side.join(s));                              // weights are needed from here on

    // row permutations
    DevBuf<int32_t> perm0, perm1, d_cnt, d_chunk;
    const size_t perm_len = (size_t)T * std::max<int64_t>(n, 1);
    HAR_CUDA(perm0.alloc(perm_len, s)); HAR_CUDA(perm1.alloc(perm_len, s));
    std::vector<int32_t> root_cnt(T, (int32_t)n);
    // the per-tree row counts come back through page-locked memory behind an event, so that the host can go on
    // preparing (buffers of the level loop) while the bagging kernel still runs; wait_roots() is the sync point
    static thread_local Pinned<int32_t> h_root_cnt;
    struct EventHolder { cudaEvent_t ev = nullptr; ~EventHolder() { if (ev) cudaEventDestroy(ev); } } ev_root;
    bool roots_pending = false;
    if (n) {
        if (need_w) {
            const int nchunks = (int)ceil_div(n, INIT_CHUNK);
            HAR_CUDA(d_cnt.alloc(T, s)); HAR_CUDA(d_chunk.alloc((size_t)T * nchunks, s));
            k_init_count<<<dim3(nchunks, T), 256, 0, s>>>(d_w, n, d_chunk.p, nchunks); HAR_LAUNCH_CHECK();
            k_init_scan<<<(unsigned)ceil_div(T, 4), 128, 0, s>>>(d_chunk.p, nchunks, T, d_cnt.p); HAR_LAUNCH_CHECK();
            k_init_scatter<<<dim3(nchunks, T), 256, 0, s>>>(d_w, n, d_chunk.p, nchunks, perm0.p); HAR_LAUNCH_CHECK();
            HAR_CUDA(h_root_cnt.reserve(T));
            HAR_CUDA(cudaMemcpyAsync(h_root_cnt.p, d_cnt.p, sizeof(int32_t) * T, cudaMemcpyDeviceToHost, s));
            HAR_CUDA(cudaEventCreateWithFlags(&ev_root.ev, cudaEventDisableTiming));
            HAR_CUDA(cudaEventRecord(ev_root.ev, s));
            roots_pending = true;
        } else {
            k_iota<<<(unsigned)ceil_div(n, 256), 256, 0, s>>>(perm0.p, n); HAR_LAUNCH_CHECK();
        }
    }

    mark("perm init launched");
    // device-side metadata
    DevBuf<uint8_t> d_kind; DevBuf<int32_t> d_nsplits, d_hbins, d_loff_all;
    HAR_CUDA(d_kind.alloc(F, s)); HAR_CUDA(d_nsplits.alloc(F, s)); HAR_CUDA(d_hbins.alloc(F, s));
    HAR_CUDA(cudaMemcpyAsync(d_kind.p, m.kind.data(), F, cudaMemcpyHostToDevice, s));
    HAR_CUDA(cudaMemcpyAsync(d_nsplits.p, m.num_splits.data(), sizeof(int32_t) * F, cudaMemcpyHostToDevice, s));
    HAR_CUDA(cudaMemcpyAsync(d_hbins.p, m.hist_bins.data(), sizeof(int32_t) * F, cudaMemcpyHostToDevice, s));
    FitMeta fm;
    fm.kind = d_kind.p; fm.nsplits = d_nsplits.p; fm.hist_bins = d_hbins.p;
    fm.C = C; fm.F = F; fm.max_depth = p->max_depth; fm.min_inst = p->min_instances_per_node; fm.impurity = p->impurity;
    fm.min_gain = p->min_info_gain;

    const bool subsample = p->features_per_node < F;
    const int K = p->features_per_node;
    int64_t total_bins_all = 0;
    for (int f = 0; f < F; f++) total_bins_all += m.num_bins[f];
    const int64_t full_len = m.loff[F];

    // trees + node stack (RandomForest.run)
    std::vector<std::vector<LNode>> trees(T);
    std::vector<std::vector<uint32_t>> tree_stats(T);      // [node][C] class counts, kept out of the hot LNode
    auto wait_roots = [&]() -> int {
        if (!roots_pending) return HAR_OK;
        HAR_CUDA(cudaEventSynchronize(ev_root.ev));
        std::copy(h_root_cnt.p, h_root_cnt.p + T, root_cnt.begin());
        roots_pending = false;
        return HAR_OK;
    };
    JavaRand rng(p->seed);

    // batch buffers
    int64_t budget_words = 96LL << 20;                       // 384 MB of counters per device batch
    if (const char* e = std::getenv("HAR_HIST_BUDGET_WORDS")) budget_words = std::max<int64_t>(1, std::atoll(e));   // tests
    // sibling subtraction (all-features fits only: with per-node feature subsets parent and child count different features)
    bool subtract = !subsample;
    if (const char* e = std::getenv("HAR_HIST_SUBTRACT")) subtract = subtract && std::atoi(e) != 0;
    if (!subsample) {
        HAR_CUDA(d_loff_all.alloc(F + 1, s));
        HAR_CUDA(cudaMemcpyAsync(d_loff_all.p, m.loff.data(), sizeof(int32_t) * (F + 1), cudaMemcpyHostToDevice, s));
    }
    const size_t hist_smem = (HIST_TILE_WORDS + 32 + HIST_FTAB + HIST_FTAB / 16) * sizeof(uint32_t);
    HAR_CUDA(cudaFuncSetAttribute(k_node_hist<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hist_smem));
    HAR_CUDA(cudaFuncSetAttribute(k_node_hist<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hist_smem));
    // lane = feature histogram kernel: every node sees the same <= 48 features, few enough (bin, class) pairs per feature
    // for two [key][32] tiles, and row * ldb fits 32 bits
    int max_hbins = 0;
    for (int f = 0; f < F; f++) max_hbins = std::max(max_hbins, (int)m.hist_bins[f]);
    const int histl_keys = (max_hbins * C + 15) & ~15;
    bool hist_lanes = !subsample && C <= 8 && F <= HISTL_MAX_FEATURES && histl_keys <= HISTL_MAX_KEYS && (int64_t)n * ldb < (1LL << 32);
    if (const char* e = std::getenv("HAR_HIST_LANES")) hist_lanes = hist_lanes && std::atoi(e) != 0;
    const size_t histl_smem = histl_smem_bytes(histl_keys, F);
    // packed exchange (two 16-bit counters per word) for batches of small nodes: needs every feature to start on an even counter
    bool pack16_ok = hist_lanes;
    for (int f = 0; f <= F; f++) pack16_ok = pack16_ok && (m.loff[f] % 2 == 0);
    pack16_ok = pack16_ok && budget_words < (1LL << 31);      // exchange offsets are 32-bit
    if (const char* e = std::getenv("HAR_PACK16")) pack16_ok = pack16_ok && std::atoi(e) != 0;
    bool any_cat = false;
    for (int f = 0; f < F; f++) any_cat = any_cat || m.kind[f] != KIND_CONT;
    auto* hist_lanes_fn = any_cat ? k_node_hist_lanes<true> : k_node_hist_lanes<false>;
    if (hist_lanes) HAR_CUDA(cudaFuncSetAttribute(hist_lanes_fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)histl_smem));
    const size_t bs_smem = ((size_t)BS_ROWS * (C + 1) + 64 * C) * 4 * (BS_THREADS / 32) + MAX_CAT * (BS_THREADS / 32) * 12 + 128 +
                           sizeof(Cand) * (BS_THREADS / 32 + 1) + 64;
    HAR_CUDA(cudaFuncSetAttribute(k_best_split, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bs_smem));

    // ---- tree sets -----------------------------------------------------------------------------------
    // With all features per node a node's split depends on nothing but its own rows, so the order in which
    // nodes are trained is free (only the *count* of Spark node groups is observable; it is recomputed from the
    // finished trees, see simulate_groups).  The trees are then dealt into two sets that run the
    // selectNodesToSplit / findBestSplits loop independently and take turns on the one stream: while the GPU
    // works on a batch of one set, the host does the driver-side bookkeeping and the descriptors of the other.
    // With per-node feature subsets the pop order feeds the subset RNG, so there is one set and Spark's order.
    struct PinnedScratch {
        Pinned<NodeDesc> h_nodes; Pinned<NodeResult> h_res; Pinned<uint32_t> h_res_stats;
        Pinned<int32_t> h_item_off, h_hitem_off, h_sub;
        Pinned<BatchInfo> h_binfo; Pinned<int32_t> h_n, h_own;
        Pinned<PoolMeta> h_meta;
    };
    // page-locked staging, kept across fits (cudaMallocHost is far too slow to repeat per call)
    static thread_local PinnedScratch pinned[2];
    struct FitSet {
        std::vector<GroupItem> stack, group;
        std::vector<NodeResult> group_res;
        std::vector<uint32_t> group_stats;
        std::vector<uint32_t*> group_slot;                  // per node of the group: the slot its counters may be kept in
        std::vector<int> tkeys;                             // driver-side scratch, reused across groups
        std::vector<char> seen;
        std::vector<std::vector<int>> by_tree, pushes;
        std::vector<std::vector<uint32_t*>> released;
        const NodeResult* res_ptr = nullptr;
        const uint32_t* stat_ptr = nullptr;
        int64_t ng_total = 0, g0 = 0, cur_g0 = 0;
        int cur_ng = 0;
        bool inflight = false, done = false;
        DevBuf<uint32_t> d_hist, d_res_stats;               // d_res_stats: [node][left | right | parent][C]
        DevBuf<uint32_t> d_hist16;                          // packed counters of a batch of small nodes (multi-GPU exchange)
        size_t cap_hist16 = 0;
        DevBuf<NodeDesc> d_nodes;
        DevBuf<NodeResult> d_res;
        DevBuf<int32_t> d_item_off, d_hitem_off, d_chunk_left, d_flist_all;
        DevBuf<uint32_t> d_lbits;                            // routing bits of the rows being partitioned
        DevBuf<int32_t> d_own;                               // per batch: first node / accumulated rank of every rank's run
        size_t cap_own = 0;
        size_t cap_nodes = 0, cap_hist = 0, cap_items = 0, cap_sub = 0, cap_lbits = 0;
        HistStore store;
        int64_t batch_stamp = 0;
        // device-driven growth (all-features fits): node tables of two consecutive levels, per-batch launch data,
        // the pool of finished nodes and the kept counters of two consecutive levels
        std::vector<int> my_trees;
        std::vector<std::pair<int64_t, int>> levels;         // (first pool index, nodes) of every trained level
        DevBuf<NodeDesc> lv_nodes[2];
        DevBuf<int32_t> lv_item_off[2], lv_hitem_off[2], d_cnt, d_total;      // d_total: [children in total, copy for the host]
        DevBuf<BatchInfo> d_binfo;
        size_t cap_lv[2] = {0, 0}, cap_cnt = 0, cap_binfo = 0;
        DevBuf<NodeResult> pool_res;
        DevBuf<uint32_t> pool_stats, kept[2];
        DevBuf<PoolMeta> pool_meta;
        size_t cap_pool = 0, cap_kept[2] = {0, 0};
        int64_t pool_used = 0, kept_slots[2] = {0, 0};
        int level = 0, ncur = 0, cur = 0;
        PinnedScratch* pin = nullptr;
        cudaStream_t st = nullptr;                          // the set's stream (forked from the caller's)
        cudaEvent_t ev = nullptr;
        ~FitSet() { if (ev) cudaEventDestroy(ev); }
    };
    int num_sets = (!subsample && T >= 2) ? 2 : 1;
    if (const char* e = std::getenv("HAR_FIT_SETS")) num_sets = std::max(1, std::min(num_sets, std::atoi(e)));
    // each set on its own stream: the all-reduce of one set's batch overlaps the other set's kernels
    bool own_streams = num_sets > 1;
    if (const char* e = std::getenv("HAR_FIT_STREAMS")) own_streams = own_streams && std::atoi(e) > 1;
    // device-driven growth (all features per node), see further down
    bool device_grow = !subsample;
    if (const char* e = std::getenv("HAR_FIT_DEVICE_GROW")) device_grow = device_grow && std::atoi(e) != 0;
    // Several ranks, the library's own NCCL communicator: the best-split scan is SHARED.  Every tree has an owner rank
    // (contiguous blocks of each set's tree order); a batch's accumulated counters are summed onto the owners of their
    // nodes (one ncclReduce per rank and batch instead of an all-reduce: half the bytes), the owner derives the
    // siblings, scans its nodes, keeps their counters, and the 72-byte results + class counts are broadcast back.
    // Same integers, same fp64 sequence per node, same tie order: the forest does not change.  HAR_SHARD_SCAN=0 or a
    // callback transport keep the replicated scan behind an all-reduce.
    bool shard_scan = device_grow && world > 1 && world <= MAX_OWNERS;
    if (const char* e = std::getenv("HAR_SHARD_SCAN")) shard_scan = shard_scan && std::atoi(e) != 0;
    if (shard_scan) shard_scan = dist_has_nccl(dist, s);      // collective decision, identical on every rank
    // tests: HAR_DEBUG_XBUF=1 takes the sharded path (owner table, exchange buffer with packed small nodes, unpack) with
    // a single rank, where every exchange is the identity; that is how a one-GPU box covers this code
    if (world == 1 && device_grow)
        if (const char* e = std::getenv("HAR_DEBUG_XBUF")) shard_scan = std::atoi(e) != 0;
    const int my_rank = dist ? dist->rank : 0;
    DevBuf<int32_t> d_owner;                                  // [T] owner rank of every tree
    std::vector<int32_t> h_owner;
    const int64_t slot_words = (int64_t)full_len + C;
    int batch_cap = (int)std::min<int64_t>(1 << 20, std::max<int64_t>(2, budget_words / std::max<int64_t>(1, slot_words)));
    batch_cap &= ~1;                                                    // sibling pairs never straddle by parity alone
    int64_t kept_cap_slots = (int64_t)((2LL << 30) / (slot_words * 4)); // <= 2 GB of kept counters per level buffer
    if (const char* e = std::getenv("HAR_KEPT_SLOTS")) kept_cap_slots = std::max<int64_t>(0, std::atoll(e));   // tests
    SideStream set_side[2];
    FitSet sets[2];
    for (int k = 0; k < num_sets; k++) {
        sets[k].pin = &pinned[k];
        sets[k].st = s;
        if (device_grow) {
            // Everything the level loop needs is sized from bounds and allocated here, on the caller's stream, before
            // the set streams fork: growing a buffer inside the loop makes the stream-ordered allocator reuse memory
            // across the two set streams, which serialises them (measured: 40 ms of host time per fit).
            FitSet& S = sets[k];
            for (int t = k; t < T; t += num_sets) S.my_trees.push_back(t);
            const int64_t Ts = std::max<int64_t>(1, (int64_t)S.my_trees.size());
            const int64_t rows_g = std::max<int64_t>(1, (dist && dist->world_size > 1) ? dist->global_rows : n);
            int64_t lvl = std::min<int64_t>(Ts << std::min(p->max_depth, 22), rows_g * Ts);
            lvl = std::max<int64_t>(Ts, std::min<int64_t>(lvl, 4 << 20));
            const int64_t nbl = ceil_div(lvl, batch_cap) + 2;
            for (int w = 0; w < 2; w++) {
                S.cap_lv[w] = (size_t)lvl;
                HAR_CUDA(S.lv_nodes[w].alloc(lvl, s));
                HAR_CUDA(S.lv_item_off[w].alloc(lvl + nbl + 1, s));
                HAR_CUDA(S.lv_hitem_off[w].alloc(lvl + nbl + 1, s));
            }
            S.cap_binfo = (size_t)nbl;
            HAR_CUDA(S.d_binfo.alloc(nbl, s)); HAR_CUDA(S.pin->h_binfo.reserve(nbl)); HAR_CUDA(S.pin->h_n.reserve(2));
            S.cap_cnt = (size_t)lvl;
            HAR_CUDA(S.d_cnt.alloc(lvl, s)); HAR_CUDA(S.d_total.alloc(2, s));
            const int64_t per_batch = std::min<int64_t>(batch_cap, lvl);
            S.cap_hist = (size_t)(per_batch * slot_words);
            HAR_CUDA(S.d_hist.alloc(S.cap_hist, s));
            if (shard_scan) { S.cap_own = (size_t)nbl * 3 * (world + 1); HAR_CUDA(S.d_own.alloc(S.cap_own, s)); HAR_CUDA(S.pin->h_own.reserve(S.cap_own)); }
            S.cap_items = (size_t)(ceil_div(n * Ts, 256) + per_batch + 1);
            HAR_CUDA(S.d_chunk_left.alloc(S.cap_items, s));
            S.cap_lbits = (size_t)((n * Ts + per_batch * 4096) / 32 + 1024);   // items x chunk_rows <= rows + nodes x chunk_rows
            HAR_CUDA(S.d_lbits.alloc(S.cap_lbits, s));
            int64_t kslots = 0;
            if (subtract && p->max_depth >= 2) kslots = std::min(kept_cap_slots, std::min(lvl, Ts << std::min(p->max_depth - 2, 22)));
            for (int w = 0; w < 2; w++) {
                S.cap_kept[w] = (size_t)kslots;
                if (kslots > 0) HAR_CUDA(S.kept[w].alloc((size_t)kslots * (size_t)slot_words, s));
            }
            int64_t pool = std::min<int64_t>(Ts << std::min(p->max_depth + 1, 23), 2 * rows_g * Ts);
            pool = std::max<int64_t>(Ts, std::min<int64_t>(pool, 4 << 20));
            S.cap_pool = (size_t)pool;
            HAR_CUDA(S.pool_res.alloc(pool, s, true)); HAR_CUDA(S.pool_stats.alloc((size_t)pool * 3 * C, s, true)); HAR_CUDA(S.pool_meta.alloc(pool, s, true));   // move into the forest
        }
        if (own_streams) { HAR_CUDA(set_side[k].fork(s)); sets[k].st = set_side[k].s; }
        sets[k].store.slot_words = (size_t)full_len + C;
        HAR_CUDA(cudaEventCreateWithFlags(&sets[k].ev, cudaEventDisableTiming));
    }
    for (int t = 0; t < T; t++) sets[t % num_sets].stack.push_back({t, 0});
    if (shard_scan) {
        h_owner.assign(T, 0);
        for (int k = 0; k < num_sets; k++) {
            const int nr = (int)sets[k].my_trees.size();
            for (int q = 0; q < nr; q++) h_owner[sets[k].my_trees[q]] = (int32_t)(((int64_t)q * world) / nr);
        }
        HAR_CUDA(d_owner.alloc(T, s));
        HAR_CUDA(cudaMemcpyAsync(d_owner.p, h_owner.data(), sizeof(int32_t) * T, cudaMemcpyHostToDevice, s));
        HAR_CUDA(cudaStreamSynchronize(s));                   // h_owner is read again only by the host
    }
    const int32_t* owner_p = shard_scan ? d_owner.p : nullptr;
    // exchange buffer (equal segments per rank, reduce-scatter): with the lane = feature histogram kernel only
    const int xmode = (shard_scan && hist_lanes) ? (pack16_ok ? 2 : 1) : 0;
    // derived siblings are subtracted on the fly by the split scan and k_hist_save instead of being written out first
    int fuse_derive = 1;
    if (const char* e = std::getenv("HAR_FUSE_DERIVE")) fuse_derive = std::atoi(e) != 0;
    int64_t groups_run = 0;

    mark("buffers");
    if (trace) cudaStreamSynchronize(s);
    const double t_setup = ms_since(t_begin);
    if (trace) fprintf(stderr, "[har_forest_fit]   setup marks (ms):%s | synced %.2f\n", setup_marks.c_str(), t_setup);

    // ---- selectNodesToSplit ----
    auto select_group = [&](FitSet& S) {
        S.group.clear();
        int64_t mem = 0;
        while (!S.stack.empty()) {
            GroupItem top = S.stack.back();
            std::vector<int32_t> subset;
            int64_t tb = total_bins_all;
            if (subsample) {
                subset.resize(K);
                reservoir_features(F, K, rng.next_long(), subset.data());
                tb = 0;
                for (int j = 0; j < K; j++) tb += m.num_bins[subset[j]];
            }
            const int64_t cost = (int64_t)C * tb * 8;
            if (mem + cost <= p->max_memory_bytes || mem == 0) {
                S.stack.pop_back();
                trees[top.tree][top.node].subset.swap(subset);
                S.group.push_back(top);
                mem += cost;
            } else break;
        }
        groups_run++;
        S.ng_total = (int64_t)S.group.size();
        S.res_ptr = nullptr; S.stat_ptr = nullptr;
        if (subtract) S.group_slot.assign(S.ng_total, nullptr);
        S.g0 = 0;
    };

    // ---- findBestSplits: enqueue one device batch of the set's group (bounded by the counter budget) ----
    auto enqueue_batch = [&](FitSet& S) -> int {
        auto t_a = now();
        {
            int64_t g1 = S.g0, words = 0, rows_total = 0;
            bool all_last_level = true;
            while (g1 < S.ng_total) {
                const LNode& nd = trees[S.group[g1].tree][S.group[g1].node];
                int64_t len = full_len;
                if (subsample) { len = 0; for (int f : nd.subset) len += (int64_t)m.hist_bins[f] * C; }
                if (g1 > S.g0 && (words + len > budget_words || g1 - S.g0 >= (1 << 20))) break;
                words += len; rows_total += nd.count;
                if (nd.level + 1 < p->max_depth) all_last_level = false;
                g1++;
            }
            const int ng = (int)(g1 - S.g0);
            fo->num_batches++;
            int chunk_rows = (int)std::min<int64_t>(4096, std::max<int64_t>(256, rows_total / (kNumSMs * 8)));
            chunk_rows = (chunk_rows + 255) & ~255;
            // host-side descriptors
            HAR_CUDA(S.pin->h_nodes.reserve(ng)); HAR_CUDA(S.pin->h_item_off.reserve(ng + 1)); HAR_CUDA(S.pin->h_res.reserve(ng)); HAR_CUDA(S.pin->h_res_stats.reserve((size_t)ng * 3 * C));
            if (subsample) HAR_CUDA(S.pin->h_sub.reserve((size_t)ng * (K + 1) * 2));
            HAR_CUDA(S.pin->h_hitem_off.reserve(ng + 1));
            // sibling subtraction: pair up children whose parent's counters were kept and that sit in this batch;
            // the child with fewer weighted instances (global class totals, so every rank picks the same one) is
            // accumulated from rows, the other derived.  Counter layout of the batch buffer:
            //   [accumulated nodes][class totals of all nodes][derived nodes]   (memset / all-reduce cover the first two)
            S.batch_stamp++;
            int n_derived = 0;
            if (subtract) {
                for (int g = 0; g < ng; g++) {
                    LNode& nd = trees[S.group[S.g0 + g].tree][S.group[S.g0 + g].node];
                    nd.batch_pos = g; nd.batch_stamp = S.batch_stamp;
                }
            }
            for (int g = 0; g < ng; g++) {                          // who is derived, who may be kept
                NodeDesc& d = S.pin->h_nodes.p[g];
                d.sib = -1; d.from = nullptr; d.save = nullptr;
                if (!subtract) continue;
                const int t = S.group[S.g0 + g].tree, ni = S.group[S.g0 + g].node;
                const LNode& nd = trees[t][ni];
                if (nd.level + 1 < p->max_depth) {                 // its children can still be split: counters may be kept
                    HAR_CUDA(S.store.get(&S.group_slot[S.g0 + g], S.st));
                    d.save = S.group_slot[S.g0 + g];
                }
                if (nd.parent < 0) continue;
                const LNode& par = trees[t][nd.parent];
                if (!par.kept) continue;
                const LNode& sb = trees[t][par.left == ni ? par.right : par.left];
                if (sb.batch_stamp != S.batch_stamp) continue;
                if (nd.derive_pref) { d.sib = sb.batch_pos; d.from = par.kept; n_derived++; }
            }
            int64_t hb = 0, items = 0, hitems = 0, hist_rows = 0;
            int64_t derived_base = (int64_t)(ng - n_derived) * full_len + (int64_t)ng * C;   // used when !subsample
            for (int g = 0; g < ng; g++) {
                const LNode& nd = trees[S.group[S.g0 + g].tree][S.group[S.g0 + g].node];
                NodeDesc& d = S.pin->h_nodes.p[g];
                d.start = nd.start; d.count = nd.count; d.tree = S.group[S.g0 + g].tree; d.buf = nd.buf;
                d.level = nd.level; d.id = nd.id; d.wtotal = 0xffffffffu;
                if (subsample) {
                    d.sub_off = g * (K + 1); d.nf = K;
                    int32_t* fl = S.pin->h_sub.p + (size_t)g * (K + 1);
                    int32_t* lo = S.pin->h_sub.p + (size_t)ng * (K + 1) + (size_t)g * (K + 1);
                    int32_t off = 0;
                    for (int j = 0; j < K; j++) { fl[j] = nd.subset[j]; lo[j] = off; off += m.hist_bins[nd.subset[j]] * C; }
                    fl[K] = 0; lo[K] = off;
                    d.hist_len = off;
                } else { d.sub_off = 0; d.nf = F; d.hist_len = (int32_t)full_len; }
                if (d.sib < 0) { d.hist_base = hb; hb += d.hist_len; }
                else { d.hist_base = derived_base; derived_base += d.hist_len; }
                S.pin->h_item_off.p[g] = (int32_t)items;
                items += ceil_div(nd.count, chunk_rows);
                S.pin->h_hitem_off.p[g] = (int32_t)hitems;
                if (d.sib < 0) { hitems += ceil_div(nd.count, chunk_rows); hist_rows += nd.count; }
            }
            S.pin->h_item_off.p[ng] = (int32_t)items;
            S.pin->h_hitem_off.p[ng] = (int32_t)hitems;
            // [0, hb): accumulated counters; [hb, hb + ng*C): class totals; then the derived nodes' counters
            const size_t reduce_words = (size_t)hb + (size_t)ng * C;
            const size_t need_hist = reduce_words + (size_t)n_derived * full_len;
            if (items > 0x7fffffff) return fail(HAR_E_UNSUPPORTED, "too many work items in one batch");
            // device buffers (grow-only)
            if ((size_t)ng > S.cap_nodes) {
                S.cap_nodes = (size_t)ng * 2;
                HAR_CUDA(S.d_nodes.alloc(S.cap_nodes, S.st)); HAR_CUDA(S.d_res.alloc(S.cap_nodes, S.st)); HAR_CUDA(S.d_item_off.alloc(S.cap_nodes + 1, S.st));
                HAR_CUDA(S.d_res_stats.alloc(S.cap_nodes * 3 * C, S.st));
                HAR_CUDA(S.d_hitem_off.alloc(S.cap_nodes + 1, S.st));
            }
            if (need_hist > S.cap_hist) { S.cap_hist = need_hist + need_hist / 4; HAR_CUDA(S.d_hist.alloc(S.cap_hist, S.st)); }
            if ((size_t)items + 1 > S.cap_items) { S.cap_items = (size_t)items * 2 + 1; HAR_CUDA(S.d_chunk_left.alloc(S.cap_items, S.st)); }
            {
                const size_t lw = (size_t)items * (size_t)(chunk_rows >> 5) + 1;
                if (lw > S.cap_lbits) { S.cap_lbits = lw + lw / 2; HAR_CUDA(S.d_lbits.alloc(S.cap_lbits, S.st)); }
            }
            if (subsample && (size_t)ng * (K + 1) * 2 > S.cap_sub) {
                S.cap_sub = (size_t)ng * (K + 1) * 4;
                HAR_CUDA(S.d_flist_all.alloc(S.cap_sub, S.st));
            }
            const int32_t* flist_p = nullptr;
            const int32_t* loff_p = d_loff_all.p;
            if (subsample) {
                HAR_CUDA(cudaMemcpyAsync(S.d_flist_all.p, S.pin->h_sub.p, sizeof(int32_t) * (size_t)ng * (K + 1) * 2, cudaMemcpyHostToDevice, S.st));
                flist_p = S.d_flist_all.p;
                loff_p = S.d_flist_all.p + (size_t)ng * (K + 1);
            }
            HAR_CUDA(cudaMemcpyAsync(S.d_nodes.p, S.pin->h_nodes.p, sizeof(NodeDesc) * ng, cudaMemcpyHostToDevice, S.st));
            HAR_CUDA(cudaMemcpyAsync(S.d_item_off.p, S.pin->h_item_off.p, sizeof(int32_t) * (ng + 1), cudaMemcpyHostToDevice, S.st));
            HAR_CUDA(cudaMemcpyAsync(S.d_hitem_off.p, S.pin->h_hitem_off.p, sizeof(int32_t) * (ng + 1), cudaMemcpyHostToDevice, S.st));
            HAR_CUDA(cudaMemsetAsync(S.d_hist.p, 0, sizeof(uint32_t) * reduce_words, S.st));
            uint32_t* d_parent = S.d_hist.p + hb;
            if (hitems > 0) {
                prof_begin(S.st);
                if (subsample)
                    k_node_hist<true><<<(unsigned)hitems, HIST_THREADS, hist_smem, S.st>>>(S.d_nodes.p, S.d_hitem_off.p, ng, chunk_rows, d_bins.p, ldb,
                        d_lab.p, d_w, n, perm0.p, perm1.p, flist_p, loff_p, d_kind.p, C, S.d_hist.p, d_parent);
                else if (hist_lanes)
                    hist_lanes_fn<<<(unsigned)hitems, HIST_THREADS, histl_smem, S.st>>>(S.d_nodes.p, S.d_hitem_off.p, ng, chunk_rows, d_bins.p, (uint32_t)ldb,
                        d_lab.p, d_w, n, perm0.p, perm1.p, loff_p, d_kind.p, C, F, histl_keys, 1, S.d_hist.p, d_parent, nullptr);
                else
                    k_node_hist<false><<<(unsigned)hitems, HIST_THREADS, hist_smem, S.st>>>(S.d_nodes.p, S.d_hitem_off.p, ng, chunk_rows, d_bins.p, ldb,
                        d_lab.p, d_w, n, perm0.p, perm1.p, nullptr, loff_p, d_kind.p, C, S.d_hist.p, d_parent);
                HAR_LAUNCH_CHECK();
                prof_end(PROF_NODE_HIST, S.st, (double)hist_rows);
            }
            // the only exchange on the path: reduceByKey(_.merge(_)) -> all-reduce of the batch's counters
            if (int rc = dist_allreduce(dist, S.d_hist.p, (int64_t)reduce_words, S.st)) return rc;
            const unsigned copy_slices = (unsigned)std::max<int64_t>(1, std::min<int64_t>(8, full_len / 2048));
            if (n_derived > 0) {
                k_hist_derive<<<dim3((unsigned)ng, copy_slices), 256, 0, S.st>>>(S.d_nodes.p, C, S.d_hist.p, d_parent, nullptr, 0);
                HAR_LAUNCH_CHECK();
            }
            prof_begin(S.st);
            k_best_split<<<ng, BS_THREADS, bs_smem, S.st>>>(S.d_nodes.p, ng, S.d_hist.p, d_parent, flist_p, loff_p, subsample ? 1 : 0, fm, S.d_res.p, S.d_res_stats.p);
            HAR_LAUNCH_CHECK();
            prof_end(PROF_BEST_SPLIT, S.st, (double)ng);
            if (subtract && !all_last_level) {
                k_hist_save<<<dim3((unsigned)ng, copy_slices), 256, 0, S.st>>>(S.d_nodes.p, S.d_res.p, C, S.d_hist.p, d_parent, nullptr, 0);
                HAR_LAUNCH_CHECK();
            }
            if (!all_last_level && items > 0) {
                prof_begin(S.st);
                k_part_count<<<(unsigned)items, HIST_THREADS, 0, S.st>>>(S.d_nodes.p, S.d_item_off.p, ng, chunk_rows, S.d_res.p, d_bins.p, ldb, n,
                                                                      perm0.p, perm1.p, S.d_chunk_left.p, S.d_lbits.p, d_binsT.p, nT);
                HAR_LAUNCH_CHECK();
                prof_end(PROF_PART_COUNT, S.st, (double)rows_total);
                k_part_scan<<<(unsigned)ceil_div(ng, 4), 128, 0, S.st>>>(S.d_item_off.p, ng, S.d_res.p, S.d_chunk_left.p);
                HAR_LAUNCH_CHECK();
                prof_begin(S.st);
                k_part_scatter<<<(unsigned)items, HIST_THREADS, 0, S.st>>>(S.d_nodes.p, S.d_item_off.p, ng, chunk_rows, S.d_res.p, d_bins.p, ldb, n,
                                                                        perm0.p, perm1.p, S.d_chunk_left.p, S.d_lbits.p);
                HAR_LAUNCH_CHECK();
                prof_end(PROF_PART_SCATTER, S.st, (double)rows_total);
            }
            HAR_CUDA(cudaMemcpyAsync(S.pin->h_res.p, S.d_res.p, sizeof(NodeResult) * ng, cudaMemcpyDeviceToHost, S.st));
            HAR_CUDA(cudaMemcpyAsync(S.pin->h_res_stats.p, S.d_res_stats.p, sizeof(uint32_t) * (size_t)ng * 3 * C, cudaMemcpyDeviceToHost, S.st));
            HAR_CUDA(cudaEventRecord(S.ev, S.st));
            S.cur_g0 = S.g0; S.cur_ng = ng; S.g0 = g1; S.inflight = true;
            t_prep += ms_since(t_a);
        }
        return HAR_OK;
    };

    auto collect_batch = [&](FitSet& S) -> int {
        auto t_a = now();
        HAR_CUDA(cudaEventSynchronize(S.ev));
        t_wait += ms_since(t_a);
        t_a = now();
        {
            if (S.cur_ng == S.ng_total) { S.res_ptr = S.pin->h_res.p; S.stat_ptr = S.pin->h_res_stats.p; }   // whole S.group in one batch: read the staging buffers in place
            else {
                if ((int64_t)S.group_res.size() < S.ng_total) S.group_res.resize(S.ng_total);
                std::copy(S.pin->h_res.p, S.pin->h_res.p + S.cur_ng, S.group_res.begin() + S.cur_g0);
                S.res_ptr = S.group_res.data();
                if (S.group_stats.size() < (size_t)S.ng_total * 3 * C) S.group_stats.resize((size_t)S.ng_total * 3 * C);
                std::copy(S.pin->h_res_stats.p, S.pin->h_res_stats.p + (size_t)S.cur_ng * 3 * C, S.group_stats.begin() + (size_t)S.cur_g0 * 3 * C);
                S.stat_ptr = S.group_stats.data();
            }
        }
        S.inflight = false;
        t_post += ms_since(t_a);
        return HAR_OK;
    };

    auto post_group = [&](FitSet& S) {
        auto t_a = now();
        // ---- driver side of findBestSplits: children, leaf rules, S.stack pushes ----
        // Trees are independent, so they are processed by a few host threads; the pushes are then appended
        // in the order Spark's `nodesForGroup.foreach` would produce (Scala mutable.HashMap order of the tree
        // indices, pop order inside a tree).
        std::vector<int>& tkeys = S.tkeys;
        std::vector<std::vector<int>>&by_tree = S.by_tree, &pushes = S.pushes;
        std::vector<std::vector<uint32_t*>>& released = S.released;
        if (by_tree.empty()) { by_tree.resize(T); pushes.resize(T); released.resize(T); S.seen.assign(T, 0); }
        tkeys.clear();
        for (auto& gi : S.group) if (!S.seen[gi.tree]) { S.seen[gi.tree] = 1; tkeys.push_back(gi.tree); }
        for (int t : tkeys) { S.seen[t] = 0; by_tree[t].clear(); pushes[t].clear(); released[t].clear(); }
        std::vector<int> torder = scala_hashmap_order(tkeys);
        for (int g = 0; g < (int)S.ng_total; g++) by_tree[S.group[g].tree].push_back(g);
        auto process_tree = [&](int t) {
            std::vector<LNode>& tr = trees[t];
            std::vector<uint32_t>& ts = tree_stats[t];
            tr.reserve(tr.size() + 2 * by_tree[t].size());
            ts.reserve((tr.size() + 2 * by_tree[t].size()) * C);
            for (int g : by_tree[t]) {
                const NodeResult& r = S.res_ptr[g];
                const uint32_t* rs = S.stat_ptr + (size_t)g * 3 * C;
                const int ni = S.group[g].node;
                const int level = tr[ni].level;
                const bool is_leaf = !(r.gain > 0.0) || level == p->max_depth;
                uint32_t* slot = subtract ? S.group_slot[g] : nullptr;
                {
                    LNode& nd = tr[ni];
                    if (nd.parent >= 0) {                           // this child is done: drop the parent's kept counters after both
                        LNode& par = tr[nd.parent];
                        if (par.kept && --par.kids_pending == 0) { released[t].push_back(par.kept); par.kept = nullptr; }
                    }
                    nd.is_leaf = is_leaf; nd.valid = r.valid != 0; nd.gain = r.gain; nd.impurity = r.impurity;
                    for (int c = 0; c < C; c++) ts[(size_t)ni * C + c] = rs[2 * C + c];
                    if (!nd.subset.empty()) std::vector<int32_t>().swap(nd.subset);
                }
                if (is_leaf) { if (slot) released[t].push_back(slot); continue; }
                const bool child_leaf = level + 1 == p->max_depth;
                const bool ll = child_leaf || r.left_imp == 0.0, rl = child_leaf || r.right_imp == 0.0;
                const bool parted = !(ll && rl);
                const int li = (int)tr.size();
                tr.emplace_back(); tr.emplace_back();
                LNode& nd = tr[ni];
                LNode& L = tr[li]; LNode& R = tr[li + 1];
                L.id = nd.id * 2; R.id = L.id + 1;
                L.parent = R.parent = ni;
                L.level = R.level = level + 1;
                {   // global class totals, so every rank derives the same child
                    uint64_t wl = 0, wr = 0;
                    for (int c = 0; c < C; c++) { wl += rs[c]; wr += rs[C + c]; }
                    R.derive_pref = wl <= wr; L.derive_pref = !R.derive_pref;
                }
                if (slot) {
                    if (r.retain) { nd.kept = slot; nd.kids_pending = 2; }
                    else released[t].push_back(slot);
                }
                L.is_leaf = ll; R.is_leaf = rl; L.impurity = r.left_imp; R.impurity = r.right_imp;
                ts.resize((size_t)(li + 2) * C);
                for (int c = 0; c < C; c++) { ts[(size_t)li * C + c] = rs[c]; ts[(size_t)(li + 1) * C + c] = rs[C + c]; }
                L.start = nd.start; L.count = parted ? r.n_left : 0; L.buf = nd.buf ^ 1;
                R.start = nd.start + L.count; R.count = parted ? nd.count - r.n_left : 0; R.buf = nd.buf ^ 1;
                nd.left = li; nd.right = li + 1;
                nd.feature = r.feature; nd.kind = r.kind; nd.split = r.split; nd.mask = r.left_mask; nd.arity = m.arity[r.feature];
                nd.threshold = r.kind == KIND_CONT ? fo->thresholds[(size_t)r.feature * MB + r.split] : 0.0;
                L.pushed = !ll; R.pushed = !rl;
                if (!ll) pushes[t].push_back(li);
                if (!rl) pushes[t].push_back(li + 1);
            }
        };
        parallel_for((int)tkeys.size(), S.ng_total >= 512, [&](int k) { process_tree(tkeys[k]); });
        for (int oi : torder) {
            const int t = tkeys[oi];
            for (int ni : pushes[t]) S.stack.push_back({t, ni});
            for (uint32_t* sl : released[t]) S.store.put(sl);
        }
        t_post += ms_since(t_a);
    };

    // one step of a set: finish its batch in flight, then enqueue the next one (of this group, or of the next group)
    auto advance = [&](FitSet& S) -> int {
        if (S.inflight) { if (int rc = collect_batch(S)) return rc; }
        if (S.g0 == S.ng_total) {
            if (S.ng_total > 0) post_group(S);
            if (S.stack.empty()) { S.done = true; S.ng_total = 0; S.g0 = 0; return HAR_OK; }
            auto t_a = now();
            select_group(S);
            t_prep += ms_since(t_a);
        }
        return enqueue_batch(S);
    };

    // ---- device-driven growth (all features per node) ------------------------------------------------------
    // The host enqueues a whole level of a set at a time (its device batches, then the kernels that build the next
    // level's node table) and reads back only the per-batch launch data; see k_grow_* in har_tree_kernels.cuh.

    auto dd_ensure_level = [&](FitSet& S, int which, int64_t nodes_cap) -> int {
        const int64_t nb = ceil_div(std::max<int64_t>(nodes_cap, 1), batch_cap);
        if ((size_t)nodes_cap > S.cap_lv[which]) {
            S.cap_lv[which] = (size_t)nodes_cap * 2;
            const int64_t nb2 = ceil_div((int64_t)S.cap_lv[which], batch_cap);
            HAR_CUDA(S.lv_nodes[which].alloc(S.cap_lv[which], S.st));
            HAR_CUDA(S.lv_item_off[which].alloc(S.cap_lv[which] + nb2 + 1, S.st));
            HAR_CUDA(S.lv_hitem_off[which].alloc(S.cap_lv[which] + nb2 + 1, S.st));
        }
        if ((size_t)nb > S.cap_binfo) {
            S.cap_binfo = (size_t)nb * 2;
            HAR_CUDA(S.d_binfo.alloc(S.cap_binfo, S.st));
            HAR_CUDA(S.pin->h_binfo.reserve(S.cap_binfo));
        }
        if (shard_scan && (size_t)nb * 3 * (world + 1) > S.cap_own) {
            S.cap_own = (size_t)nb * 4 * (world + 1);
            HAR_CUDA(S.d_own.alloc(S.cap_own, S.st));
            HAR_CUDA(S.pin->h_own.reserve(S.cap_own));
        }
        return HAR_OK;
    };
    auto dd_ensure_kept = [&](FitSet& S, int which, int level_of_nodes, int64_t nodes_upper) -> int {
        // nodes past the preallocated slots simply do not keep their counters (their children are accumulated from rows)
        const int64_t slots = (subtract && level_of_nodes + 1 < p->max_depth) ? std::min<int64_t>(nodes_upper, (int64_t)S.cap_kept[which]) : 0;
        S.kept_slots[which] = slots;
        return HAR_OK;
    };
    // launch data of level `S.level` (already laid out on the device) -> host, behind an event
    auto dd_fetch_info = [&](FitSet& S, int64_t nb_upper) -> int {
        HAR_CUDA(S.pin->h_n.reserve(2));
        HAR_CUDA(cudaMemcpyAsync(S.pin->h_binfo.p, S.d_binfo.p, sizeof(BatchInfo) * nb_upper, cudaMemcpyDeviceToHost, S.st));
        HAR_CUDA(cudaMemcpyAsync(S.pin->h_n.p, S.d_total.p + 1, sizeof(int32_t), cudaMemcpyDeviceToHost, S.st));
        if (shard_scan)
            HAR_CUDA(cudaMemcpyAsync(S.pin->h_own.p, S.d_own.p, sizeof(int32_t) * nb_upper * 3 * (world + 1), cudaMemcpyDeviceToHost, S.st));
        HAR_CUDA(cudaEventRecord(S.ev, S.st));
        S.inflight = true;
        return HAR_OK;
    };
    auto dd_start = [&](FitSet& S) -> int {
        auto t_a = now();
        const int nr = (int)S.my_trees.size();
        S.level = 0; S.cur = 0; S.ncur = nr;
        if (nr == 0) { S.done = true; return HAR_OK; }
        if (int rc = dd_ensure_level(S, 0, nr)) return rc;
        if (int rc = dd_ensure_kept(S, 0, 0, nr)) return rc;
        HAR_CUDA(S.pin->h_nodes.reserve(nr));
        for (int q = 0; q < nr; q++) {
            NodeDesc& d = S.pin->h_nodes.p[q];
            const int t = S.my_trees[q];
            d.start = 0; d.hist_base = 0; d.count = root_cnt[t]; d.tree = t; d.buf = 0; d.level = 0;
            d.sub_off = 0; d.nf = F; d.hist_len = (int32_t)full_len; d.sib = -1; d.from = nullptr;
            d.save = q < S.kept_slots[0] ? S.kept[0].p + (int64_t)q * slot_words : nullptr;
            d.id = 1; d.wtotal = 0xffffffffu;
        }
        HAR_CUDA(cudaMemcpyAsync(S.lv_nodes[0].p, S.pin->h_nodes.p, sizeof(NodeDesc) * nr, cudaMemcpyHostToDevice, S.st));
        HAR_CUDA(S.pin->h_n.reserve(2));
        S.pin->h_n.p[1] = nr;
        HAR_CUDA(cudaMemcpyAsync(S.d_total.p, S.pin->h_n.p + 1, sizeof(int32_t), cudaMemcpyHostToDevice, S.st));
        const int64_t nb = ceil_div(nr, batch_cap);
        k_grow_layout<<<(unsigned)nb, 256, 0, S.st>>>(S.lv_nodes[0].p, S.d_total.p, batch_cap, C, kNumSMs, S.lv_item_off[0].p,
                                                     S.lv_hitem_off[0].p, S.d_binfo.p, S.d_total.p + 1, owner_p, world, S.d_own.p, xmode);
        HAR_LAUNCH_CHECK();
        t_prep += ms_since(t_a);
        return dd_fetch_info(S, nb);
    };
    // one level of a set: wait for its launch data, enqueue its batches and the construction of the next level
    auto dd_advance = [&](FitSet& S) -> int {
        auto t_a = now();
        HAR_CUDA(cudaEventSynchronize(S.ev));
        S.inflight = false;
        t_wait += ms_since(t_a);
        t_a = now();
        const int ncur = S.pin->h_n.p[0];
        S.ncur = ncur;
        if (ncur == 0) { S.done = true; return HAR_OK; }
        const int L = S.level, cur = S.cur, nxt = cur ^ 1;
        const int64_t nb = ceil_div(ncur, batch_cap);
        const int64_t base = S.pool_used;
        if ((size_t)(base + ncur) > S.cap_pool) {                       // grow the pool of finished nodes, keeping its content
            const size_t ncap = std::max<size_t>((size_t)(base + ncur) * 2, 4096);
            DevBuf<NodeResult> nr_; DevBuf<uint32_t> ns_; DevBuf<PoolMeta> nm_;
            HAR_CUDA(nr_.alloc(ncap, S.st, true)); HAR_CUDA(ns_.alloc(ncap * 3 * C, S.st, true)); HAR_CUDA(nm_.alloc(ncap, S.st, true));
            if (base > 0) {
                HAR_CUDA(cudaMemcpyAsync(nr_.p, S.pool_res.p, sizeof(NodeResult) * base, cudaMemcpyDeviceToDevice, S.st));
                HAR_CUDA(cudaMemcpyAsync(ns_.p, S.pool_stats.p, sizeof(uint32_t) * base * 3 * C, cudaMemcpyDeviceToDevice, S.st));
                HAR_CUDA(cudaMemcpyAsync(nm_.p, S.pool_meta.p, sizeof(PoolMeta) * base, cudaMemcpyDeviceToDevice, S.st));
            }
            S.pool_res.swap(nr_); S.pool_stats.swap(ns_); S.pool_meta.swap(nm_);
            S.cap_pool = ncap;
        }
        const bool last_level = L + 1 >= p->max_depth;                  // no child of this level can be split
        const int32_t* loff_p = d_loff_all.p;
        for (int64_t b = 0; b < nb; b++) {
            const BatchInfo bi = S.pin->h_binfo.p[b];
            const int ng = bi.ng;
            if (ng <= 0) continue;
            fo->num_batches++;
            if (trace) fprintf(stderr, "[har_forest_fit]   level %d batch %lld: %d nodes (%d derived)\n", L, (long long)b, ng, bi.n_derived);
            const size_t reduce_words = (size_t)bi.hb_direct + (size_t)ng * C;        // accumulated counters + class totals
            const size_t need_hist = reduce_words + (size_t)bi.n_derived * full_len;
            if (need_hist > S.cap_hist) { S.cap_hist = need_hist + need_hist / 4; HAR_CUDA(S.d_hist.alloc(S.cap_hist, S.st)); }
            if ((size_t)bi.items + 1 > S.cap_items) { S.cap_items = (size_t)bi.items * 2 + 1; HAR_CUDA(S.d_chunk_left.alloc(S.cap_items, S.st)); }
            {
                const size_t lw = (size_t)bi.items * (size_t)(bi.chunk_rows >> 5) + 1;
                if (lw > S.cap_lbits) { S.cap_lbits = lw + lw / 2; HAR_CUDA(S.d_lbits.alloc(S.cap_lbits, S.st)); }
            }
            const NodeDesc* nodes = S.lv_nodes[cur].p + b * batch_cap;
            const int32_t* io = S.lv_item_off[cur].p + b * (batch_cap + 1);
            const int32_t* ho = S.lv_hitem_off[cur].p + b * (batch_cap + 1);
            NodeResult* res = S.pool_res.p + base + b * batch_cap;
            uint32_t* rstats = S.pool_stats.p + (base + b * batch_cap) * 3 * C;
            uint32_t* d_parent = S.d_hist.p + bi.hb_direct;
            const bool pack16 = xmode != 0;               // counters travel through the exchange buffer
            if (pack16) {
                const size_t need16 = (size_t)bi.xwords;
                if (need16 > S.cap_hist16) { S.cap_hist16 = need16 + need16 / 4; HAR_CUDA(S.d_hist16.alloc(S.cap_hist16, S.st)); }
                HAR_CUDA(cudaMemsetAsync(S.d_hist16.p, 0, sizeof(uint32_t) * need16, S.st));
                HAR_CUDA(cudaMemsetAsync(d_parent, 0, sizeof(uint32_t) * (size_t)ng * C, S.st));
            } else {
                HAR_CUDA(cudaMemsetAsync(S.d_hist.p, 0, sizeof(uint32_t) * reduce_words, S.st));
            }
            if (bi.hitems > 0) {
                prof_begin(S.st);
                if (hist_lanes)
                    hist_lanes_fn<<<(unsigned)bi.hitems, HIST_THREADS, histl_smem, S.st>>>(nodes, ho, ng, bi.chunk_rows, d_bins.p, (uint32_t)ldb,
                        d_lab.p, d_w, n, perm0.p, perm1.p, loff_p, d_kind.p, C, F, histl_keys, 1, S.d_hist.p, d_parent,
                        pack16 ? S.d_hist16.p : nullptr);
                else
                    k_node_hist<false><<<(unsigned)bi.hitems, HIST_THREADS, hist_smem, S.st>>>(nodes, ho, ng, bi.chunk_rows, d_bins.p, ldb,
                        d_lab.p, d_w, n, perm0.p, perm1.p, nullptr, loff_p, d_kind.p, C, S.d_hist.p, d_parent);
                HAR_LAUNCH_CHECK();
                prof_end(PROF_NODE_HIST, S.st, (double)bi.hist_rows);
            }
            // the exchange (reduceByKey(_.merge(_))): counters onto the owners of their nodes, or an all-reduce
            const int32_t* own = shard_scan ? S.pin->h_own.p + b * 3 * (world + 1) : nullptr;
            int g_lo = 0, g_hi = ng;
            if (shard_scan) {
                int32_t* ptrs[2 * MAX_OWNERS]; int64_t cnts[2 * MAX_OWNERS]; int roots[2 * MAX_OWNERS];
                if (pack16) {
                    int32_t* tot = reinterpret_cast<int32_t*>(d_parent);
                    const int64_t tot_cnt = (int64_t)ng * C;
                    if (int rc = dist_scatter_exchange(dist, reinterpret_cast<int32_t*>(S.d_hist16.p), bi.xseg, 1, &tot, &tot_cnt, S.st)) return rc;
                } else {
                    for (int r = 0; r < world; r++) {
                        ptrs[2 * r] = reinterpret_cast<int32_t*>(S.d_hist.p) + (int64_t)own[world + 1 + r] * full_len;
                        cnts[2 * r] = (int64_t)(own[world + 1 + r + 1] - own[world + 1 + r]) * full_len;
                        ptrs[2 * r + 1] = reinterpret_cast<int32_t*>(d_parent) + (int64_t)own[r] * C;
                        cnts[2 * r + 1] = (int64_t)(own[r + 1] - own[r]) * C;
                        roots[2 * r] = roots[2 * r + 1] = r;
                    }
                    if (int rc = dist_reduce_list(dist, 2 * world, ptrs, cnts, roots, S.st)) return rc;
                }
                g_lo = own[my_rank]; g_hi = own[my_rank + 1];
                if (pack16 && g_hi > g_lo) {                        // the owned nodes' counters to their 32-bit places
                    const unsigned sl = (unsigned)std::max<int64_t>(1, std::min<int64_t>(8, full_len / 2048));
                    k_hist_unpack<<<dim3((unsigned)(g_hi - g_lo), sl), 256, 0, S.st>>>(nodes, g_lo, S.d_hist16.p, S.d_hist.p);
                    HAR_LAUNCH_CHECK();
                }
            } else if (int rc = dist_allreduce(dist, S.d_hist.p, (int64_t)reduce_words, S.st)) return rc;
            const unsigned copy_slices = (unsigned)std::max<int64_t>(1, std::min<int64_t>(8, full_len / 2048));
            if (bi.n_derived > 0) {
                k_hist_derive<<<dim3((unsigned)ng, fuse_derive ? 1u : copy_slices), fuse_derive ? 32 : 256, 0, S.st>>>(
                    nodes, C, S.d_hist.p, d_parent, owner_p, my_rank, fuse_derive);
                HAR_LAUNCH_CHECK();
            }
            if (shard_scan) {
                // results travel as one sum over the ranks: every rank but the owner of a node contributes zero bits
                HAR_CUDA(cudaMemsetAsync(res, 0, sizeof(NodeResult) * (size_t)ng, S.st));
                HAR_CUDA(cudaMemsetAsync(rstats, 0, sizeof(uint32_t) * (size_t)ng * 3 * C, S.st));
            }
            if (g_hi > g_lo) {
                prof_begin(S.st);
                k_best_split<<<g_hi - g_lo, BS_THREADS, bs_smem, S.st>>>(nodes, ng, S.d_hist.p, d_parent, nullptr, loff_p, 0, fm, res, rstats, g_lo, fuse_derive);
                HAR_LAUNCH_CHECK();
                prof_end(PROF_BEST_SPLIT, S.st, (double)(g_hi - g_lo));
            }
            if (!last_level && subtract && S.kept_slots[cur] > 0) {
                // before the results travel: a rank keeps the counters of the nodes it owns, from its own results
                k_hist_save<<<dim3((unsigned)ng, copy_slices), 256, 0, S.st>>>(nodes, res, C, S.d_hist.p, d_parent, owner_p, my_rank, fuse_derive);
                HAR_LAUNCH_CHECK();
            }
            if (shard_scan) {
                static_assert(sizeof(NodeResult) % 4 == 0, "NodeResult travels as int32 words");
                int32_t* ptrs[2] = {reinterpret_cast<int32_t*>(res), reinterpret_cast<int32_t*>(rstats)};
                const int64_t cnts[2] = {(int64_t)ng * (int64_t)(sizeof(NodeResult) / 4), (int64_t)ng * 3 * C};
                if (int rc = dist_scatter_exchange(dist, nullptr, 0, 2, ptrs, cnts, S.st)) return rc;
            }
            if (!last_level) {
                if (bi.items > 0) {
                    prof_begin(S.st);
                    k_part_count<<<(unsigned)bi.items, HIST_THREADS, 0, S.st>>>(nodes, io, ng, bi.chunk_rows, res, d_bins.p, ldb, n,
                                                                               perm0.p, perm1.p, S.d_chunk_left.p, S.d_lbits.p, d_binsT.p, nT);
                    HAR_LAUNCH_CHECK();
                    prof_end(PROF_PART_COUNT, S.st, (double)bi.rows_total);
                    k_part_scan<<<(unsigned)ceil_div(ng, 4), 128, 0, S.st>>>(io, ng, res, S.d_chunk_left.p);
                    HAR_LAUNCH_CHECK();
                    prof_begin(S.st);
                    k_part_scatter<<<(unsigned)bi.items, HIST_THREADS, 0, S.st>>>(nodes, io, ng, bi.chunk_rows, res, d_bins.p, ldb, n,
                                                                                 perm0.p, perm1.p, S.d_chunk_left.p, S.d_lbits.p);
                    HAR_LAUNCH_CHECK();
                    prof_end(PROF_PART_SCATTER, S.st, (double)bi.rows_total);
                }
            }
        }
        S.pool_used = base + ncur;
        S.levels.emplace_back(base, ncur);
        // ---- the next level's node table, on the device ----
        const int64_t next_upper = 2LL * ncur;
        if (int rc = dd_ensure_level(S, nxt, next_upper)) return rc;
        if (int rc = dd_ensure_kept(S, nxt, L + 1, next_upper)) return rc;
        if ((size_t)ncur > S.cap_cnt) { S.cap_cnt = (size_t)ncur * 2; HAR_CUDA(S.d_cnt.alloc(S.cap_cnt, S.st)); }
        const unsigned gb = (unsigned)ceil_div(ncur, 256);
        k_grow_flags<<<gb, 256, 0, S.st>>>(S.lv_nodes[cur].p, S.pool_res.p + base, ncur, p->max_depth, S.d_cnt.p);
        HAR_LAUNCH_CHECK();
        k_grow_scan<<<1, 1024, 0, S.st>>>(S.d_cnt.p, ncur, S.d_total.p);
        HAR_LAUNCH_CHECK();
        k_grow_emit<<<gb, 256, 0, S.st>>>(S.lv_nodes[cur].p, S.pool_res.p + base, S.pool_stats.p + base * 3 * C, S.d_cnt.p, ncur, C,
                                          p->max_depth, batch_cap, (int32_t)full_len, F, (int)(base + ncur), S.kept[nxt].p,
                                          S.kept_slots[nxt], slot_words, S.pool_meta.p + base, S.lv_nodes[nxt].p);
        HAR_LAUNCH_CHECK();
        const int64_t nb_next = ceil_div(next_upper, batch_cap);
        k_grow_layout<<<(unsigned)nb_next, 256, 0, S.st>>>(S.lv_nodes[nxt].p, S.d_total.p, batch_cap, C, kNumSMs, S.lv_item_off[nxt].p,
                                                          S.lv_hitem_off[nxt].p, S.d_binfo.p, S.d_total.p + 1, owner_p, world, S.d_own.p,
                                                          xmode);
        HAR_LAUNCH_CHECK();
        S.level = L + 1; S.cur = nxt;
        groups_run++;
        t_prep += ms_since(t_a);
        return dd_fetch_info(S, nb_next);
    };
    if (int rc = wait_roots()) return rc;
    if (!device_grow) {
        for (int t = 0; t < T; t++) {
            LNode r; r.id = 1; r.start = 0; r.count = root_cnt[t]; r.buf = 0;
            trees[t].reserve(256);
            trees[t].push_back(r);
            tree_stats[t].assign(C, 0);
        }
    }
    if (device_grow) {
        for (int k = 0; k < num_sets; k++) if (int rc = dd_start(sets[k])) return rc;
        for (;;) {
            bool any = false;
            for (int k = 0; k < num_sets; k++) {
                if (sets[k].done) continue;
                if (int rc = dd_advance(sets[k])) return rc;
                any = any || !sets[k].done;
            }
            if (!any) break;
        }
    } else
    for (;;) {
        bool any = false;
        for (int k = 0; k < num_sets; k++) {
            if (sets[k].done) continue;
            if (int rc = advance(sets[k])) return rc;
            any = any || !sets[k].done;
        }
        if (!any) break;
    }
    for (int k = 0; k < num_sets; k++) HAR_CUDA(set_side[k].join(s));
    const bool spark_order = num_sets == 1 && !device_grow;       // the groups that ran are Spark's groups
    if (spark_order) fo->num_groups = groups_run;
    auto t_emit = now();

    // ---- toNode(prune = true): per tree in parallel, then concatenated in tree order ----
    if (!device_grow) {
        emit_forest(fo.get(), T, C, trees, tree_stats, trace);
        if (!spark_order) {
            // Spark's node-group count is not needed to train or predict: recount it on demand (har_forest_fit_info)
            auto kept = std::make_shared<std::vector<std::vector<LNode>>>(std::move(trees));
            const int64_t cost = (int64_t)C * total_bins_all * 8, max_mem = p->max_memory_bytes;
            fo->count_groups = [kept, cost, max_mem]() { return simulate_groups(*kept, cost, max_mem); };
        }
    } else {
        // ---- toNode(prune = true) on the device, straight into the predict tables (k_fin_*) ----
        int64_t npool_all = 0;
        for (int k = 0; k < num_sets; k++) npool_all += sets[k].pool_used;
        const int64_t slots_cap = (int64_t)T + 2 * npool_all;
        har_forest* f = fo.get();
        int dev_now = 0;
        HAR_CUDA(cudaGetDevice(&dev_now));
        HAR_CUDA(cudaMallocAsync(&f->d_nodes, sizeof(PNode) * slots_cap, s));
        HAR_CUDA(cudaMallocAsync((void**)&f->d_slot_stats, sizeof(uint32_t) * slots_cap * C, s));
        HAR_CUDA(cudaMallocAsync((void**)&f->d_roots, sizeof(int64_t) * T, s));
        f->dev = dev_now; f->leaf_mode = -1;
        DevBuf<double> d_thr; DevBuf<int32_t> d_ar, d_tot;
        HAR_CUDA(d_thr.alloc((size_t)F * MB, s)); HAR_CUDA(d_ar.alloc(F, s)); HAR_CUDA(d_tot.alloc(2, s));
        HAR_CUDA(cudaMemcpyAsync(d_thr.p, fo->thresholds.data(), sizeof(double) * (size_t)F * MB, cudaMemcpyHostToDevice, s));
        HAR_CUDA(cudaMemcpyAsync(d_ar.p, m.arity.data(), sizeof(int32_t) * F, cudaMemcpyHostToDevice, s));
        HAR_CUDA(cudaMemsetAsync(d_tot.p, 0, sizeof(int32_t) * 2, s));
        std::vector<int64_t> roots(T);
        for (int t = 0; t < T; t++) roots[t] = t;
        HAR_CUDA(cudaMemcpyAsync(f->d_roots, roots.data(), sizeof(int64_t) * T, cudaMemcpyHostToDevice, s));
        DevBuf<uint8_t> d_leafflag[2], d_alive[2];
        DevBuf<int32_t> d_pred[2], d_rank[2], d_slot[2];
        for (int k = 0; k < num_sets; k++) {
            FitSet& S = sets[k];
            const int np = (int)S.pool_used, nroots = (int)S.my_trees.size();
            if (np == 0) continue;
            HAR_CUDA(d_leafflag[k].alloc(np, s)); HAR_CUDA(d_alive[k].alloc(np, s));
            HAR_CUDA(d_pred[k].alloc(np, s)); HAR_CUDA(d_rank[k].alloc(np, s)); HAR_CUDA(d_slot[k].alloc(np, s));
            HAR_CUDA(cudaMemsetAsync(d_alive[k].p, 0, np, s));
            HAR_CUDA(cudaMemsetAsync(d_alive[k].p, 1, nroots, s));
            for (int L = (int)S.levels.size() - 1; L >= 0; L--) {
                const int base = (int)S.levels[L].first, nl = S.levels[L].second;
                k_fin_prune<<<(unsigned)ceil_div(nl, 256), 256, 0, s>>>(S.pool_res.p, S.pool_stats.p, S.pool_meta.p, base, nl, L,
                                                                       p->max_depth, C, d_leafflag[k].p, d_pred[k].p);
                HAR_LAUNCH_CHECK();
            }
            for (int L = 0; L < (int)S.levels.size(); L++) {
                const int base = (int)S.levels[L].first, nl = S.levels[L].second;
                k_fin_alive<<<(unsigned)ceil_div(nl, 256), 256, 0, s>>>(S.pool_meta.p, base, nl, d_leafflag[k].p, d_alive[k].p, d_rank[k].p);
                HAR_LAUNCH_CHECK();
            }
            k_grow_scan<<<1, 1024, 0, s>>>(d_rank[k].p, np, d_tot.p + k);
            HAR_LAUNCH_CHECK();
            const int32_t* rank_base = k == 0 ? nullptr : d_tot.p;
            const unsigned gb = (unsigned)ceil_div(np, 256);
            k_fin_children<<<gb, 256, 0, s>>>(S.pool_meta.p, S.pool_stats.p, np, nroots, d_rank[k].p, rank_base, T, C, d_alive[k].p,
                                              d_leafflag[k].p, d_slot[k].p, (PNode*)f->d_nodes, f->d_slot_stats);
            HAR_LAUNCH_CHECK();
            k_fin_nodes<<<gb, 256, 0, s>>>(S.pool_res.p, S.pool_stats.p, np, d_rank[k].p, rank_base, T, C, d_alive[k].p, d_leafflag[k].p,
                                           d_slot[k].p, d_thr.p, MB, d_ar.p, (PNode*)f->d_nodes, f->d_slot_stats);
            HAR_LAUNCH_CHECK();
        }
        int32_t h_tot[2] = {0, 0};
        HAR_CUDA(cudaMemcpyAsync(h_tot, d_tot.p, sizeof h_tot, cudaMemcpyDeviceToHost, s));
        HAR_CUDA(cudaStreamSynchronize(s));
        f->total_nodes = (int64_t)T + 2 * ((int64_t)h_tot[0] + h_tot[1]);
        f->num_slots = f->total_nodes;
        // the pools move into the forest: flat host arrays and Spark's group count are rebuilt from them on demand
        auto lazy = std::make_shared<LazyPools>();
        lazy->T = T; lazy->C = C; lazy->MB = MB; lazy->max_depth = p->max_depth; lazy->num_sets = num_sets; lazy->dev = dev_now;
        lazy->arity = m.arity; lazy->thresholds = fo->thresholds;
        for (int k = 0; k < num_sets; k++) {
            FitSet& S = sets[k];
            LazyPools::SetPool& sp = lazy->set[k];
            sp.my_trees = S.my_trees; sp.N = S.pool_used;
            sp.res = S.pool_res.detach(); sp.stats = S.pool_stats.detach(); sp.meta = S.pool_meta.detach();
        }
        const int64_t cost = (int64_t)C * total_bins_all * 8, max_mem = p->max_memory_bytes;
        fo->count_groups = [lazy, cost, max_mem]() -> int64_t {
            if (lazy->build() != HAR_OK) return -1;
            return simulate_groups(lazy->trees, cost, max_mem);
        };
        fo->materialize = [lazy, f, T, C, trace]() -> int {
            if (int rc = lazy->build()) return rc;
            emit_forest(f, T, C, lazy->trees, lazy->tree_stats, trace);
            return HAR_OK;
        };
    }
    fo->num_launches = launches_since_reset() - launches0;
    if (trace)
        fprintf(stderr, "[har_forest_fit] setup %.2f ms | loop: host prep+launch %.2f, gpu wait %.2f, driver-side %.2f | toNode %.2f | "
                        "total %.2f ms, %lld groups run, %lld batches, %zu nodes\n", t_setup, t_prep, t_wait, t_post, ms_since(t_emit),
                ms_since(t_begin), (long long)groups_run, (long long)fo->num_batches, (size_t)fo->total_nodes);
    (void)world;
    *out = fo.release();
    return HAR_OK;
}

HAR_API int har_forest_shape(const har_forest_t* f, int32_t* num_trees, int32_t* num_classes, int32_t* num_features,
                             int64_t* total_nodes) {
    if (!f) return fail(HAR_E_INVALID, "har_forest_shape: NULL forest");
    if (num_trees) *num_trees = f->T;
    if (num_classes) *num_classes = f->C;
    if (num_features) *num_features = f->F;
    if (total_nodes) *total_nodes = f->total_nodes >= 0 ? f->total_nodes : (int64_t)f->feature.size();
    return HAR_OK;
}

HAR_API int har_forest_export(const har_forest_t* f, int64_t* tree_off, int32_t* feature, int32_t* left, int32_t* right,
                              int32_t* node_id, int32_t* arity, double* threshold, uint64_t* left_mask, double* prediction,
                              double* impurity, double* gain, double* stats) {
    if (!f) return fail(HAR_E_INVALID, "har_forest_export: NULL forest");
    if (int rc = har_forest_ensure_host(f)) return rc;
    const size_t N = f->feature.size();
    if (tree_off) std::copy(f->tree_off.begin(), f->tree_off.end(), tree_off);
    if (feature) std::copy(f->feature.begin(), f->feature.end(), feature);
    if (left) std::copy(f->left.begin(), f->left.end(), left);
    if (right) std::copy(f->right.begin(), f->right.end(), right);
    if (node_id && f->node_id.size() == N) std::copy(f->node_id.begin(), f->node_id.end(), node_id);
    if (arity) std::copy(f->arity.begin(), f->arity.end(), arity);
    if (threshold) std::copy(f->threshold.begin(), f->threshold.end(), threshold);
    if (left_mask) std::copy(f->left_mask.begin(), f->left_mask.end(), left_mask);
    if (prediction && f->prediction.size() == N) std::copy(f->prediction.begin(), f->prediction.end(), prediction);
    if (impurity && f->impurity.size() == N) std::copy(f->impurity.begin(), f->impurity.end(), impurity);
    if (gain && f->gain.size() == N) std::copy(f->gain.begin(), f->gain.end(), gain);
    if (stats) std::copy(f->stats.begin(), f->stats.end(), stats);
    return HAR_OK;
}

HAR_API int har_forest_fit_info(const har_forest_t* f, double* thresholds, int32_t* nsplits, int32_t max_bins,
                                int64_t* num_groups, int64_t* num_batches, int64_t* num_kernel_launches) {
    if (!f) return fail(HAR_E_INVALID, "har_forest_fit_info: NULL forest");
    if ((thresholds || nsplits) && (max_bins != f->max_bins || f->nsplits.empty()))
        return fail(HAR_E_INVALID, "har_forest_fit_info: forest was not fitted with max_bins = %d", max_bins);
    if (thresholds) std::copy(f->thresholds.begin(), f->thresholds.end(), thresholds);
    if (nsplits) std::copy(f->nsplits.begin(), f->nsplits.end(), nsplits);
    if (num_groups) {
        std::lock_guard<std::mutex> lk(f->count_mu);
        if (f->count_groups) { f->num_groups = f->count_groups(); f->count_groups = nullptr; }
        *num_groups = f->num_groups;
    }
    if (num_batches) *num_batches = f->num_batches;
    if (num_kernel_launches) *num_kernel_launches = f->num_launches;
    return HAR_OK;
}

// bit-for-bit comparison of the count division used by k_best_split with __ddiv_rn over n pseudo-random operand pairs
HAR_API int har_debug_div_check(long long n, unsigned long long seed, unsigned long long* mismatches, har_stream_t stream) {
    cudaStream_t s = (cudaStream_t)stream;
    DevBuf<unsigned long long> d;
    HAR_CUDA(d.alloc(1, s));
    HAR_CUDA(cudaMemsetAsync(d.p, 0, sizeof(unsigned long long), s));
    if (n > 0) { k_div_check<<<(unsigned)ceil_div(n, 256), 256, 0, s>>>(n, seed, d.p); HAR_LAUNCH_CHECK(); }
    HAR_CUDA(cudaMemcpyAsync(mismatches, d.p, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    HAR_CUDA(cudaStreamSynchronize(s));
    return HAR_OK;
}
