// The opaque har_forest handle: flat host copy of the (pruned) trees + lazily built device tables.
#pragma once
#include <cstdint>
#include <functional>
#include <mutex>
#include <vector>

struct har_forest {
    int T = 0, C = 0, F = 0;
    // flat, pre-order per tree (see har_forest_export in include/har_b200.h)
    std::vector<int64_t> tree_off;
    std::vector<int32_t> feature, left, right, node_id, arity;
    std::vector<double> threshold, prediction, impurity, gain, stats;
    std::vector<uint64_t> left_mask;
    // fit by-products
    std::vector<double> thresholds;
    std::vector<int32_t> nsplits;
    int max_bins = 0;
    mutable int64_t num_groups = 0;
    int64_t num_batches = 0, num_launches = 0;
    // when the fit trained its nodes in another order than Spark's (see har_tree_fit.cu), the number of Spark node
    // groups is recounted from the finished trees the first time har_forest_fit_info asks for it
    mutable std::function<int64_t()> count_groups;
    mutable std::mutex count_mu;
    // device tables for predict (built on first use, on the device current at that time)
    int dev = -1;
    void* d_nodes = nullptr;      // PNode[total]
    double* d_leaf = nullptr;     // [num_leaves][C]: DT = class counts, RF = counts / total
    int64_t* d_roots = nullptr;   // [T] index of each tree's root in d_nodes
    int leaf_mode = -1;           // is_forest flag the leaf table was built for
    void* ev_pred = nullptr;      // cudaEvent_t recorded after the last predict: a mode switch on another stream waits for it
                                  // before it rewrites the leaf vectors
    // Fits that grew the trees on the device (har_tree_fit.cu) write the tables above straight from the device-side
    // node pool; per-slot class counts stay on the device so the leaf table can be rebuilt for the other predict mode.
    // The flat host arrays are then filled on first use (har_forest_export / import paths) by `materialize`.
    uint32_t* d_slot_stats = nullptr;   // [num_slots][C]
    int64_t num_slots = 0;
    int64_t total_nodes = -1;     // pruned node count, known before the host arrays exist
    mutable std::function<int()> materialize;
    int max_depth = 0;
    // binned predict path (har_predict.cu): compact nodes + category masks converted from d_nodes, and the device
    // copies of the split candidates the rows are binned with.  binnable: -1 not decided, 0 no (imported forest, a
    // threshold that is not a candidate, too many features), 1 yes.
    void* d_cnodes = nullptr;            // CNode[num compact]
    unsigned long long* d_cmask = nullptr;
    double* d_bin_thr = nullptr;         // [F][max_bins]
    int32_t* d_bin_ns = nullptr;         // [F]
    void* d_walk = nullptr;              // uint2[num compact]: branch-free walk table (no categorical nodes)
    uint32_t* d_leaf_of = nullptr;       // [num compact] leaf-vector index of a leaf node
    bool has_cat_nodes = false;
    int64_t num_cnodes = 0;
    int binnable = -1;
};

// Fill the flat host arrays if the fit left them on the device.  Returns 0 (HAR_OK) or the error of the read-back.
inline int har_forest_ensure_host(const har_forest* f) {
    std::lock_guard<std::mutex> lk(f->count_mu);
    if (!f->materialize) return 0;
    const int rc = f->materialize();
    if (rc == 0) f->materialize = nullptr;
    return rc;
}
