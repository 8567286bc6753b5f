// Device node of the predict tables (har_predict.cu walks them; har_tree_fit.cu can write them straight from the
// device-side tree pool).  16 bytes, children adjacent so that one 128-bit load serves a visit.
#pragma once
#include <cstdint>

namespace har {

struct __align__(16) PNode {
    int32_t feature;     // < 0: leaf.  bit 30: categorical, bit 29: values outside the mask domain go left
    int32_t left;        // internal: index of the left child (right = left + 1); leaf: leaf-vector index
    union { double thr; unsigned long long mask; };
};
constexpr int32_t P_CAT = 1 << 30, P_DEFLEFT = 1 << 29, P_FMASK = (1 << 29) - 1;

// CategoricalSplit keeps the smaller side: isLeft = |left| <= arity / 2 (Split.scala); the node stores the test as
// "bit v of mask set -> left", with values outside [0, arity) following the default of the stored side.
__host__ __device__ inline void pnode_set_categorical(PNode& d, int32_t feature, int arity, unsigned long long left_mask) {
    const unsigned long long dom = arity >= 64 ? ~0ULL : ((1ULL << arity) - 1);
#ifdef __CUDA_ARCH__
    const int pc = __popcll(left_mask & dom);
#else
    const int pc = __builtin_popcountll(left_mask & dom);
#endif
    const bool is_left = pc <= arity / 2;
    d.feature = feature | P_CAT;
    if (is_left) d.mask = left_mask & dom;
    else { d.mask = (left_mask & dom) | ~dom; d.feature |= P_DEFLEFT; }
}

// Compact node of the binned predict path: 8 bytes, so that the 4095 nodes of a depth-12 tree are 32 KB and the trees
// being walked stay in L1.  Rows are binned first (u8 per feature, TreePoint.findBin), and `x[f] <= threshold` becomes
// `bin[f] <= split` - every threshold of a fitted forest is one of the feature's candidate thresholds.  Categorical
// nodes keep their 64-bit mask in a side table indexed like the nodes.
struct __align__(8) CNode {
    uint16_t feature;    // leaf: unused
    uint8_t split;       // continuous: index of the threshold among the feature's candidates
    uint8_t flags;       // C_LEAF | C_CAT | C_DEFLEFT
    uint32_t left;       // internal: index of the left child (right = left + 1); leaf: leaf-vector index
};
constexpr uint8_t C_LEAF = 1, C_CAT = 2, C_DEFLEFT = 4;
constexpr uint8_t BIN_OUTSIDE = 255;   // categorical value outside [0, 64) or not an integer

// leaf vectors of the predict kernel from per-slot class counts: DT = counts, RF = counts / total
static __global__ void k_fin_leaf_table(const uint32_t* __restrict__ slot_stats, int64_t nslots, int C, int is_forest, double* __restrict__ leaf) {
    const int64_t sidx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (sidx >= nslots) return;
    const uint32_t* st = slot_stats + sidx * C;
    double total = 0.0;
    for (int c = 0; c < C; c++) total += (double)st[c];
    for (int c = 0; c < C; c++) {
        const double v = (double)st[c];
        leaf[sidx * C + c] = is_forest ? (total != 0.0 ? v / total : 0.0) : v;
    }
}

}  // namespace har
