// The exchanges on the path (reduceByKey(_.merge(_)) in RandomForest.findBestSplits): int32
// sum all-reduce of a batch's counters, and — for the node-sharded best-split scan — reduce-scatter of the counters
// by node followed by an all-gather of the per-node results.
//
// The C ABI takes the collective as a callback (har_dist.allreduce_i32) so that any transport can sit behind it.
// When the process already has NCCL loaded (torch does), the library binds to that same libnccl at run time and
// builds its own communicator over the ranks of the har_dist, then issues the collectives itself on the stream of
// the tree set that needs them: no host round trip through the caller's runtime per batch, and the transfer of one
// set overlaps the kernels of the other.  No link-time dependency: without NCCL in the process the callback is used.
//
// Which transport carries a group's exchanges is decided ONCE per (group_id, world, rank, device), collectively and
// identically on every rank, through the callback:
//   1. every rank contributes 34 words to one callback all-reduce: the 128-byte NCCL unique id (rank 0 only; zeros
//      elsewhere) and one word saying "NCCL is usable here" (libnccl loaded, HAR_NCCL != 0, id obtained);
//   2. only if every rank said yes do the ranks call ncclCommInitRank; a second callback all-reduce of one word then
//      tells everybody whether every rank got a communicator.  If not, the ranks that did abort theirs.
// A rank never leaves the callback path on its own, so the ranks cannot end up on different transports.  The reason
// for a fall-back is kept with the decision (har_dist_transport) and left in har_last_error().
#include <dlfcn.h>

#include <atomic>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "har_common.cuh"

namespace har {

namespace {
struct NcclId { char internal[128]; };
using nccl_get_id_fn = int (*)(NcclId*);
using nccl_init_fn = int (*)(void**, int, NcclId, int);
using nccl_allreduce_fn = int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t);
using nccl_reduce_scatter_fn = int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t);
using nccl_allgather_fn = int (*)(const void*, void*, size_t, int, void*, cudaStream_t);
using nccl_group_fn = int (*)();
using nccl_reduce_fn = int (*)(const void*, void*, size_t, int, int, int, void*, cudaStream_t);
using nccl_bcast_fn = int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t);
using nccl_comm_fn = int (*)(void*);
using nccl_errstr_fn = const char* (*)(int);
constexpr int kNcclInt8 = 0, kNcclInt32 = 2, kNcclSum = 0;

struct NcclApi {
    nccl_get_id_fn get_id = nullptr;
    nccl_init_fn init = nullptr;
    nccl_allreduce_fn allreduce = nullptr;
    nccl_reduce_fn reduce = nullptr;
    nccl_reduce_scatter_fn reduce_scatter = nullptr;
    nccl_bcast_fn bcast = nullptr;
    nccl_group_fn group_start = nullptr, group_end = nullptr;
    nccl_comm_fn destroy = nullptr, abort = nullptr;
    nccl_errstr_fn errstr = nullptr;
    bool ok = false;
    const char* why = "";
    NcclApi() {
        if (const char* e = std::getenv("HAR_NCCL")) if (std::atoi(e) == 0) { why = "HAR_NCCL=0"; return; }
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);      // only a libnccl that is already in the process
        if (!h) { why = "libnccl.so.2 is not loaded in this process"; return; }
        get_id = (nccl_get_id_fn)dlsym(h, "ncclGetUniqueId");
        init = (nccl_init_fn)dlsym(h, "ncclCommInitRank");
        allreduce = (nccl_allreduce_fn)dlsym(h, "ncclAllReduce");
        reduce = (nccl_reduce_fn)dlsym(h, "ncclReduce");
        reduce_scatter = (nccl_reduce_scatter_fn)dlsym(h, "ncclReduceScatter");
        bcast = (nccl_bcast_fn)dlsym(h, "ncclBroadcast");
        group_start = (nccl_group_fn)dlsym(h, "ncclGroupStart");
        group_end = (nccl_group_fn)dlsym(h, "ncclGroupEnd");
        destroy = (nccl_comm_fn)dlsym(h, "ncclCommDestroy");
        abort = (nccl_comm_fn)dlsym(h, "ncclCommAbort");
        errstr = (nccl_errstr_fn)dlsym(h, "ncclGetErrorString");
        ok = get_id && init && allreduce && reduce && reduce_scatter && bcast && group_start && group_end && destroy && abort && errstr;
        if (!ok) why = "libnccl.so.2 lacks a needed symbol";
    }
};
NcclApi& nccl() { static NcclApi a; return a; }

// One decision per group of ranks, made collectively on first use and kept until har_dist_release().
struct CommSlot {
    uint64_t group = 0;
    int world = 0, rank = -1, dev = -1;
    void* comm = nullptr;          // non-NULL: the library's NCCL communicator carries the batch exchanges
    std::string why;               // why not, when comm == NULL
};
std::mutex g_comm_mu;
std::vector<CommSlot> g_slots;
std::atomic<long long> g_calls{0}, g_bytes{0};

// 34-word exchange buffer of the transport decision: part of the library image, so that taking part in the decision
// cannot fail on an allocation on one rank only (which would leave the other ranks waiting in the collective).
constexpr int kXchgWords = 34;
__device__ int32_t g_xchg[64];

int callback_allreduce(const har_dist* d, void* buf, int64_t count, cudaStream_t s) {
    if (!d->allreduce_i32) return fail(HAR_E_INVALID, "har_dist.allreduce_i32 is NULL with world_size > 1");
    int rc = d->allreduce_i32(d->ctx, buf, count, (har_stream_t)s);
    if (rc != 0) return fail(HAR_E_COMM, "all-reduce callback failed with code %d", rc);
    return HAR_OK;
}

bool slot_matches(const CommSlot& c, const har_dist* d, int dev) {
    return c.group == d->group_id && c.world == d->world_size && c.rank == d->rank && c.dev == dev;
}

// Transport for this har_dist; collective on first use (every rank reaches its first exchange together).
// *out = communicator or NULL (callback).  Fails only if the callback itself fails.
int get_comm(const har_dist* d, cudaStream_t s, void** out) {
    *out = nullptr;
    int dev = 0;
    HAR_CUDA(cudaGetDevice(&dev));
    {
        std::lock_guard<std::mutex> lk(g_comm_mu);
        for (const CommSlot& c : g_slots)
            if (slot_matches(c, d, dev)) { *out = c.comm; return HAR_OK; }
    }
    // not decided yet: the two callback exchanges below are collective, so the mutex is not held across them
    NcclApi& api = nccl();
    CommSlot slot;
    slot.group = d->group_id; slot.world = d->world_size; slot.rank = d->rank; slot.dev = dev;
    int32_t h[kXchgWords];
    std::memset(h, 0, sizeof h);
    bool mine = api.ok;
    std::string why = api.ok ? "" : api.why;
    if (mine && d->rank == 0) {
        NcclId id;
        std::memset(&id, 0, sizeof id);
        int rc = api.get_id(&id);
        if (rc != 0) { mine = false; why = std::string("ncclGetUniqueId failed: ") + api.errstr(rc); }
        else std::memcpy(h, &id, sizeof id);
    }
    h[32] = mine ? 1 : 0;
    h[33] = 1;                                          // head count: must come back as world_size
    int32_t* d_x = nullptr;
    HAR_CUDA(cudaGetSymbolAddress((void**)&d_x, g_xchg));
    HAR_CUDA(cudaMemcpyAsync(d_x, h, sizeof h, cudaMemcpyHostToDevice, s));
    if (int rc = callback_allreduce(d, d_x, kXchgWords, s)) return rc;
    HAR_CUDA(cudaMemcpyAsync(h, d_x, sizeof h, cudaMemcpyDeviceToHost, s));
    HAR_CUDA(cudaStreamSynchronize(s));
    if (h[33] != d->world_size)
        return fail(HAR_E_COMM, "transport decision: the callback summed %d ranks, har_dist.world_size is %d", h[33], d->world_size);
    bool nonzero = false;
    for (int i = 0; i < 32; i++) nonzero = nonzero || h[i] != 0;
    if (h[32] != d->world_size || !nonzero) {
        char buf[256];
        snprintf(buf, sizeof buf, "NCCL is usable on %d of %d ranks%s%s", h[32], d->world_size, why.empty() ? "" : "; here: ", why.c_str());
        slot.why = buf;
    } else {
        NcclId id;
        std::memcpy(&id, h, sizeof id);
        void* comm = nullptr;
        int rc = api.init(&comm, d->world_size, id, d->rank);
        int32_t ok = (rc == 0 && comm) ? 1 : 0;
        if (!ok) why = std::string("ncclCommInitRank failed: ") + api.errstr(rc);
        HAR_CUDA(cudaMemcpyAsync(d_x, &ok, sizeof ok, cudaMemcpyHostToDevice, s));
        if (int rc2 = callback_allreduce(d, d_x, 1, s)) { if (comm) api.abort(comm); return rc2; }
        int32_t all = 0;
        HAR_CUDA(cudaMemcpyAsync(&all, d_x, sizeof all, cudaMemcpyDeviceToHost, s));
        HAR_CUDA(cudaStreamSynchronize(s));
        if (all == d->world_size) slot.comm = comm;
        else {
            if (comm) api.abort(comm);
            char buf[256];
            snprintf(buf, sizeof buf, "ncclCommInitRank succeeded on %d of %d ranks%s%s", all, d->world_size,
                     why.empty() ? "" : "; here: ", why.c_str());
            slot.why = buf;
        }
    }
    if (!slot.comm) last_error() = "har_b200: batch exchanges stay on the caller's all-reduce callback (" + slot.why + ")";
    std::lock_guard<std::mutex> lk(g_comm_mu);
    g_slots.push_back(slot);
    *out = slot.comm;
    return HAR_OK;
}
}  // namespace

int dist_allreduce(const har_dist* d, void* buf, int64_t count, cudaStream_t s) {
    if (!d || d->world_size <= 1 || count == 0) return HAR_OK;
    g_calls += 1; g_bytes += 4 * count;
    void* comm = nullptr;
    if (int rc = get_comm(d, s, &comm)) return rc;
    if (comm) {
        int rc = nccl().allreduce(buf, buf, (size_t)count, kNcclInt32, kNcclSum, comm, s);
        if (rc != 0) return fail(HAR_E_COMM, "ncclAllReduce failed: %s", nccl().errstr(rc));
        return HAR_OK;
    }
    return callback_allreduce(d, buf, count, s);
}

bool dist_has_nccl(const har_dist* d, cudaStream_t s) {
    if (!d || d->world_size <= 1) return false;
    void* comm = nullptr;
    return get_comm(d, s, &comm) == HAR_OK && comm != nullptr;
}

// Node-sharded exchange (NCCL only).  reduce: every op sums `cnt[k]` int32 at `ptr[k]` over all ranks onto rank
// `root[k]` (in place; the other ranks keep partial sums).  bcast: every op sends `bytes[k]` bytes at `ptr[k]` from rank
// `root[k]` to everybody (in place).  All ops of a call travel in one NCCL group; empty ops are skipped.
int dist_reduce_list(const har_dist* d, int n_ops, int32_t* const* ptr, const int64_t* cnt, const int* root, cudaStream_t s) {
    if (!d || d->world_size <= 1) return HAR_OK;
    void* comm = nullptr;
    if (int rc = get_comm(d, s, &comm)) return rc;
    if (!comm) return fail(HAR_E_COMM, "dist_reduce_list needs the library's NCCL communicator");
    NcclApi& api = nccl();
    g_calls += 1;
    int rc = api.group_start();
    for (int k = 0; k < n_ops && rc == 0; k++) {
        if (cnt[k] <= 0) continue;
        g_bytes += 4 * cnt[k];
        rc = api.reduce(ptr[k], ptr[k], (size_t)cnt[k], kNcclInt32, kNcclSum, root[k], comm, s);
    }
    int rc2 = api.group_end();
    if (rc == 0) rc = rc2;
    if (rc != 0) return fail(HAR_E_COMM, "ncclReduce (node-sharded counters) failed: %s", api.errstr(rc));
    return HAR_OK;
}

// One group: reduce-scatter of `seg` int32 per rank (rank r ends up with the sums of xbuf[r * seg, (r + 1) * seg), in
// place) plus all-reduces of the listed int32 arrays.  Two or three operations per batch instead of one per owner.
int dist_scatter_exchange(const har_dist* d, int32_t* xbuf, int64_t seg, int n_all, int32_t* const* all_ptr,
                          const int64_t* all_cnt, cudaStream_t s) {
    if (!d || d->world_size <= 1) return HAR_OK;
    void* comm = nullptr;
    if (int rc = get_comm(d, s, &comm)) return rc;
    if (!comm) return fail(HAR_E_COMM, "dist_scatter_exchange needs the library's NCCL communicator");
    NcclApi& api = nccl();
    g_calls += 1;
    int rc = api.group_start();
    if (rc == 0 && seg > 0) {
        g_bytes += 4 * seg * d->world_size;
        rc = api.reduce_scatter(xbuf, xbuf + (int64_t)d->rank * seg, (size_t)seg, kNcclInt32, kNcclSum, comm, s);
    }
    for (int k = 0; k < n_all && rc == 0; k++) {
        if (all_cnt[k] <= 0) continue;
        g_bytes += 4 * all_cnt[k];
        rc = api.allreduce(all_ptr[k], all_ptr[k], (size_t)all_cnt[k], kNcclInt32, kNcclSum, comm, s);
    }
    int rc2 = api.group_end();
    if (rc == 0) rc = rc2;
    if (rc != 0) return fail(HAR_E_COMM, "NCCL exchange (reduce-scatter / all-reduce) failed: %s", api.errstr(rc));
    return HAR_OK;
}

int dist_bcast_list(const har_dist* d, int n_ops, void* const* ptr, const int64_t* bytes, const int* root, cudaStream_t s) {
    if (!d || d->world_size <= 1) return HAR_OK;
    void* comm = nullptr;
    if (int rc = get_comm(d, s, &comm)) return rc;
    if (!comm) return fail(HAR_E_COMM, "dist_bcast_list needs the library's NCCL communicator");
    NcclApi& api = nccl();
    g_calls += 1;
    int rc = api.group_start();
    for (int k = 0; k < n_ops && rc == 0; k++) {
        if (bytes[k] <= 0) continue;
        g_bytes += bytes[k];
        rc = api.bcast(ptr[k], ptr[k], (size_t)bytes[k], kNcclInt8, root[k], comm, s);
    }
    int rc2 = api.group_end();
    if (rc == 0) rc = rc2;
    if (rc != 0) return fail(HAR_E_COMM, "ncclBroadcast (node results) failed: %s", api.errstr(rc));
    return HAR_OK;
}

}  // namespace har

HAR_API int har_dist_transport(const har_dist* dist, char* why, int32_t cap) {
    if (why && cap > 0) why[0] = 0;
    if (!dist || dist->world_size <= 1) return 0;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return -1;
    std::lock_guard<std::mutex> lk(har::g_comm_mu);
    for (const har::CommSlot& c : har::g_slots)
        if (har::slot_matches(c, dist, dev)) {
            if (why && cap > 0) snprintf(why, (size_t)cap, "%s", c.comm ? "library NCCL communicator" : c.why.c_str());
            return c.comm ? 1 : 0;
        }
    return -1;
}

HAR_API void har_dist_release(void) {
    std::vector<har::CommSlot> old;
    {
        std::lock_guard<std::mutex> lk(har::g_comm_mu);
        old.swap(har::g_slots);
    }
    for (har::CommSlot& c : old)
        if (c.comm) har::nccl().destroy(c.comm);
}

// all-reduce calls / bytes issued by this process since the last reset (bench.py; declared in include/har_b200_debug.h)
HAR_API void har_debug_comm_stats(long long* calls, long long* bytes, int* via_nccl, int reset) {
    if (calls) *calls = har::g_calls.load();
    if (bytes) *bytes = har::g_bytes.load();
    if (via_nccl) {
        std::lock_guard<std::mutex> lk(har::g_comm_mu);
        *via_nccl = 0;
        for (const har::CommSlot& c : har::g_slots) if (c.comm) *via_nccl = 1;
    }
    if (reset) { har::g_calls = 0; har::g_bytes = 0; }
}
