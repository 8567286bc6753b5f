// The one exchange on the path: int32 sum all-reduce of a batch's counters (reduceByKey(_.merge(_)) in
// RandomForest.findBestSplits).
//
// The C ABI takes the collective as a callback (har_dist.allreduce_i32) so that any transport can sit behind it.
// When the process already has NCCL loaded (torch does), the library binds to that same libnccl at run time,
// builds its own communicator over the ranks of the har_dist — the 128-byte unique id is produced on rank 0 and
// exchanged ONCE through the callback (a sum with zeros from the other ranks) — and from then on issues
// ncclAllReduce itself, on the stream of the tree set that needs it: no host round trip through the caller's
// runtime per batch, and the transfer of one set overlaps the kernels of the other.  HAR_NCCL=0 keeps every
// exchange on the callback.  No link-time dependency: without NCCL in the process the callback is used.
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
using nccl_errstr_fn = const char* (*)(int);
constexpr int kNcclInt32 = 2, kNcclSum = 0;

struct NcclApi {
    nccl_get_id_fn get_id = nullptr;
    nccl_init_fn init = nullptr;
    nccl_allreduce_fn allreduce = nullptr;
    nccl_errstr_fn errstr = nullptr;
    bool ok = false;
    NcclApi() {
        if (const char* e = std::getenv("HAR_NCCL")) if (std::atoi(e) == 0) return;
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);      // only a libnccl that is already in the process
        if (!h) return;
        get_id = (nccl_get_id_fn)dlsym(h, "ncclGetUniqueId");
        init = (nccl_init_fn)dlsym(h, "ncclCommInitRank");
        allreduce = (nccl_allreduce_fn)dlsym(h, "ncclAllReduce");
        errstr = (nccl_errstr_fn)dlsym(h, "ncclGetErrorString");
        ok = get_id && init && allreduce && errstr;
    }
};
NcclApi& nccl() { static NcclApi a; return a; }

struct CommSlot { void* comm = nullptr; int world = 0, rank = -1, dev = -1; bool failed = false; };
std::mutex g_comm_mu;
CommSlot g_comm;                       // one communicator per process: keyed by (world, rank, device)
std::atomic<long long> g_calls{0}, g_bytes{0};

int callback_allreduce(const har_dist* d, void* buf, int64_t count, cudaStream_t s) {
    if (!d->allreduce_i32) return fail(HAR_E_INVALID, "har_dist.allreduce_i32 is NULL with world_size > 1");
    int rc = d->allreduce_i32(d->ctx, buf, count, (har_stream_t)s);
    if (rc != 0) return fail(HAR_E_COMM, "all-reduce callback failed with code %d", rc);
    return HAR_OK;
}

// communicator for this har_dist, created on first use (collective: every rank reaches its first exchange together)
void* get_comm(const har_dist* d, cudaStream_t s) {
    NcclApi& api = nccl();
    if (!api.ok) return nullptr;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
    std::lock_guard<std::mutex> lk(g_comm_mu);
    if (g_comm.comm && g_comm.world == d->world_size && g_comm.rank == d->rank && g_comm.dev == dev) return g_comm.comm;
    if (g_comm.failed) return nullptr;
    NcclId id;
    std::memset(&id, 0, sizeof id);
    if (d->rank == 0 && api.get_id(&id) != 0) { g_comm.failed = true; }
    // every rank must take part in the id exchange even if rank 0 failed (all zeros then: nobody initialises)
    int32_t* d_id = nullptr;
    bool good = cudaMalloc((void**)&d_id, sizeof id) == cudaSuccess;
    good = good && cudaMemcpyAsync(d_id, &id, sizeof id, cudaMemcpyHostToDevice, s) == cudaSuccess;
    good = good && callback_allreduce(d, d_id, (int64_t)(sizeof id / 4), s) == HAR_OK;
    good = good && cudaMemcpyAsync(&id, d_id, sizeof id, cudaMemcpyDeviceToHost, s) == cudaSuccess;
    good = good && cudaStreamSynchronize(s) == cudaSuccess;
    if (d_id) cudaFree(d_id);
    bool nonzero = false;
    for (char c : id.internal) nonzero = nonzero || c != 0;
    if (!good || !nonzero) { g_comm.failed = true; return nullptr; }
    void* comm = nullptr;
    int rc = api.init(&comm, d->world_size, id, d->rank);
    if (rc != 0 || !comm) { g_comm.failed = true; return nullptr; }
    g_comm.comm = comm; g_comm.world = d->world_size; g_comm.rank = d->rank; g_comm.dev = dev;
    return comm;
}
}  // namespace

int dist_allreduce(const har_dist* d, void* buf, int64_t count, cudaStream_t s) {
    if (!d || d->world_size <= 1 || count == 0) return HAR_OK;
    g_calls += 1; g_bytes += 4 * count;
    if (void* comm = get_comm(d, s)) {
        int rc = nccl().allreduce(buf, buf, (size_t)count, kNcclInt32, kNcclSum, comm, s);
        if (rc != 0) return fail(HAR_E_COMM, "ncclAllReduce failed: %s", nccl().errstr(rc));
        return HAR_OK;
    }
    return callback_allreduce(d, buf, count, s);
}

}  // namespace har

// all-reduce calls / bytes issued by this process since the last reset (bench.py; not part of the reference-facing surface)
HAR_API void har_debug_comm_stats(long long* calls, long long* bytes, int* via_nccl, int reset) {
    if (calls) *calls = har::g_calls.load();
    if (bytes) *bytes = har::g_bytes.load();
    if (via_nccl) { std::lock_guard<std::mutex> lk(har::g_comm_mu); *via_nccl = har::g_comm.comm ? 1 : 0; }
    if (reset) { har::g_calls = 0; har::g_bytes = 0; }
}
