"""Multi-GPU plumbing: one process per GPU, rows sharded by Spark partition, the only exchange is
the all-reduce of node-group histograms done by torch.distributed (NCCL over
NVLink) through the har_allreduce_i32_fn callback of include/har_b200.h."""
import ctypes as C

import numpy as np
import torch

from . import _native as N


def shard_partitions(total_partitions, rank, world_size):
    """Contiguous, balanced partition range [p0, p1) owned by `rank`."""
    base, rem = divmod(total_partitions, world_size)
    p0 = rank * base + min(rank, rem)
    return p0, p0 + base + (1 if rank < rem else 0)


def shard_rows(global_part_off, rank, world_size):
    """-> (row0, row1, local part_off, partition_offset) for this rank."""
    gpo = np.ascontiguousarray(global_part_off, dtype=np.int64)
    p0, p1 = shard_partitions(len(gpo) - 1, rank, world_size)
    r0, r1 = int(gpo[p0]), int(gpo[p1])
    return r0, r1, gpo[p0:p1 + 1] - r0, p0


class _DevPtr:
    def __init__(self, ptr, count):
        self.__cuda_array_interface__ = {"shape": (count,), "typestr": "<i4", "data": (ptr, False), "version": 2}


def wrap_i32(ptr, count, device=None):
    """int32 view of `count` elements at a raw pointer: CUDA (device given) or host memory."""
    if device is not None:
        return torch.as_tensor(_DevPtr(ptr, count), device=device)
    buf = (C.c_int32 * count).from_address(ptr)
    return torch.frombuffer(buf, dtype=torch.int32)


_GROUP_IDS = {}


def _group_id(td, group):
    """Stable id of a torch process group inside this process (the library keys its transport decision on it)."""
    key = id(group) if group is not None else 0
    if key not in _GROUP_IDS:
        _GROUP_IDS[key] = len(_GROUP_IDS) + 1
    return _GROUP_IDS[key]


class Dist:
    """Builds the har_dist struct; keeps the ctypes callback alive.

    The callback hands the library's buffer to torch.distributed.all_reduce.  With an NCCL group the device buffer is
    reduced in place on the stream the library names.  With a group that cannot reduce device memory (gloo) the
    buffer is staged through the host: copy down on that stream, reduce on the CPU, copy back — slow, but it lets
    several ranks share one GPU (tests) and keeps the collective path usable without NCCL.  host_buffers=True: the
    pointers the callback receives are host pointers (CPU-only unit tests of the callback)."""

    def __init__(self, global_rows, partition_offset, total_partitions, group=None, host_buffers=False):
        import torch.distributed as td
        self._td, self._group = td, group
        self.rank, self.world = td.get_rank(group), td.get_world_size(group)
        self._device = None if host_buffers else torch.device("cuda", torch.cuda.current_device())
        self._staged = (not host_buffers) and td.get_backend(group) != "nccl"
        self.calls = 0
        self.bytes = 0

        streams = {}

        def _cb(ctx, ptr, count, stream):
            try:
                t = wrap_i32(ptr, count, self._device)
                if self._device is not None:
                    # the library may call from a stream of its own (one per tree set): order the collective there
                    st = streams.get(stream)
                    if st is None:
                        st = streams[stream] = (torch.cuda.ExternalStream(stream, device=self._device) if stream
                                                else torch.cuda.default_stream(self._device))
                    with torch.cuda.stream(st):
                        if self._staged:
                            h = t.cpu()                       # synchronises `st`
                            td.all_reduce(h, op=td.ReduceOp.SUM, group=self._group)
                            t.copy_(h)
                            st.synchronize()
                        else:
                            td.all_reduce(t, op=td.ReduceOp.SUM, group=self._group)
                else:
                    td.all_reduce(t, op=td.ReduceOp.SUM, group=self._group)
                self.calls += 1
                self.bytes += 4 * count
                return 0
            except Exception as e:      # never let an exception cross the C boundary
                print("har_b200 all-reduce callback failed:", repr(e))
                return 1

        self._cb = N.ALLREDUCE_FN(_cb)
        self.struct = N.Dist(self.rank, self.world, int(global_rows), int(partition_offset), int(total_partitions),
                             self._cb, None, _group_id(td, group))

    def transport(self):
        """(1 library NCCL communicator | 0 callback | -1 undecided, reason) — har_dist_transport."""
        buf = C.create_string_buffer(512)
        rc = N.lib().har_dist_transport(C.byref(self.struct), buf, 512)
        return rc, buf.value.decode("utf-8", "replace")


def release():
    """Destroys the library's cached NCCL communicators (har_dist_release); call before destroy_process_group."""
    N.lib().har_dist_release()


def exchange_shard_layout(local_rows, local_partitions, group=None):
    """-> (global_rows, this rank's first global partition index, total partitions), by all-gather."""
    import torch.distributed as td
    dev = torch.device("cuda", torch.cuda.current_device()) if td.get_backend(group) == "nccl" else torch.device("cpu")
    mine = torch.tensor([int(local_rows), int(local_partitions)], dtype=torch.int64, device=dev)
    allv = [torch.zeros_like(mine) for _ in range(td.get_world_size(group))]
    td.all_gather(allv, mine, group=group)
    rows = [int(v[0]) for v in allv]
    parts = [int(v[1]) for v in allv]
    r = td.get_rank(group)
    return sum(rows), sum(parts[:r]), sum(parts)
