"""Multi-GPU check (run under torchrun on N GPUs of one box; not collected by pytest):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tests/multi_gpu_check.py

Every rank windows its own shard of users, the forest is fitted collectively (row-sharded histograms,
NCCL all-reduce through the har_dist callback) and must be bit-identical on every rank and identical to the
forest a single GPU fits on all rows with the same Spark partitioning."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "activity-recognition-using-apache-spark_b200")); sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as td
from har_b200 import core, synth
from har_b200.dist import Dist, release

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
td.init_process_group("nccl", device_id=torch.device("cuda", local))
n_per, per_user, P = 4_000_000, 100_000, 16          # samples, samples per user, partitions per rank


def windows(first, n):
    u, a, ts, x, y, z = synth.synth_stream(n, first_sample=first, per_user=per_user, seed=2018)
    ws, wu, wa = core.window_segments(u, a, ts)
    return core.window_features(x, y, z, ts, ws), wa.to(torch.float64)


feats, labels = windows(rank * n_per, n_per)
nw = feats.shape[0]
counts = [torch.zeros(1, dtype=torch.int64, device="cuda") for _ in range(world)]
td.all_gather(counts, torch.tensor([nw], dtype=torch.int64, device="cuda"))
counts = [int(c.item()) for c in counts]
po = [(nw * k) // P for k in range(P + 1)]
for kw in (dict(num_trees=20, max_depth=8, feature_subset="all"), dict(num_trees=10, max_depth=6, feature_subset="sqrt")):
    dist = Dist(sum(counts), rank * P, world * P)
    fo = core.fit_forest(feats, labels, num_classes=6, max_bins=32, seed=2018, part_off=po, dist=dist, **kw)
    # identical on every rank
    sig = torch.tensor([fo.total_nodes, int(fo.feature.astype(np.int64).sum()), int(fo.threshold.view(np.int64).sum() % (1 << 40)),
                        int(fo.stats.sum())], dtype=torch.int64, device="cuda")
    sigs = [torch.zeros_like(sig) for _ in range(world)]
    td.all_gather(sigs, sig)
    assert all(bool((s == sigs[0]).all()) for s in sigs), "ranks disagree on the forest"
    # identical to a single-GPU fit on all rows (rank 0 rebuilds every shard locally)
    if rank == 0:
        allf, alll, gpo, off = [], [], [0], 0
        for r in range(world):
            f, l = windows(r * n_per, n_per)
            assert f.shape[0] == counts[r]
            allf.append(f); alll.append(l)
            gpo += [off + (counts[r] * k) // P for k in range(1, P + 1)]
            off += counts[r]
        ref = core.fit_forest(torch.cat(allf), torch.cat(alll), num_classes=6, max_bins=32, seed=2018, part_off=gpo, **kw)
        for k in core.Forest.FIELDS:
            a, b = np.asarray(getattr(ref, k)), np.asarray(getattr(fo, k))
            assert a.shape == b.shape and np.array_equal(a.view(np.uint8), b.view(np.uint8)), "field %s differs between 1 and %d GPUs" % (k, world)
        calls, nbytes, via = core.comm_stats()
        print("ok: %s -> %d nodes, %d all-reduce calls so far (%d through the callback), %.1f MB reduced, %s; identical on %d "
              "ranks and to the 1-GPU forest" % (kw, fo.total_nodes, calls, dist.calls, nbytes / 1e6,
                                                 "library NCCL communicator" if via else "callback only", world))
    td.barrier()
release()
td.destroy_process_group()
