"""One rank of tests/test_gpu_configs.py::test_two_ranks_on_one_gpu_fit_the_single_rank_forest.

    RANK=r WORLD_SIZE=2 MASTER_ADDR=127.0.0.1 MASTER_PORT=p HAR_NCCL=0 python _two_rank_worker.py <all|sqrt> <tmpdir>

Both ranks use cuda:0.  The collective is torch.distributed on gloo, staged through the host by har_b200.dist.Dist:
no kernel of one rank ever waits for the other rank's kernels (the host waits), so sharing the GPU is safe."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "activity-recognition-using-apache-spark_b200")); sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
import torch.distributed as td

from conftest import synth_tree_data
from har_b200 import core, dist as hd

subset, tmp = sys.argv[1], sys.argv[2]
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
assert os.environ.get("HAR_NCCL") == "0"
torch.cuda.set_device(0)
td.init_process_group("gloo", rank=rank, world_size=world)

n, F, Cn, P = 24000, 12, 4, 10                    # P Spark partitions over all rows, uneven on purpose
X, y, arity = synth_tree_data(n, F=F, C=Cn, seed=41, n_cat_unordered=1, n_cat_ordered=1)
gpo = np.array([0, 1000, 1000, 5200, 7000, 11000, 13500, 13501, 19000, 22000, n], dtype=np.int64)
r0, r1, lpo, p0 = hd.shard_rows(gpo, rank, world)
grow, poff, ptot = hd.exchange_shard_layout(r1 - r0, len(lpo) - 1)
assert (grow, ptot, poff) == (n, P, p0)
kw = dict(arity=arity, num_classes=Cn, num_trees=9, max_depth=8, max_bins=32, feature_subset=subset, seed=7)
d = hd.Dist(grow, poff, ptot)
fo = core.fit_forest(X[r0:r1], y[r0:r1], part_off=lpo, dist=d, **kw)
via, why = d.transport()
assert via == 0 and "HAR_NCCL=0" in why, (via, why)
assert d.calls > 5
arrays = {k: np.asarray(getattr(fo, k)) for k in core.Forest.FIELDS}
np.savez(os.path.join(tmp, "forest_rank%d.npz" % rank), **arrays)
td.barrier()
if rank == 0:
    other = np.load(os.path.join(tmp, "forest_rank1.npz"))
    ref = core.fit_forest(X, y, part_off=gpo, **kw)                 # one rank, all rows, same Spark partitioning
    for k in core.Forest.FIELDS:
        a, b, c = arrays[k], other[k], np.asarray(getattr(ref, k))
        assert a.shape == b.shape == c.shape, k
        assert np.array_equal(a.view(np.uint8), b.view(np.uint8)), "ranks disagree on %s" % k
        assert np.array_equal(a.view(np.uint8), c.view(np.uint8)), "%s differs between 2 ranks and 1 rank" % k
    assert ref.total_nodes > 500
    print("%d nodes, %d callback all-reduces: identical to the single-rank forest" % (ref.total_nodes, d.calls))
td.barrier()
hd.release()
td.destroy_process_group()
print("rank %d ok" % rank)
