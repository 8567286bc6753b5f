"""CPU suite: the C-ABI library loads and exports every symbol include/har_b200.h declares (no compute
without a GPU), the pyspark-shaped Params surface, and the N>1 host logic on world_size-2 gloo."""
import ctypes as C
import os
import re
import socket
import subprocess
import sys

import numpy as np
import pytest

from conftest import PKG, ROOT


def _header_symbols():
    txt = open(os.path.join(ROOT, "include", "har_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(har_[a-z0-9_]+)\s*\(", txt)) - {"har_allreduce_i32_fn"})


def test_library_exports_every_declared_symbol():
    from har_b200 import _native
    _native.build()
    lib = C.CDLL(_native._SO)
    syms = _header_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), "include/har_b200.h declares %s but libhar_b200.so does not export it" % s
    # and the ctypes binding knows all of them
    assert set(syms) <= set(_native.EXPORTED)


def test_no_cpu_fallback_without_a_gpu():
    import torch
    from har_b200 import core
    from har_b200._native import HarError
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(HarError, match="no CPU fallback"):
        core.fit_forest(np.zeros((4, 2)), np.zeros(4), num_classes=2)
    with pytest.raises(HarError, match="no CPU fallback"):
        core.window_segments(np.zeros(4, np.int32), np.zeros(4, np.int32))


def test_product_package_never_imports_the_oracle():
    for base, _, files in os.walk(PKG):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(base, f), errors="ignore").read()
                assert "import oracle" not in src and "from oracle" not in src and "libhar_oracle" not in src, os.path.join(base, f)


def test_struct_layouts_match_the_header():
    from har_b200 import _native
    assert C.sizeof(_native.FitParams) == 8 * 4 + 2 * 8 + 2 * 8
    assert C.sizeof(_native.Dist) == 4 + 4 + 8 + 4 + 4 + 8 + 8 + 8


def test_estimator_params_follow_spark_2_4():
    from pyspark.ml.classification import DecisionTreeClassifier, RandomForestClassifier
    dt = DecisionTreeClassifier(featuresCol="features", labelCol="label", maxDepth=3)
    assert dt.getMaxDepth() == 3 and dt.getMaxBins() == 32 and dt.getImpurity() == "gini"
    assert dt.getMinInstancesPerNode() == 1 and dt.getMinInfoGain() == 0.0 and dt.getMaxMemoryInMB() == 256
    assert re.fullmatch(r"DecisionTreeClassifier_[0-9a-f]{20}", dt.uid)
    assert dt.isSet(dt.maxDepth) and not dt.isSet(dt.maxBins) and dt.isDefined(dt.maxBins) and dt.hasParam("seed")
    rf = RandomForestClassifier(numTrees=100, maxDepth=4, maxBins=32)
    assert rf.getNumTrees() == 100 and rf.getFeatureSubsetStrategy() == "auto" and rf.getSubsamplingRate() == 1.0
    rf2 = rf.copy({rf.maxDepth: 7})
    assert rf2.getMaxDepth() == 7 and rf.getMaxDepth() == 4 and rf2.uid == rf.uid
    assert "maxDepth: Maximum depth of the tree" in rf.explainParams()
    pm = rf.extractParamMap()
    assert pm[rf.numTrees] == 100
    with pytest.raises(TypeError):
        RandomForestClassifier(numTrees="many")


def test_feature_subset_resolution_matches_mllib():
    from har_b200.core import resolve_features_per_node as r
    assert r("auto", 3100, 1) == 3100 and r("auto", 3100, 100) == 56 and r("sqrt", 43, 5) == 7
    assert r("log2", 43, 5) == 6 and r("onethird", 43, 5) == 15 and r("10", 43, 5) == 10 and r("0.5", 43, 5) == 22
    assert r("all", 43, 5) == 43 and r("1000", 43, 5) == 43
    with pytest.raises(ValueError):
        r("most", 43, 5)


def test_shard_partitions_cover_and_balance():
    from har_b200.dist import shard_partitions, shard_rows
    for P, W in [(1024, 8), (7, 2), (3, 4), (10, 3)]:
        got = [shard_partitions(P, r, W) for r in range(W)]
        assert got[0][0] == 0 and got[-1][1] == P
        assert all(a[1] == b[0] for a, b in zip(got, got[1:]))
        sizes = [b - a for a, b in got]
        assert max(sizes) - min(sizes) <= 1
    gpo = np.array([0, 10, 25, 25, 60, 100])
    r0, r1, lpo, p0 = shard_rows(gpo, 1, 2)
    assert (r0, r1, p0) == (25, 100, 3) and lpo.tolist() == [0, 35, 75]


_WORKER = r'''
import os, sys
sys.path.insert(0, %(pkg)r); sys.path.insert(0, %(root)r)
import ctypes as C
import numpy as np, torch, torch.distributed as td
from har_b200 import _native as N
from har_b200.dist import Dist, exchange_shard_layout, shard_rows
from oracle import oracle as orc
td.init_process_group("gloo", init_method="tcp://127.0.0.1:%(port)d", rank=int(sys.argv[1]), world_size=2)
rank = td.get_rank()
rng = np.random.RandomState(0)
n, F, Cn = 6000, 5, 3
bins = rng.randint(0, 32, size=(n, F)).astype(np.uint8)
labels = rng.randint(0, Cn, size=n).astype(np.uint8)
gpo = np.array([0, 1500, 2500, 4700, 6000])
r0, r1, lpo, p0 = shard_rows(gpo, rank, 2)
w_full = orc.bag_weights(n, 2, 1.0, seed=5, part_off=gpo)
hist_off = np.arange(F + 1, dtype=np.int64) * 32 * Cn
full, _ = orc.node_histogram(bins, labels, hist_off, Cn, weights=w_full[1])
mine, _ = orc.node_histogram(bins[r0:r1], labels[r0:r1], hist_off, Cn, weights=w_full[1][r0:r1])
grow, poff, ptot = exchange_shard_layout(r1 - r0, len(lpo) - 1)
assert (grow, ptot) == (n, 4) and poff == p0, (grow, poff, ptot)
d = Dist(grow, poff, ptot, host_buffers=True)
buf = mine.astype(np.int32)
rc = d.struct.allreduce_i32(None, buf.ctypes.data_as(C.c_void_p), buf.size, None)   # the callback the library calls
assert rc == 0 and d.calls == 1 and d.bytes == 4 * buf.size
assert np.array_equal(buf.astype(np.int64), full), "sum of shard histograms != full histogram"
# all-gather through the integer all-reduce (how the findSplits sample is assembled)
vals = np.zeros(8); vals[rank * 4:(rank + 1) * 4] = np.array([1.5, -2.25, 1e-300, 3e200]) * (rank + 1)
iv = vals.view(np.int32).copy()
assert d.struct.allreduce_i32(None, iv.ctypes.data_as(C.c_void_p), iv.size, None) == 0
want = np.concatenate([np.array([1.5, -2.25, 1e-300, 3e200]), np.array([1.5, -2.25, 1e-300, 3e200]) * 2])
assert np.array_equal(iv.view(np.float64), want)
td.barrier(); td.destroy_process_group()
print("rank", rank, "ok")
'''


def test_world_size_2_gloo_allreduce_callback_sums_shard_histograms(tmp_path):
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    script = tmp_path / "worker.py"
    script.write_text(_WORKER % {"pkg": PKG, "root": ROOT, "port": port})
    procs = [subprocess.Popen([sys.executable, str(script), str(r)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(2)]
    outs = [p.communicate(timeout=180)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0 and ("rank %d ok" % r) in o, o


def test_bench_reference_arm_prints_the_contract_line():
    """`bench.py --impl reference` (the CPU arm the driver runs beside ours) needs no GPU: one JSON line with the
    metric, the cpu_baseline description and a zero-copy e2e object."""
    import json, os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                          "--cpu-sample", "1000000", "--trees", "10", "--depth", "5"], stdout=subprocess.PIPE,
                         stderr=subprocess.PIPE, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "samples/s" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"] == {"value": line["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    # the other ranks of a torchrun launch print nothing and exit 0
    env = dict(os.environ, RANK="1", WORLD_SIZE="2")
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1"], env=env,
                         stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_forest_import_rejects_malformed_models():
    """har_forest_import validates the flat arrays on the host (no GPU needed): a truncated or cyclic model must not
    reach the predict kernel."""
    from har_b200 import core
    from har_b200._native import HarError
    ok = dict(tree_off=[0, 3], feature=[0, -1, -1], left=[1, -1, -1], right=[2, -1, -1], arity=[0, 0, 0],
              threshold=[0.5, 0, 0], left_mask=[0, 0, 0], stats=np.ones((3, 2)))
    f = core.Forest.from_arrays(1, 2, 1, **ok)
    assert f.total_nodes == 3
    for field, value, msg in (("left", [2, -1, -1], "two parents"), ("right", [5, -1, -1], "child outside"),
                              ("left", [0, -1, -1], "child outside"), ("feature", [3, -1, -1], "tests feature"),
                              ("arity", [70, 0, 0], "arity"), ("tree_off", [0, 0], "empty")):
        bad = dict(ok); bad[field] = value
        with pytest.raises(HarError, match=msg):
            core.Forest.from_arrays(1, 2, 1, **bad)
    # a node that nothing points to
    bad = dict(ok, tree_off=[0, 4], feature=[0, -1, -1, -1], left=[1, -1, -1, -1], right=[2, -1, -1, -1], arity=[0] * 4,
               threshold=[0.5, 0, 0, 0], left_mask=[0] * 4, stats=np.ones((4, 2)))
    with pytest.raises(HarError, match="not reachable"):
        core.Forest.from_arrays(1, 2, 1, **bad)
