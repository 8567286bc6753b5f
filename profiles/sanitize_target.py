"""Small pass over every kernel family for compute-sanitizer (memcheck / racecheck / initcheck / synccheck, one tool
per run):  compute-sanitizer --tool racecheck python profiles/sanitize_target.py

Window transform (fused call, the two separate calls, the static-scratch segmentation twice in a row, a second window
length), raw-text ingest, split finding, binning, bagging (Poisson pipeline and Bernoulli), the device-driven fit
(all features, sibling subtraction, multi-batch levels), the host-driven fit (feature subsets), both predict modes."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "activity-recognition-using-apache-spark_b200")); sys.path.insert(0, ROOT)
import numpy as np
import torch
from har_b200 import _native, core, synth

n = 120_000
u, a, ts, x, y, z = synth.synth_stream(n, per_user=30_000, seed=7)
f, wu, wa, ws = core.window_transform(u, a, ts, x, y, z)
ws2, _, _ = core.window_segments(u, a, ts)
f2 = core.window_features(x, y, z, ts, ws2)
assert torch.equal(ws, ws2) and torch.equal(torch.nan_to_num(f), torch.nan_to_num(f2))
core.window_features(x, y, z, ts, core.window_segments(u, a, ts, window_len=64)[0], window_len=64)
lib = _native.lib()
for k in range(2):
    cap = n // 200
    o = torch.empty(cap, dtype=torch.int64, device="cuda"); ou = torch.empty(cap, dtype=torch.int32, device="cuda"); oa = torch.empty_like(ou)
    cnt = C.c_int64()
    rc = lib.har_window_segments_static(C.c_void_p(u.data_ptr()), C.c_void_p(a.data_ptr()), C.c_void_p(ts.data_ptr()), n - 1000 * k, 200, 0,
                                        C.c_void_p(o.data_ptr()), C.c_void_p(ou.data_ptr()), C.c_void_p(oa.data_ptr()), C.byref(cnt),
                                        C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
text = b"".join(b"%d,Walking,%d,%.2f,%.2f,%.2f;\n" % (i % 5, 1000 * i, 0.1 * i, -0.2 * i, 9.8) for i in range(3000))
core.parse_wisdm_raw(text)
feats = torch.nan_to_num(f, nan=0.0)
labels = wa.to(torch.float64)
nw = feats.shape[0]
po = [(nw * k) // 7 for k in range(8)]
fo = core.fit_forest(feats, labels, num_classes=6, num_trees=6, max_depth=7, feature_subset="all", seed=3, part_off=po)
fo.predict(feats, is_forest=True); fo.predict(feats, is_forest=False)
os.environ["HAR_HIST_BUDGET_WORDS"] = "60000"
core.fit_forest(feats, labels, num_classes=6, num_trees=4, max_depth=6, feature_subset="all", seed=3, part_off=po)
del os.environ["HAR_HIST_BUDGET_WORDS"]
fs = core.fit_forest(feats, labels, num_classes=6, num_trees=5, max_depth=5, feature_subset="sqrt", seed=3, part_off=po, subsampling_rate=0.7)
fs.predict(feats, is_forest=True)
core.fit_forest(feats, labels, num_classes=6, num_trees=1, max_depth=5, subsampling_rate=0.5, seed=3, part_off=po)
rng = np.random.RandomState(0)
X = rng.randn(3000, 6); X[:, 0] = rng.randint(0, 4, 3000); X[:, 1] = rng.randint(0, 9, 3000)
yy = (X[:, 2] > 0).astype(np.float64) + (X[:, 0] > 1)
fc = core.fit_forest(X, yy, arity=[4, 9, 0, 0, 0, 0], num_classes=3, num_trees=3, max_depth=5, feature_subset="all", seed=1)
fc.predict(X)
len(fo.feature); len(fs.feature)
torch.cuda.synchronize()
print("sanitize target ok: %d windows, %d + %d + %d nodes" % (nw, fo.total_nodes, fs.total_nodes, fc.total_nodes))
