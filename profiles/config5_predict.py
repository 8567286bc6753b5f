"""batched predict of a 500-tree depth-12 forest over 1e8 windows x 43 f64, batch-size sweep.

The forest is trained with the library itself on 1e6 windows of the synthetic stream (RandomForest 500 x depth 12,
featureSubsetStrategy sqrt), then 1e8 rows are scored as batches of the sizes below (the 1e6 distinct windows tiled;
scoring does not depend on row values repeating).  Prints one JSON line; run on one B200:
    python profiles/config5_predict.py [--trees 500 --depth 12 --rows 100000000]
"""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "activity-recognition-using-apache-spark_b200")); sys.path.insert(0, ROOT)
import torch
from har_b200 import core, synth
from har_b200.window import WisdmWindowFeaturizer

ap = argparse.ArgumentParser()
ap.add_argument("--trees", type=int, default=500)
ap.add_argument("--depth", type=int, default=12)
ap.add_argument("--rows", type=int, default=100_000_000)
ap.add_argument("--train-samples", type=int, default=200_000_000)
args = ap.parse_args()

user, act, ts, x, y, z = synth.synth_stream(args.train_samples, 0, args.train_samples // 1000, 2018)
feats, wu, wa, ws = WisdmWindowFeaturizer().transform_device(user, act, ts, x, y, z)
del user, act, ts, x, y, z
nw = feats.shape[0]
labels = wa.to(torch.float64)
P = 1024
t0 = time.perf_counter()
forest = core.fit_forest(feats, labels, num_classes=6, num_trees=args.trees, max_depth=args.depth, max_bins=32,
                         feature_subset="sqrt", seed=2018, part_off=[(nw * k) // P for k in range(P + 1)])
torch.cuda.synchronize()
fit_s = time.perf_counter() - t0
out = {"config": "5: predict, forest %d trees depth %d (sqrt features), %d windows x 43 f64" % (args.trees, args.depth, args.rows),
       "train_windows": nw, "fit_s": fit_s, "forest_nodes": forest.total_nodes, "batches": []}
for batch in (100_000, 1_000_000, 10_000_000):
    reps = (batch + nw - 1) // nw
    X = feats.repeat(reps, 1)[:batch].contiguous() if reps > 1 else feats[:batch].contiguous()
    nb = max(1, min(args.rows // batch, 40))              # bounded number of timed batches, scaled to args.rows below
    forest.predict(X, is_forest=True, want=("pred",))     # warm-up
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(nb):
        raw, prob, pred = forest.predict(X, is_forest=True)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / nb
    out["batches"].append({"batch_rows": batch, "ms_per_batch": ms, "rows_per_s": batch / (ms * 1e-3),
                           "tree_rows_per_s": batch * args.trees / (ms * 1e-3),
                           "seconds_for_all_rows": args.rows / (batch / (ms * 1e-3)), "timed_batches": nb})
    del X
print(json.dumps(out))
