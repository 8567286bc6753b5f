"""Config 5 alone (500 trees x depth 12 scored over synthetic windows): python profiles/config5_predict.py [rows]
Prints bench.py's `config5` object (batch sweep, rows/s, tree-rows/s).  HAR_PREDICT_SHAPE=RK picks the rows x trees
interleave of the binned predict kernel for A/B runs; HAR_PREDICT_BINS=0 the f64 walk."""
import json
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "activity-recognition-using-apache-spark_b200")); sys.path.insert(0, ROOT)
import torch
import bench
from har_b200 import synth
from har_b200.window import WisdmWindowFeaturizer

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
n = 100_000_000
stream = synth.synth_stream(n, first_sample=0, per_user=n // 1000, seed=2018)
args = types.SimpleNamespace(config5_rows=rows)
out = bench.config5_predict(args, WisdmWindowFeaturizer(), stream, None)
out["shape"] = os.environ.get("HAR_PREDICT_SHAPE", "default")
print(json.dumps(out))
