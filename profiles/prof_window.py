"""Window stage only (for ncu): python profiles/prof_window.py [samples]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "activity-recognition-using-apache-spark_b200")); sys.path.insert(0, ROOT)
import torch
from har_b200 import core, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 50_000_000
u, a, ts, x, y, z = synth.synth_stream(n, per_user=100_000, seed=2018)
for _ in range(3):
    ws, wu, wa = core.window_segments(u, a, ts)
    f = core.window_features(x, y, z, ts, ws)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); f = core.window_features(x, y, z, ts, ws); e1.record(); torch.cuda.synchronize()
print("windows", ws.numel(), "features kernel ms", e0.elapsed_time(e1), "GB/s", (24 * n + 344 * ws.numel()) / e0.elapsed_time(e1) / 1e6)
