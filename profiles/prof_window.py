"""Window stage only (for timing and ncu): python profiles/prof_window.py [samples] [reps]

Prints one JSON line: CUDA-event times of the segmentation pass, the feature kernel and the fused
har_window_transform call (segments + features + the one host wait), with their fractions of the measured HBM peak
(MEASURED_PEAKS.json) at the algorithmic byte counts of DESIGN.md: 8 B/sample for segmentation, 24 B/sample + 344
B/window for the features, 41.78 B/sample for the stage."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "activity-recognition-using-apache-spark_b200")); sys.path.insert(0, ROOT)
import torch
from har_b200 import core, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    peak = 6650.0
u, a, ts, x, y, z = synth.synth_stream(n, per_user=max(n // 1000, 1), seed=2018)
for _ in range(3):
    ws, wu, wa = core.window_segments(u, a, ts)
    f = core.window_features(x, y, z, ts, ws)
    core.window_transform(u, a, ts, x, y, z)
torch.cuda.synchronize()


def timed(fn):
    best, tot = 1e30, 0.0
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        best, tot = min(best, ms), tot + ms
    return best, tot / reps


nw = ws.numel()
seg = timed(lambda: core.window_segments(u, a, ts))
feat = timed(lambda: core.window_features(x, y, z, ts, ws))
fused = timed(lambda: core.window_transform(u, a, ts, x, y, z))
core.profile(enable=True, reset=True)
for _ in range(reps):
    core.window_transform(u, a, ts, x, y, z)
torch.cuda.synchronize()
prof = core.profile(enable=False)
k_seg = prof["seg_tiles"][0] / max(prof["seg_tiles"][1], 1)
k_feat = prof["window_features"][0] / max(prof["window_features"][1], 1)
gb = lambda b, ms: b / ms / 1e6
out = {"samples": n, "windows": nw, "reps": reps, "hbm_peak_gbs": peak,
       "segments_call_ms": {"best": seg[0], "mean": seg[1]}, "features_call_ms": {"best": feat[0], "mean": feat[1]},
       "transform_call_ms": {"best": fused[0], "mean": fused[1]},
       "k_seg_pass_ms": k_seg, "k_seg_pass_frac": gb(8.0 * n, k_seg) / peak,
       "k_window_features_ms": k_feat, "k_window_features_frac": gb(24.0 * n + 344.0 * nw, k_feat) / peak,
       "stage_frac_mean": gb(41.78 * n, fused[1]) / peak, "stage_frac_best": gb(41.78 * n, fused[0]) / peak}
print(json.dumps(out))
