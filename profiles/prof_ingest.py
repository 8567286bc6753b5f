"""Throughput of the raw-text ingest: python profiles/prof_ingest.py  (text resident in HBM, two passes: size + parse)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "activity-recognition-using-apache-spark_b200")); sys.path.insert(0, ROOT)
import numpy as np, torch
from har_b200 import core
n = 1_000_000
rng = np.random.RandomState(1)
u = rng.randint(1, 37, n); a = rng.randint(0, 6, n); ts = 49105962326000 + np.arange(n) * 50000000
xyz = (rng.randn(n, 3) * 5).astype(np.float32)
names = core.WISDM_ACTIVITIES
block = "".join("%d,%s,%d,%s,%s,%s;\n" % (u[i], names[a[i]], ts[i], xyz[i, 0], xyz[i, 1], xyz[i, 2]) for i in range(n)).encode()
text = torch.from_numpy(np.frombuffer(block, np.uint8).copy()).cuda().repeat(20)       # 2e7 records
torch.cuda.synchronize()
core.parse_wisdm_raw(text)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3):
    cols, dropped = core.parse_wisdm_raw(text)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 3
print("%d records, %.1f MB text: %.2f ms per call (sizing pass + parse) = %.1f GB/s of text, %.1f M records/s, dropped %d"
      % (cols[0].numel(), text.numel() / 1e6, ms, text.numel() / ms / 1e6, cols[0].numel() / ms / 1e3, dropped))
