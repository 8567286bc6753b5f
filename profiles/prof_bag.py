"""Bagging kernel only (for ncu / timing): python profiles/prof_bag.py  [HAR_BAG_KERNEL=v1 for the single-warp kernel]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "activity-recognition-using-apache-spark_b200")); sys.path.insert(0, ROOT)
import torch
from har_b200 import core
n, T = 484573, 100
for P in (1024, 256, 4096):
    po = [(n * k) // P for k in range(P + 1)]
    w = core.bag_weights(n, T, 1.0, seed=2018, part_off=po)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        w = core.bag_weights(n, T, 1.0, seed=2018, part_off=po)
    e1.record(); torch.cuda.synchronize()
    print("P=%d  %.3f ms/call  mean weight %.6f" % (P, e0.elapsed_time(e1) / 3, float(w.float().mean())))
