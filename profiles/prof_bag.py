"""Bagging kernel only (for ncu): python profiles/prof_bag.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "activity-recognition-using-apache-spark_b200")); sys.path.insert(0, ROOT)
import torch
from har_b200 import core
n, T, P = 484573, 100, 1024
po = [(n * k) // P for k in range(P + 1)]
for _ in range(2):
    w = core.bag_weights(n, T, 1.0, seed=2018, part_off=po)
torch.cuda.synchronize()
print("mean weight", float(w.float().mean()))
