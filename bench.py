#!/usr/bin/env python
"""bench.py — BASELINE.json's metric on its config 3 shape, one process per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A step = one pass of the hot path over one batch of synthetic input: the WISDM window transform over a
tri-axial stream (per GPU: --samples samples, 20 Hz, 1000 users, 6 classes), then RandomForest(100 trees,
depth 10, maxBins 32, subsamplingRate 1, featureSubsetStrategy all) fit on the resulting windows, then
batched predict over them.  Weak scaling: every rank holds its own --samples-sized shard of users; the only
exchange is the all-reduce of node-group histograms.

`value` = samples of all ranks / device time of a step with inputs resident in HBM (CUDA events, max over
ranks); `e2e` = the same through the pyspark-shaped public API from pinned HOST columns (H2D of the stream,
D2H of features / predictions inside the timed region).  `--impl reference` times the reference's CPU
implementation of the path: Spark itself cannot run here (no JVM), so this is the C/OpenMP oracle port of
MLlib's algorithm on all host cores, on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "activity-recognition-using-apache-spark_b200")
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "accel samples/s (window features) + RF fit time"
UNIT = "samples/s"
WINDOW = 200
F = 43
BYTES_PER_SAMPLE = 41.78          # 40 B read (i32 user, i32 activity, i64 ts, 3 x f64) + 356/200 written


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--samples", type=int, default=100_000_000, help="stream samples per GPU (config 3: 1e8)")
    ap.add_argument("--users", type=int, default=1000, help="users per GPU shard")
    ap.add_argument("--trees", type=int, default=100)
    ap.add_argument("--depth", type=int, default=10)
    ap.add_argument("--partitions", type=int, default=1024, help="Spark partitions per GPU shard (bagging streams)")
    ap.add_argument("--cpu-sample", type=int, default=30_000_000, help="stream samples of the bounded CPU baseline")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------------
def cpu_pass(n_samples, trees, depth, per_user, seed=2018):
    """One pass of the path on the host cores with the oracle; returns (seconds, windows)."""
    from oracle import oracle as orc
    import numpy as np
    u, a, ts, x, y, z = orc.synth_stream(n_samples, 0, per_user, seed)
    t0 = time.perf_counter()
    ws = orc.window_segments(u, a, ts, L=WINDOW)
    feats = orc.window_features(x, y, z, ts, ws, L=WINDOW)
    labels = a[ws].astype(np.float64)
    parts = max(1, len(ws) // 500)
    po = [(len(ws) * k) // parts for k in range(parts + 1)]
    fo = orc.fit_forest(feats, labels, num_classes=6, num_trees=trees, max_depth=depth, max_bins=32,
                        feature_subset="all", seed=seed, part_off=po)
    orc.predict(fo, feats, is_forest=True)
    return time.perf_counter() - t0, len(ws)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import oracle as orc
    cores = orc.num_threads()
    n = args.cpu_sample
    per_user = max(n // max(args.users // 10, 1), 20000)
    for _ in range(args.warmup):
        cpu_pass(min(n, 1_000_000), args.trees, args.depth, per_user)
    times = []
    for _ in range(args.steps):
        dt, nw = cpu_pass(n, args.trees, args.depth, per_user)
        times.append(dt)
    total = sum(times)
    value = n * args.steps / total
    sample = ("%d samples (%d windows) per step of the config-3 stream shape: window transform + RF(%d trees, depth %d, "
              "all features) fit + predict; C/OpenMP oracle port of MLlib 2.4 (Spark/JVM unavailable on this box)"
              % (n, nw, args.trees, args.depth))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, args.gpus),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "spark": "unavailable on this box (no JVM / pyspark 2.4)"}
    print(json.dumps(line))
    return 0


def workload_config(args, world):
    return {"workload": "config 3: synthetic tri-axial stream (20 Hz, %d users/GPU, 6 classes) -> 10 s window features -> "
                        "RandomForest(%d trees, maxDepth %d, maxBins 32, subsamplingRate 1, featureSubsetStrategy all) fit + "
                        "predict" % (args.users, args.trees, args.depth),
            "samples_per_gpu": args.samples, "global_samples": args.samples * world, "window": WINDOW, "features": F,
            "trees": args.trees, "max_depth": args.depth, "partitions_per_gpu": args.partitions,
            "parallelism": "row-sharded x%d, int32 histogram all-reduce" % world,
            "l2": "inputs (%.1f GB/GPU) exceed L2; no flush needed" % (args.samples * 40 / 1e9)}


class ClockSampler:
    """SM clock / throttle reasons sampled DURING the timed region (B200_PROFILING.md clocks line).  Uses NVML
    in-process (cheap queries from a thread; a looping `nvidia-smi -lms` process was seen to stall CUDA API calls
    for tens of ms); falls back to one-shot nvidia-smi calls if pynvml is missing."""
    REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def __init__(self, gpu_index, period_s=0.1):
        self.idx, self.period, self.rows, self._stop, self._thr = gpu_index, period_s, [], threading.Event(), None
        self._nvml = None

    def _physical_index(self):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            ids = [v for v in vis.split(",") if v.strip() != ""]
            if self.idx < len(ids) and ids[self.idx].strip().isdigit():
                return int(ids[self.idx])
        return self.idx

    def prepare(self):
        """NVML init + one throw-away query, done before the warm-up steps (the first queries are slow and take
        driver locks; they must not land inside the timed region)."""
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nvml = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(self._physical_index())
            self._sample()
            self.rows.clear()
        except Exception:
            self._nvml = None

    def start(self):
        self._thr = threading.Thread(target=self._run, daemon=True)
        self._thr.start()

    def sample_now(self):
        t0 = time.perf_counter()
        try:
            self._sample()
        except Exception:
            pass
        self.query_ms = getattr(self, "query_ms", []) + [round(1e3 * (time.perf_counter() - t0), 3)]

    def _sample(self):
        if self._nvml is not None:
            n = self._nvml
            sm = n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM)
            mx = n.nvmlDeviceGetMaxClockInfo(self._h, n.NVML_CLOCK_SM)
            try:
                mask = n.nvmlDeviceGetCurrentClocksEventReasons(self._h)
            except Exception:
                mask = n.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
            self.rows.append((float(sm), float(mx), int(mask)))
        else:
            q = "clocks.sm,clocks.max.sm,clocks_event_reasons.active"
            out = subprocess.run(["nvidia-smi", "--query-gpu=" + q, "--format=csv,noheader,nounits", "-i", str(self._physical_index())],
                                 stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=10).stdout.strip().split(",")
            self.rows.append((float(out[0]), float(out[1]), int(out[2].strip(), 16)))

    def _run(self):
        while not self._stop.is_set():
            try:
                self._sample()
            except Exception:
                pass
            self._stop.wait(self.period if self._nvml is not None else 1.0)

    def stop(self):
        self._stop.set()
        if self._thr:
            self._thr.join(timeout=5)
        sm = sorted(r[0] for r in self.rows)
        reasons = set()
        for _, _, mask in self.rows:
            for name, bit in self.REASONS:
                if mask & bit:
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max((r[1] for r in self.rows), default=None),
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvml" if self._nvml is not None else "nvidia-smi",
                "query_ms": getattr(self, "query_ms", []), "sampled": "between steps, inside the timed region"}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return json.load(fh), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic():
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            return json.load(fh)
    except Exception:
        return {}


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as td
    from har_b200 import core, synth
    from har_b200.dist import Dist
    from har_b200.window import WisdmWindowFeaturizer
    from pyspark.ml.classification import RandomForestClassifier
    from pyspark.sql import DataFrame

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        td.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = args.samples
    per_user = max(n // args.users, 1)
    user, act, ts, x, y, z = synth.synth_stream(n, first_sample=rank * n, per_user=per_user, seed=2018)
    torch.cuda.synchronize()
    feat = WisdmWindowFeaturizer()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.prepare()

    def device_step(record=None):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        ev[0].record()
        feats, wu, wa, ws = feat.transform_device(user, act, ts, x, y, z)
        ev[1].record()
        nw = feats.shape[0]
        labels = wa.to(torch.float64)
        P = args.partitions
        po = [(nw * k) // P for k in range(P + 1)]
        dist = None
        if world > 1:
            cnt = torch.tensor([nw], dtype=torch.int64, device="cuda")
            td.all_reduce(cnt)
            dist = Dist(int(cnt.item()), rank * P, world * P)
        core.comm_stats(reset=True)
        forest = core.fit_forest(feats, labels, num_classes=6, num_trees=args.trees, max_depth=args.depth, max_bins=32,
                                 feature_subset="all", seed=2018, part_off=po, dist=dist)
        comm = core.comm_stats()
        ev[2].record()
        raw, prob, pred = forest.predict(feats, is_forest=True)
        ev[3].record()
        if record is not None:
            record.append((ev, nw, forest.total_nodes, forest, forest.num_batches,
                           comm))
        return forest, pred, labels

    def barrier():
        if world > 1:
            td.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        forest, pred, labels = device_step()
    barrier()
    train_acc = float((pred == labels).double().mean().item())
    core.profile(enable=True, reset=True)
    core.launch_count(reset=True)
    if rank == 0:
        pass                       # clocks are sampled inside the loop below
    rec = []
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    e0.record()
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    # with several ranks only after the last step: a query in the middle would make the other ranks wait for rank 0
    sample_at = ({args.steps // 2, args.steps - 1} if world == 1 else {args.steps - 1}) if rank == 0 else set()
    query_spans = []
    for i in range(args.steps):
        starts[i].record()
        device_step(rec)
        marks[i + 1].record()
        if i in sample_at:
            # NVML queries from a concurrent thread were seen to stall CUDA calls that create resources (one step of
            # 260 ms instead of 22); the clocks are therefore read here, inside the timed region, right after a step
            # (the GPU has not left its load state).  A query takes 1.5-70 ms of host time with the GPU idle; that
            # span is bracketed by its own events and subtracted from the timed total (reported in "clocks").
            qa, qb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            qa.record()
            clocks.sample_now()
            qb.record()
            query_spans.append((qa, qb))
    e1.record()
    barrier()
    launches = core.launch_count()
    prof = core.profile(enable=False)
    clk = clocks.stop() if rank == 0 else None
    excluded_ms = sum(a.elapsed_time(b) for a, b in query_spans)
    ms_total = e0.elapsed_time(e1) - excluded_ms
    if clk is not None:
        clk["query_ms_excluded_from_timing"] = round(excluded_ms, 3)
    t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
    if world > 1:
        td.all_reduce(t, op=td.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    stage = {"window_ms": 0.0, "fit_ms": 0.0, "predict_ms": 0.0}
    for ev, *_ in rec:
        stage["window_ms"] += ev[0].elapsed_time(ev[1]) / args.steps
        stage["fit_ms"] += ev[1].elapsed_time(ev[2]) / args.steps
        stage["predict_ms"] += ev[2].elapsed_time(ev[3]) / args.steps
    nw = rec[-1][1]
    value = n * world / (ms_step * 1e-3)

    # ---- end to end through the pyspark-shaped API from pinned host columns -------------------------
    e2e = None
    if not args.no_e2e:
        host = {}
        for name, tns in (("user", user), ("activity", act), ("timestamp", ts), ("x", x), ("y", y), ("z", z)):
            h = torch.empty(tns.shape, dtype=tns.dtype, pin_memory=True)
            h.copy_(tns)
            host[name] = h.numpy()
        types = {"user": "int", "activity": "int", "timestamp": "bigint", "x": "double", "y": "double", "z": "double"}
        raw_df = DataFrame(host, types)
        fz = WisdmWindowFeaturizer(outputCol="features", labelCol="label")
        rf = RandomForestClassifier(featuresCol="features", labelCol="label", numTrees=args.trees, maxDepth=args.depth,
                                    maxBins=32, featureSubsetStrategy="all", seed=2018)

        def host_step():
            windows = fz.transform(raw_df).repartition(args.partitions)
            model = rf.fit(windows)
            out = model.transform(windows)
            return out._cols["prediction"]

        host_step()                                    # warm-up (allocator, pinned staging)
        barrier()
        core.COUNTERS["h2d_bytes"] = core.COUNTERS["d2h_bytes"] = 0
        k_e2e = max(1, min(args.steps, 2))
        t0 = time.perf_counter()
        e0.record()
        for _ in range(k_e2e):
            host_step()
        e1.record()
        barrier()
        wall = (time.perf_counter() - t0) / k_e2e
        tt = torch.tensor([wall], dtype=torch.float64, device="cuda")
        if world > 1:
            td.all_reduce(tt, op=td.ReduceOp.MAX)
        e2e = {"value": n * world / float(tt.item()), "unit": UNIT,
               "h2d_bytes_per_step": core.COUNTERS["h2d_bytes"] // k_e2e, "d2h_bytes_per_step": core.COUNTERS["d2h_bytes"] // k_e2e,
               "ms_per_step": 1e3 * float(tt.item()), "api": "WisdmWindowFeaturizer.transform -> RandomForestClassifier.fit -> "
               "model.transform on host (pinned) columns", "steps": k_e2e}
        del raw_df, host

    # ---- CPU baseline (rank 0, N = 1 only) ------------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        from oracle import oracle as orc
        cs = args.cpu_sample
        dt, cw = cpu_pass(cs, args.trees, args.depth, max(cs // max(args.users // 10, 1), 20000))
        cpu = {"value": cs / dt, "unit": UNIT, "cores": orc.num_threads(), "kind": "port",
               "sample": "%d samples (%d windows) of the same stream shape, same RF settings, one pass (%.1f s); C/OpenMP oracle "
                         "port of MLlib 2.4 — Spark local[N] is unavailable on this box (no JVM)" % (cs, cw, dt)}

    if rank != 0:
        if world > 1:
            td.destroy_process_group()
        return 0

    # ---- roofline of the dominant kernel (live CUDA-event durations from inside the library) ----------
    peaks, peak_src = measured_peaks()
    peak = float(peaks.get("hbm_gbs", 6650.0))
    traffic = ncu_traffic()
    kern = {}
    for name, (ms, ln, units) in prof.items():
        if ln:
            kern[name] = {"ms_per_step": ms / args.steps, "launches_per_step": ln / args.steps, "units_per_step": units / args.steps}
    alg_bytes = {   # algorithmic bytes per unit, DESIGN.md §kernels
        "node_hist": F + 6,                 # per (row, tree) visit: 43 bin bytes + label + weight + 4 B row index
        "window_features": 24 + 356.0 / WINDOW,   # per sample: x, y, z f64 read + 43 f64 / window written
        "seg_tiles": 8,                     # per sample: user + activity i32
        "predict": F * 8 + 104,             # per row: 43 f64 read + raw/prob/pred written
        "part_count": 4 + 1, "part_scatter": 4 + 1 + 4,
        "bin_features": F * 8 + F,
        "bag_weights": 1,                   # per (row, tree): one weight byte written
        "best_split": (F * 32 * 6 + 6) * 4 + 456}   # per node: its counters read once + one result record written
    limiter = {
        "node_hist": "shared-memory atomics (red.shared.add.u32), not HBM",
        "bag_weights": "one serial Well19937c stream per Spark partition (latency chain of the generator), not HBM; "
                       "runs on a side stream under split finding and binning",
        "best_split": "fp64 gini evaluation of every (feature, threshold) candidate (issue-bound), not HBM",
        "window_features": "issue/latency-bound (ncu: no pipe or memory unit above 60 %), see profiles/"}
    dominant = max(kern, key=lambda k: kern[k]["ms_per_step"]) if kern else None
    roofs = {}
    for name, k in kern.items():
        if name in alg_bytes and k["ms_per_step"] > 0:
            ach = alg_bytes[name] * k["units_per_step"] / (k["ms_per_step"] * 1e-3) / 1e9
            roofs[name] = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                           "traffic": traffic.get(name), "ms_per_step": k["ms_per_step"],
                           "launches_per_step": k["launches_per_step"],
                           "avg_launch_ms": k["ms_per_step"] / k["launches_per_step"],
                           "algorithmic_bytes_per_launch": alg_bytes[name] * k["units_per_step"] / k["launches_per_step"]}
    for name, k in kern.items():               # kernels without a byte model still report their time
        roofs.setdefault(name, {"ms_per_step": k["ms_per_step"], "launches_per_step": k["launches_per_step"]})
    roofline = dict(roofs.get(dominant, {}))
    roofline["kernel"] = dominant
    roofline["peak_source"] = peak_src
    if dominant in limiter:
        roofline["limiter"] = limiter[dominant]
    if dominant == "node_hist":
        k = kern["node_hist"]
        roofline["smem_atomics_per_s"] = k["units_per_step"] * F / (k["ms_per_step"] * 1e-3)
        try:       # measured micro-benchmark peaks of red.shared.add.u32 on this GPU (same tile size as the kernel)
            pk_free, pk_rand = core.smem_atomic_peak(0), core.smem_atomic_peak(1)
            roofline["smem_atomic_peak_conflict_free_per_s"] = pk_free
            roofline["smem_atomic_peak_random_per_s"] = pk_rand
            roofline["smem_atomic_frac_of_random_peak"] = roofline["smem_atomics_per_s"] / pk_rand
        except Exception as e:
            roofline["smem_atomic_peak_error"] = repr(e)
    win_ms = stage["window_ms"]
    window_roof = {"bound": "hbm", "achieved": BYTES_PER_SAMPLE * n / (win_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                   "bytes_per_sample": BYTES_PER_SAMPLE, "stage_ms": win_ms}
    window_roof["frac"] = window_roof["achieved"] / peak

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": workload_config(args, world),
            "stages": {"window_features_samples_per_s": n / (win_ms * 1e-3), "window_ms": win_ms, "rf_fit_s": stage["fit_ms"] * 1e-3,
                       "predict_ms": stage["predict_ms"], "windows_per_gpu": nw, "forest_nodes": rec[-1][2],
                       "node_groups": rec[-1][3].num_groups, "device_batches": rec[-1][4], "allreduce_calls": rec[-1][5][0], "allreduce_via": "library NCCL communicator" if rec[-1][5][2] else "caller's callback (torch.distributed)",
                       "allreduce_bytes": rec[-1][5][1], "train_accuracy": train_acc},
            "roofline": roofline, "window_stage_roofline": window_roof, "kernels": roofs,
            "step_ms": [starts[i].elapsed_time(marks[i + 1]) for i in range(args.steps)],
            "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clk}
    print(json.dumps(line))
    if world > 1:
        td.destroy_process_group()
    return 0


def main():
    args = parse()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3            # timing rule: at least three warm-up steps
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
