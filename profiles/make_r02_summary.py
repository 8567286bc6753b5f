"""Builds profiles/r02_summary.md and profiles/traffic.json from the files one run of profiles/r02_profile_commands.sh leaves
in gpurun_out/ (launch list, three ncu --set full reports, the plain bench line of the same command).

    python profiles/make_r02_summary.py
"""
import csv
import io
import json
import os
import subprocess
import sys
from contextlib import redirect_stdout

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "gpurun_out")
sys.path.insert(0, os.path.join(ROOT, "profiles"))
import summarize_ncu


def table(rep, *subs):
    buf = io.StringIO()
    argv = sys.argv
    sys.argv = ["summarize_ncu.py", os.path.join(G, rep)] + list(subs)
    with redirect_stdout(buf):
        summarize_ncu.main()
    sys.argv = argv
    return buf.getvalue()


def raw_rows(rep):
    out = subprocess.run(["ncu", "-i", os.path.join(G, rep), "--page", "raw", "--csv"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    idx = {h: i for i, h in enumerate(rows[0])}
    return idx, rows[2:]


def launch_list():
    rows = list(csv.reader(open(os.path.join(G, "r02_launches.csv"))))
    for i, r in enumerate(rows):
        if "Kernel Name" in r:
            hdr, start = r, i
            break
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = {}
    for r in rows[start + 1:]:
        if len(r) <= vi:
            continue
        name = r[ki].split("(")[0].replace("void ", "").replace("har::", "").replace("<unnamed>::", "")
        if name.startswith("at::") or name.startswith("vectorized") or "elementwise" in name or "nccl" in name.lower():
            name = "(torch kernels: synthetic stream, label cast, signature sums)"
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += float(r[vi].replace(",", "")) / 1e6
    tot = sum(v[1] for v in agg.values())
    lines = ["| kernel | launches | total ms | share |", "|---|---|---|---|"]
    for name, (c, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        if ms / tot < 0.002:
            continue
        lines.append("| %s | %d | %.3f | %.1f%% |" % (name, c, ms, 100 * ms / tot))
    return "\n".join(lines), tot


def main():
    plain = json.loads(open(os.path.join(G, "r02_prof_plain.json")).read().strip().splitlines()[-1])
    k = {n: v["ms_per_step"] for n, v in plain["kernels"].items()}
    ll, tot = launch_list()
    # traffic per launch for the bench line
    traffic = {}
    for rep, mapping in (("r02_full_window_predict.ncu-rep", {"k_window_features_v4": "window_features", "k_seg_tiles": "seg_tiles", "k_predict_walk": "predict"}),
                         ("r02_full_fit.ncu-rep", {"k_node_hist_lanes": "node_hist", "k_best_split": "best_split", "k_part_count": "part_count", "k_part_scatter": "part_scatter"})):
        idx, body = raw_rows(rep)
        acc = {}
        for r in body:
            for sub, key in mapping.items():
                if sub in r[idx["Kernel Name"]]:
                    def b(col):
                        v = float(r[idx[col]].replace(",", ""))
                        u = body and rows_unit[col]
                        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
                    rows_unit = units(rep)
                    acc.setdefault(key, []).append(b("dram__bytes_read.sum") + b("dram__bytes_write.sum"))
        for key, vals in acc.items():
            traffic[key] = sum(vals) / len(vals)
    traffic["_note"] = ("dram__bytes_read.sum + dram__bytes_write.sum per launch from ncu --set full (profiles/r02_summary.md, round 2); tree "
                        "kernels = mean over the captured launches of one mid-depth level, window / seg / predict = single launches of the config-3 step")
    old = {}
    try:
        old = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    except Exception:
        pass
    for key in ("bag_weights",):
        if key in old and key not in traffic:
            traffic[key] = old[key]
    json.dump(traffic, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)

    md = []
    md.append("# Round-2 profiles (B200, sm_100a, driver 580, `--clock-control none`)\n")
    md.append("Command profiled (`profiles/r02_profile_commands.sh`): `python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-wisdm --no-config5 --no-config4`\n"
              "(config 3: 1e8 samples -> 484,573 windows -> RF 100 x depth 10 -> predict).  It was run plain first\n"
              "(`r02_bench_profiled_command.json`: %.2f ms per step: window %.3f ms, fit %.2f ms, predict %.3f ms), then under ncu.\n"
              "The tables are written by `profiles/make_r02_summary.py` from the reports of that run.\n"
              % (plain["ms_per_step"], plain["stages"]["window_ms"], plain["stages"]["rf_fit_s"] * 1e3, plain["stages"]["predict_ms"]))
    md.append("## Launch list (`r02_launches.csv`, ncu `--metrics gpu__time_duration.sum`, 1200 launches = 5 steps + setup)\n")
    md.append("Per-launch times are cold-cache and serialised: compare SHARES with the live CUDA-event numbers of the bench line\n"
              "(node_hist %.2f, best_split %.2f, bag %.2f, part %.2f, find_splits %.2f, predict %.2f, window_features %.2f, seg_tiles %.2f ms\n"
              "per step; bagging, split finding and binning overlap on a side stream and the two tree sets overlap each other).\n"
              % (k["node_hist"], k["best_split"], k["bag_weights"], k["part_count"] + k["part_scatter"], k["find_splits"], k["predict"],
                 k["window_features"], k["seg_tiles"]))
    md.append(ll + "\n")
    md.append("## `ncu --set full`: window stage and predict (`gpurun_out/r02_full_window_predict.ncu-rep`, one launch each of a timed step)\n")
    md.append(table("r02_full_window_predict.ncu-rep"))
    md.append(open(os.path.join(ROOT, "profiles", "r02_notes_window.md")).read())
    md.append("## `ncu --set full`: fit kernels of a mid-depth level (`gpurun_out/r02_full_fit.ncu-rep`)\n")
    md.append(table("r02_full_fit.ncu-rep"))
    md.append("## `ncu --set full`: the histogram kernel at the root level (`gpurun_out/r02_full_hist_root.ncu-rep`)\n")
    md.append(table("r02_full_hist_root.ncu-rep"))
    md.append(open(os.path.join(ROOT, "profiles", "r02_notes_fit.md")).read())
    open(os.path.join(ROOT, "profiles", "r02_summary.md"), "w").write("\n".join(md))
    print("wrote profiles/r02_summary.md, profiles/traffic.json; launch list total %.1f ms" % tot)


_units_cache = {}


def units(rep):
    if rep not in _units_cache:
        out = subprocess.run(["ncu", "-i", os.path.join(G, rep), "--page", "raw", "--csv"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
        rows = list(csv.reader(out.splitlines()))
        _units_cache[rep] = dict(zip(rows[0], rows[1]))
    return _units_cache[rep]


if __name__ == "__main__":
    main()
