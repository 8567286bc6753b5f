"""ncu report -> markdown table of the metrics quoted in profiles/*_summary.md.

    python profiles/summarize_ncu.py gpurun_out/r02_full_fit.ncu-rep [kernel-substring ...]

One column per distinct (kernel, grid) launch kept (first occurrence of each kernel name unless substrings are given, in
which case every launch whose name contains one of them is listed)."""
import csv
import subprocess
import sys

METRICS = [
    ("gpu__time_duration.sum", "us"),
    ("dram__bytes_read.sum", ""),
    ("dram__bytes_write.sum", ""),
    ("launch__grid_size", ""),
    ("launch__block_size", ""),
    ("launch__registers_per_thread", ""),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "%"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "%"),
    ("smsp__inst_executed.sum", ""),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "%"),
    ("lts__t_sector_hit_rate.pct", "%"),
    ("l1tex__t_sector_hit_rate.pct", "%"),
    ("smsp__inst_executed_op_shared_atom.sum", ""),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_atom.sum", ""),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", ""),
    ("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "%"),
    ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "%"),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", ""),
    ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", ""),
    ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", ""),
    ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", ""),
    ("smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", ""),
]


def main():
    rep, subs = sys.argv[1], sys.argv[2:]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, body = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    kept, seen = [], set()
    for r in body:
        name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "").replace("har::", "")
        if subs:
            if not any(s_ in name for s_ in subs):
                continue
        elif name in seen:
            continue
        seen.add(name)
        kept.append((name + " grid " + r[idx["Grid Size"]].split(",")[0].strip("("), r))
    print("| metric | " + " | ".join(k for k, _ in kept) + " | unit |")
    print("|---|" + "---|" * (len(kept) + 1))
    for m, _ in METRICS:
        if m not in idx:
            continue
        vals = []
        for _, r in kept:
            v = r[idx[m]]
            try:
                f = float(v.replace(",", ""))
                v = ("%d" % f) if f == int(f) and abs(f) >= 100 else ("%.4g" % f)
            except ValueError:
                pass
            vals.append(v)
        print("| %s | %s | %s |" % (m, " | ".join(vals), units[idx[m]]))


if __name__ == "__main__":
    main()
