"""Summarise ncu outputs brought back in gpurun_out/ into small text files under profiles/ (tracked).

  python tools/summarize_ncu.py launches <launches.csv> <out.md>      # per-kernel totals and shares of a launch list
  python tools/summarize_ncu.py full <report.ncu-rep> <out.md>        # key counters of an `ncu --set full` capture
"""
import collections
import csv
import re
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "l1tex__t_bytes.sum", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct",
        "smsp__average_warp_latency_per_inst_issued.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "sm__cycles_elapsed.avg.per_second"]


def short(name):
    m = re.search(r"([A-Za-z_0-9]+)(<[^>]*>)?\s*\(", name)
    return (m.group(1) + (m.group(2) or "")) if m else name[:40]


def launches(path, out):
    rows = list(csv.reader(open(path)))
    i = next(k for k, r in enumerate(rows) if "Kernel Name" in r)
    h = rows[i]
    ki, vi = h.index("Kernel Name"), h.index("Metric Value")
    agg = collections.OrderedDict()
    for r in rows[i + 1:]:
        if len(r) <= vi:
            continue
        n = short(r[ki])
        v = float(r[vi].replace(",", ""))
        a = agg.setdefault(n, [0, 0.0, 0.0])
        a[0] += 1
        a[1] += v
        a[2] = max(a[2], v)
    tot = sum(a[1] for a in agg.values())
    with open(out, "w") as f:
        f.write("ncu launch list `%s` (gpu__time_duration.sum, --clock-control none; cold-cache, serialised: compare shares)\n\n" % path)
        f.write("| kernel | launches | total us | share | max us |\n|---|---|---|---|---|\n")
        for k, (n, v, mx) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write("| %s | %d | %.1f | %.1f%% | %.1f |\n" % (k, n, v / 1e3, 100 * v / tot, mx / 1e3))
    print(open(out).read())


def full(path, out):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    h, units = rows[0], rows[1]
    with open(out, "w") as f:
        f.write("ncu --set full capture `%s` (per launch; --clock-control none)\n\n" % path)
        names = [short(r[h.index("Kernel Name")]) for r in rows[2:]]
        f.write("| metric | unit | " + " | ".join("%s #%d" % (n, i) for i, n in enumerate(names)) + " |\n")
        f.write("|---|---|" + "---|" * len(names) + "\n")
        for k in KEYS:
            if k in h:
                i = h.index(k)
                f.write("| %s | %s | %s |\n" % (k, units[i], " | ".join(r[i] for r in rows[2:])))
    print(open(out).read())


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
