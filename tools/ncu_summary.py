#!/usr/bin/env python
"""Summarise ncu output for profiles/:  launch list (gpu__time_duration per kernel) and the raw page
of a --set full capture (one line per captured launch with the metrics DESIGN.md quotes).

  python tools/ncu_summary.py launches gpurun_out/r01_launches.csv
  python tools/ncu_summary.py full     gpurun_out/r01_full.ncu-rep
"""
import collections
import csv
import subprocess
import sys

FULL_METRICS = [
    ("gpu__time_duration.sum", "time"),
    ("launch__grid_size", "grid"),
    ("launch__registers_per_thread", "regs"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("lts__t_bytes.sum", "l2_bytes"),
    ("lts__t_sector_hit_rate.pct", "l2_hit%"),
    ("l1tex__t_sector_hit_rate.pct", "l1_hit%"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
    ("sm__issue_active.avg.pct_of_peak_sustained_elapsed", "issue%"),
    ("sm__inst_executed.avg.per_cycle_elapsed", "ipc"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "thr/inst"),
    ("smsp__inst_executed.sum", "warp_insts"),
]


def launches(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        try:
            v = float(row["Metric Value"].replace(",", ""))
        except (KeyError, ValueError):
            continue
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(row["Metric Unit"], 1.0)
        agg.setdefault(row["Kernel Name"].split("(")[0], []).append(v)
    tot = sum(sum(v) for v in agg.values())
    print("| kernel | launches | total us | avg us | share |\n|---|---|---|---|---|")
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print("| %s | %d | %.1f | %.1f | %.3f |" % (k, len(v), sum(v), sum(v) / len(v), sum(v) / tot))


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    cols = [(hdr.index(m), short) for m, short in FULL_METRICS if m in hdr]
    print("| # | kernel | " + " | ".join("%s [%s]" % (s, units[i]) if units[i] else s for i, s in cols) + " |")
    print("|---|---|" + "---|" * len(cols))
    for n, d in enumerate(data):
        name = d[hdr.index("Kernel Name")].split("(")[0]
        vals = []
        for i, _ in cols:
            try:
                x = float(d[i].replace(",", ""))
                vals.append("%.4g" % x)
            except ValueError:
                vals.append(d[i])
        print("| %d | %s | " % (n, name) + " | ".join(vals) + " |")
    # stall sampling of each kernel's first captured launch
    seen = set()
    for d in data:
        name = d[hdr.index("Kernel Name")].split("(")[0]
        if name in seen:
            continue
        seen.add(name)
        st = [(h.replace("smsp__pcsamp_warps_issue_stalled_", ""), float(d[i].replace(",", "") or 0))
              for i, h in enumerate(hdr) if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued")]
        tot = sum(v for _, v in st) or 1.0
        top = sorted(st, key=lambda x: -x[1])[:7]
        print("\nstall samples, first captured launch of %s: " % name + ", ".join("%s %.1f%%" % (k, 100 * v / tot) for k, v in top))


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
