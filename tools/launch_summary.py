"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list as a markdown table (kernel, launches,
total us, avg us, share of the listed time).  usage: python tools/launch_summary.py launches.csv > profiles/rNN_launches.md"""
import collections
import csv
import sys


def main(path):
    rows = [l for l in open(path) if not l.startswith("==")]
    tot = collections.OrderedDict()
    for r in csv.DictReader(rows):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = r["Kernel Name"].split("(")[0].split("::")[-1]
        if name.startswith("void "):
            name = name[5:]
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        us = v * {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3}.get(unit, 1e-3)
        t = tot.setdefault(name, [0, 0.0])
        t[0] += 1
        t[1] += us
    total = sum(t[1] for t in tot.values()) or 1.0
    print("| kernel | launches | total us | avg us | share |")
    print("|---|---|---|---|---|")
    for name, (n, us) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
        print("| %s | %d | %.1f | %.1f | %.3f |" % (name, n, us, us / n, us / total))


if __name__ == "__main__":
    main(sys.argv[1])
