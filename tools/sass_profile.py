#!/usr/bin/env python
"""Join an ncu SASS source page (ncu -i X.ncu-rep --page source --csv --print-source sass) with the line table of
the same cubin (nvdisasm -g -c) and aggregate executed warp instructions / stall samples by source function and line.

  python tools/sass_profile.py shade_sass.csv pt.sass _ZN4mcrt7k_shade [topN]
"""
import collections
import csv
import os
import re
import sys


def line_table(sass_path, mangled_prefix):
    tab, cur, on = {}, None, False
    for l in open(sass_path):
        if l.startswith(".text."):
            on = l.startswith(".text." + mangled_prefix)
            continue
        if not on:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
        if m:
            tab[int(m.group(1), 16)] = (cur, m.group(2).strip())
    return tab


_fn_cache = {}


def enclosing_function(path, line):
    if path not in _fn_cache:
        fns = []
        try:
            src = open(path).read().split("\n")
        except OSError:
            src = []
        for i, l in enumerate(src, 1):
            m = re.match(r"^(?:template\s*<[^>]*>\s*)?(?:HD|__device__|__global__|static|inline|__forceinline__|\s)*[\w:<>\*&\s]+?\b(\w+)\s*\([^;]*$", l)
            if m and not l.startswith(" ") and not l.startswith("//") and "=" not in l.split("(")[0]:
                fns.append((i, m.group(1)))
        _fn_cache[path] = fns
    name = "?"
    for i, n in _fn_cache[path]:
        if i <= line:
            name = n
    return name


def main():
    csv_path, sass_path, prefix = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    tab = line_table(sass_path, prefix)
    rows = list(csv.reader(open(csv_path)))
    hdr = rows[1]
    ia, ii, it, isamp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
    base = None
    by_fn, by_line, by_op = collections.defaultdict(lambda: [0, 0, 0]), collections.defaultdict(lambda: [0, 0, 0]), collections.Counter()
    tot = [0, 0, 0]
    for r in rows[2:]:
        if r and r[0] == "Kernel Name":      # a second launch in the same page: the first one is the profile
            break
        if len(r) != len(hdr) or r[ia] == "Address":
            continue
        a = int(r[ia], 16)
        base = a if base is None else base
        loc, txt = tab.get(a - base, (None, r[1]))
        v = [int(r[ii] or 0), int(r[it] or 0), int(r[isamp] or 0)]
        fn = ("%s:%s" % (os.path.basename(loc[0]), enclosing_function(loc[0], loc[1]))) if loc else "?"
        ln = ("%s:%d" % (os.path.basename(loc[0]), loc[1])) if loc else "?"
        op = txt.split()[0] if not txt.startswith("@") else txt.split()[1]
        by_op[op.split(".")[0]] += v[0]
        for d, k in ((by_fn, fn), (by_line, ln)):
            for j in range(3):
                d[k][j] += v[j]
        for j in range(3):
            tot[j] += v[j]
    print("total warp instructions %d, threads/inst %.2f, samples %d" % (tot[0], tot[1] / max(tot[0], 1), tot[2]))
    for title, d in (("function", by_fn), ("line", by_line)):
        print("\n%-52s %8s %8s %8s" % (title, "inst%", "samples%", "thr/inst"))
        for k, v in sorted(d.items(), key=lambda kv: -kv[1][0])[:top]:
            print("%-52s %8.2f %8.2f %8.1f" % (k, 100.0 * v[0] / tot[0], 100.0 * v[2] / max(tot[2], 1), v[1] / max(v[0], 1)))
    print("\nopcodes:", ", ".join("%s %.1f%%" % (k, 100.0 * v / tot[0]) for k, v in by_op.most_common(25)))


main()
