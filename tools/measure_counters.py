"""Per-kernel hardware counters of ONE step of bench.py's workload, measured under ncu ON THE GPU BOX and written to
profiles/kernel_counters.json together with the fingerprint of the build they belong to.  bench.py reads the file
for roofline.traffic / roofline.issue / roofline.l2 and refuses it when the fingerprint differs from the library's.

  gpurun -- python tools/measure_counters.py            (then copy gpurun_out/kernel_counters.json to profiles/)

One frame is issued as ONE wavefront, one frame at a time, so a launch's counters belong to that launch alone;
per-kernel numbers are means over the frame's launches of that kernel.  A number printed by the run under ncu is
never a bench value: only the counters are kept.
"""
import collections
import csv
import datetime
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
METRICS = ["dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum", "smsp__inst_executed.sum",
           "smsp__thread_inst_executed.sum", "gpu__time_duration.sum"]
ALIAS = {"k_shade_sorted": "k_shade", "k_shade_sorted_est": "k_shade", "k_shade_est": "k_shade"}


def main():
    import mcrt_b200  # noqa: F401
    from mcrt_b200 import build
    out_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    log = os.path.join(out_dir, "kernel_counters.csv")
    env = dict(os.environ, MCRT_OVERLAP_HALVES="1")
    split = os.environ.get("MCRT_SHADE_SPLIT", "1") != "0"      # PTB_DEFAULT_SHADE_SPLIT = 1
    per_frame = 16 if split else 11      # trace / shade launches of one frame: closest, 4 x mixed, shadow, 5 x (1 or 2) shade
    # warm-up 3 steps (the bench's minimum) + 1 timed + 1 profiled-stage step + e2e: skip the warm-up's launches and
    # capture the launches of the first timed frame: raygen, closest, 5 x shade, 4 x mixed, shadow, control kernels
    cmd = ["ncu", "--metrics", ",".join(METRICS), "--clock-control", "none", "--csv", "--log-file", log,
           "-k", "regex:k_trace|k_shade", "-s", str(3 * per_frame), "-c", str(per_frame),
           sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "3", "--frames-in-flight", "1",
           "--halves", "1", "--no-cpu-baseline", "--no-extras"]
    subprocess.run(cmd, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, check=True)
    rows = [l for l in open(log) if not l.startswith("==")]
    per = collections.OrderedDict()
    for r in csv.DictReader(rows):
        name = r["Kernel Name"].split("(")[0].split("::")[-1]
        if name.startswith("void "):
            name = name[5:]
        name = name.split("<")[0]          # k_trace_mixed<0> / <1> (flat / two-level) -> k_trace_mixed
        name = ALIAS.get(name, name)
        try:
            v = float(r["Metric Value"].replace(",", ""))
        except ValueError:
            continue
        unit = r.get("Metric Unit", "")
        scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9,
                 "second": 1.0, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}.get(unit, 1.0)
        per.setdefault((name, r["ID"]), {})[r["Metric Name"]] = v * scale
    kernels = {}
    for (name, _), m in per.items():
        kernels.setdefault(name, []).append(m)
    summary = {}
    for name, ms in kernels.items():
        n = len(ms)
        mean = lambda k: sum(m.get(k, 0.0) for m in ms) / n
        summary[name] = {"launches": n,
                         "dram_bytes_per_launch": mean("dram__bytes_read.sum") + mean("dram__bytes_write.sum"),
                         "l2_bytes_per_launch": mean("lts__t_bytes.sum"),
                         "warp_insts_per_launch": mean("smsp__inst_executed.sum"),
                         "threads_per_inst": mean("smsp__thread_inst_executed.sum") / max(mean("smsp__inst_executed.sum"), 1.0),
                         "ncu_time_per_launch_s": mean("gpu__time_duration.sum")}
    if "k_shade_surface" in summary and "k_shade_scatter" in summary:
        # the shade STAGE of the split build: one "launch" = one bounce = the two kernels, as bench.py's stage timer counts
        a, b = summary["k_shade_surface"], summary["k_shade_scatter"]
        n = max(a["launches"], 1)
        tot = lambda k: a[k] * a["launches"] + b[k] * b["launches"]
        wi = tot("warp_insts_per_launch")
        summary["k_shade"] = {"launches": n, "dram_bytes_per_launch": tot("dram_bytes_per_launch") / n,
                              "l2_bytes_per_launch": tot("l2_bytes_per_launch") / n, "warp_insts_per_launch": wi / n,
                              "threads_per_inst": (a["threads_per_inst"] * a["warp_insts_per_launch"] * a["launches"] +
                                                   b["threads_per_inst"] * b["warp_insts_per_launch"] * b["launches"]) / max(wi, 1.0),
                              "ncu_time_per_launch_s": tot("ncu_time_per_launch_s") / n,
                              "note": "k_shade_surface + k_shade_scatter per bounce"}
    doc = {"build_fingerprint": build._fingerprint()[:16], "when": datetime.datetime.utcnow().strftime("%Y-%m-%dT%H:%MZ"),
           "workload": "bench.py default (sponza_class 1920x1080 depth 5), one frame as one wavefront",
           "how": "ncu --metrics %s --clock-control none, the trace / shade launches of the first timed frame" % ",".join(METRICS),
           "kernels": summary}
    with open(os.path.join(out_dir, "kernel_counters.json"), "w") as f:
        json.dump(doc, f, indent=1)
    print(json.dumps(doc, indent=1))


if __name__ == "__main__":
    main()
