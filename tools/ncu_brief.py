"""One line of headline metrics + the stall breakdown for every launch in an .ncu-rep (ncu --set full).
usage: python tools/ncu_brief.py report.ncu-rep"""
import csv
import subprocess
import sys

WANT = [("time us", "gpu__time_duration.sum"), ("regs", "launch__registers_per_thread"),
        ("occ %", "sm__warps_active.avg.pct_of_peak_sustained_active"), ("warp insts", "smsp__inst_executed.sum"),
        ("thr/inst", "smsp__thread_inst_executed_per_inst_executed.ratio"), ("ipc", "sm__inst_executed.avg.per_cycle_elapsed"),
        ("dram rd MB", "dram__bytes_read.sum"), ("dram wr MB", "dram__bytes_write.sum"),
        ("l2 hit %", "lts__t_sector_hit_rate.pct"), ("l1 hit %", "l1tex__t_sector_hit_rate.pct"),
        ("dram %", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
        ("smem KB", "launch__shared_mem_per_block_dynamic")]


def main(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    stall = [(i, h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""))
             for i, h in enumerate(hdr) if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")].split("(")[0]
        out = [name]
        for label, key in WANT:
            if key in hdr:
                i = hdr.index(key)
                v, u = r[i].replace(",", ""), units[i]
                try:
                    f = float(v)
                    if u in ("ns", "nsecond"):
                        f /= 1e3
                    if u in ("ms", "msecond") and "us" in label:
                        f *= 1e3
                    if u in ("s", "second") and "us" in label:
                        f *= 1e6
                    if u == "byte" and "MB" in label:
                        f /= 1e6
                    if u == "Kbyte" and "MB" in label:
                        f /= 1e3
                    if u == "Gbyte" and "MB" in label:
                        f *= 1e3
                    if u == "byte" and "KB" in label:
                        f /= 1e3
                    out.append("%s %.4g" % (label, f))
                except ValueError:
                    out.append("%s %s" % (label, v))
        print(" | ".join(out))
        vals = [(float(r[i].replace(",", "")), n) for i, n in stall if r[i]]
        tot = sum(v for v, _ in vals) or 1.0
        print("    stalls: " + ", ".join("%s %.1f%%" % (n, 100 * v / tot) for v, n in sorted(vals, reverse=True)[:8]))


if __name__ == "__main__":
    main(sys.argv[1])
