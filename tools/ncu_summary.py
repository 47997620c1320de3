"""Summarise an ncu launch list (csv from --metrics gpu__time_duration.sum) and/or a full .ncu-rep into markdown.
usage: python tools/ncu_summary.py --launches gpurun_out/launches.csv --rep gpurun_out/prof.ncu-rep > profiles/xx.md"""
import argparse
import collections
import csv
import subprocess
import sys


def launches(path):
    rows = list(csv.reader(open(path)))
    hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
    hdr = rows[hi]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[hi + 1:]:
        if len(r) <= vi:
            continue
        k = r[ki].split("(")[0]
        v = float(r[vi].replace(",", ""))
        v = v / 1e3 if r[ui] == "ns" else (v * 1e3 if r[ui] == "ms" else v)
        agg.setdefault(k, []).append(v)
    tot = sum(sum(v) for v in agg.values())
    print("| kernel | launches | mean us | total us | share of listed time |")
    print("|---|---:|---:|---:|---:|")
    for k, v in agg.items():
        print("| %s | %d | %.1f | %.1f | %.1f%% |" % (k, len(v), sum(v) / len(v), sum(v), 100 * sum(v) / tot))
    print()


WANT = [
    ("Kernel Name", "kernel"), ("launch__grid_size", "grid"), ("launch__registers_per_thread", "regs"),
    ("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram rd"), ("dram__bytes_write.sum", "dram wr"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"),
    ("smsp__inst_executed.sum", "warp insts"), ("launch__waves_per_multiprocessor", "waves"),
]


def rep(path):
    out = subprocess.check_output(["ncu", "-i", path, "--page", "raw", "--csv"], stderr=subprocess.DEVNULL).decode()
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = [(hdr.index(k) if k in hdr else -1, n) for k, n in WANT]
    print("| " + " | ".join("%s%s" % (n, (" [" + units[i] + "]") if i >= 0 and units[i] else "") for i, n in idx) + " |")
    print("|" + "---|" * len(idx))
    for r in rows[2:]:
        cells = []
        for i, n in idx:
            v = r[i] if i >= 0 else "?"
            if n == "kernel":
                v = v.split("(")[0]
            else:
                try:
                    f = float(v.replace(",", ""))
                    v = ("%.0f" % f) if abs(f) >= 1000 or f == int(f) else ("%.2f" % f)
                except ValueError:
                    pass
            cells.append(v)
        print("| " + " | ".join(cells) + " |")
    print()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--launches")
    ap.add_argument("--rep")
    a = ap.parse_args()
    if a.launches:
        launches(a.launches)
    if a.rep:
        rep(a.rep)
