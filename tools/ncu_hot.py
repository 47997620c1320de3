"""Reads an .ncu-rep captured with `--set full --import-source on`: per kernel the top warp-stall reasons (raw page) and
the source lines with the most executed instructions / stall samples (source page).
usage: python tools/ncu_hot.py report.ncu-rep [kernel-substring] [top-n]"""
import csv
import io
import subprocess
import sys
from collections import defaultdict


def page(rep, name, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv", *extra], capture_output=True, text=True).stdout
    return out


def stalls(rep):
    r = list(csv.reader(io.StringIO(page(rep, "raw"))))
    h = r[0]
    names = [n for n in h if n.startswith("smsp__average_warps_issue_stalled_") and n.endswith("_per_issue_active.ratio")]
    seen = set()
    for row in r[2:]:
        k = row[h.index("Kernel Name")].split("(")[0] + " grid " + row[h.index("Grid Size")] if "Grid Size" in h else row[h.index("Kernel Name")]
        if k in seen:
            continue
        seen.add(k)
        d = sorted(((float(row[h.index(n)] or 0), n[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]) for n in names), reverse=True)[:6]
        print("%-34s %s" % (k[:34], "  ".join("%s=%.2f" % (n, v) for v, n in d)))


def source(rep, kern, top):
    """cuda,sass view: rows with a line number are per-source-line aggregates, rows with '-' are SASS instructions."""
    out = page(rep, "source", ["--print-source", "cuda,sass", "-k", "regex:" + kern])
    rows = list(csv.reader(io.StringIO(out)))
    h = None
    seen = set()
    cur = []
    fn = ""

    def flush():
        if not cur or h is None or fn in seen:
            return
        seen.add(fn)
        ci = h.index("Instructions Executed")
        cs = h.index("Warp Stall Sampling (All Samples)")
        ca = h.index("Avg. Threads Executed")
        tot_i = sum(float(r[ci] or 0) for r in cur) or 1
        tot_s = sum(float(r[cs] or 0) for r in cur) or 1
        print("=====", fn, "warp insts %.0f, samples %.0f" % (tot_i, tot_s))
        for r in sorted(cur, key=lambda r: -float(r[ci] or 0))[:top]:
            print("%5s %6.2f%% inst %6.2f%% stall thr %4s | %s" % (r[0], 100 * float(r[ci] or 0) / tot_i, 100 * float(r[cs] or 0) / tot_s, r[ca][:4], r[1].strip()[:120]))

    for r in rows:
        if not r:
            continue
        if r[0] == "Function Name":
            flush()
            cur = []
            fn = r[1]
        elif r[0] == "Line No":
            h = r
        elif r[0].isdigit() and h is not None and len(r) == len(h):
            cur.append(r)
    flush()


if __name__ == "__main__":
    rep = sys.argv[1]
    kern = sys.argv[2] if len(sys.argv) > 2 else ""
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    stalls(rep)
    if kern != "-":
        source(rep, kern, top)
