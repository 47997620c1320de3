"""Markdown evidence for every kernel in an .ncu-rep captured with `--set full --import-source on`:
per kernel (first launch of each name + grid): duration, DRAM bytes, DRAM / SM / issue utilisation, pipe utilisation
(ALU, FMA, LSU, XU, tensor, uniform), L1 / L2 hit rates, occupancy, registers, top stall reasons -- and a SASS opcode
histogram (executed warp instructions per opcode, from the source page) for the kernels matching --sass.
usage: python tools/ncu_report.py report.ncu-rep [--sass regex ...] > profiles/xx.md"""
import argparse
import collections
import csv
import io
import re
import subprocess


def page(rep, name, extra=()):
    return subprocess.run(["ncu", "-i", rep, "--page", name, "--csv", *extra], capture_output=True, text=True).stdout


COLS = [
    ("us", "gpu__time_duration.sum"),
    ("DRAM rd MB", "dram__bytes_read.sum"),
    ("DRAM wr MB", "dram__bytes_write.sum"),
    ("DRAM %", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
    ("SM %", "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("issue %", "sm__issue_active.avg.pct_of_peak_sustained_elapsed"),
    ("ALU %", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
    ("FMA %", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
    ("LSU %", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
    ("XU %", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
    ("tensor %", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
    ("L1 hit %", "l1tex__t_sector_hit_rate.pct"),
    ("L2 hit %", "lts__t_sector_hit_rate.pct"),
    ("occ %", "sm__warps_active.avg.pct_of_peak_sustained_active"),
    ("regs", "launch__registers_per_thread"),
    ("smem KB", "launch__shared_mem_per_block_dynamic"),
    ("warp inst", "smsp__inst_executed.sum"),
]


def to_float(v, unit, want):
    try:
        x = float(v.replace(",", ""))
    except ValueError:
        return None
    u = unit.lower()
    if want == "us":
        return x / 1e3 if u in ("ns", "nsecond") else x * 1e3 if u in ("ms", "msecond") else x * 1e6 if u in ("s", "second") else x
    if want.endswith("MB"):
        return x / 1e6 if u == "byte" else x / 1e3 if u == "kbyte" else x * 1e3 if u == "gbyte" else x
    if want == "smem KB":
        return x / 1e3 if u == "byte" else x if u == "kbyte" else x
    return x


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("--sass", nargs="*", default=[])
    ap.add_argument("--top", type=int, default=14)
    ap.add_argument("--show", nargs="*", default=["UTCIMMA", "UTMALDG", "LDTM", "UTCBAR", "UTMAPF", "UBLKCP", "POPC", "LDGSTS"],
                    help="opcodes listed even when they are not among the top ones")
    a = ap.parse_args()
    rows = list(csv.reader(io.StringIO(page(a.rep, "raw"))))
    h, units = rows[0], rows[1]
    ki = h.index("Kernel Name")
    gi = h.index("Grid Size") if "Grid Size" in h else None
    stall_cols = [(i, n[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]) for i, n in enumerate(h)
                  if n.startswith("smsp__average_warps_issue_stalled_") and n.endswith("_per_issue_active.ratio")]
    seen = collections.OrderedDict()
    for r in rows[2:]:
        name = r[ki].split("(")[0]
        key = (name, r[gi] if gi is not None else "")
        seen.setdefault(key, []).append(r)
    print("| kernel | grid | launches | " + " | ".join(c for c, _ in COLS) + " | top stalls (warps per issue) |")
    print("|---|---|---:|" + "---:|" * len(COLS) + "---|")
    for (name, grid), rs in seen.items():
        r = rs[-1]   # the last launch of this shape (warm)
        vals = []
        for label, metric in COLS:
            if metric not in h:
                vals.append("-")
                continue
            i = h.index(metric)
            x = to_float(r[i], units[i], label)
            vals.append("-" if x is None else ("%.0f" % x if label in ("regs", "warp inst") else "%.1f" % x))
        st = sorted(((float(r[i] or 0), n) for i, n in stall_cols), reverse=True)[:3]
        print("| %s | %s | %d | %s | %s |" % (name, grid.replace(" ", ""), len(rs), " | ".join(vals), ", ".join("%s %.2f" % (n, v) for v, n in st)))
    for rx in a.sass:
        out = page(a.rep, "source", ["--print-source", "sass", "-k", "regex:" + rx])
        srows = list(csv.reader(io.StringIO(out)))
        hist, total, hdr, kname, launches = collections.Counter(), 0, None, None, 0
        for r in srows:
            if r and r[0] == "Kernel Name":
                launches += 1
                continue
            if r and r[0] == "Address":
                hdr = r
                continue
            if hdr is None or len(r) < len(hdr) or launches > 1:
                continue
            try:
                n = int(float(r[hdr.index("# Warp Instructions Executed")] if "# Warp Instructions Executed" in hdr else r[hdr.index("Instructions Executed")]))
            except (ValueError, IndexError):
                continue
            src = r[hdr.index("Source")]
            mnem = re.sub(r"^@!?U?P\d+\s+", "", src.strip()).split()[0].rstrip(";") if src.strip() else "?"
            op = mnem.split(".")[0]
            hist[op] += n
            total += n
        print("\n### SASS opcode histogram: `%s` (executed warp instructions of the first matching launch, %d total)\n" % (rx, total))
        print("| opcode | warp instructions | share |")
        print("|---|---:|---:|")
        top = hist.most_common(a.top)
        for op, n in top:
            print("| %s | %d | %.1f%% |" % (op, n, 100.0 * n / max(total, 1)))
        for op in a.show:
            if op in hist and op not in dict(top):
                print("| %s | %d | %.3f%% |" % (op, hist[op], 100.0 * hist[op] / max(total, 1)))


if __name__ == "__main__":
    main()
