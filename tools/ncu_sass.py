"""Per-SASS-instruction executed counts / stall samples of one kernel from an .ncu-rep (--set full --import-source on).
usage: python tools/ncu_sass.py report.ncu-rep kernel-regex [launch-index]"""
import csv
import io
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "-k", "regex:" + kern],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
launch = -1
h = None
body = []
for r in rows:
    if r and r[0] == "Kernel Name":
        launch += 1
        continue
    if r and r[0] == "Address":
        h = r
        continue
    if launch == which and h is not None and len(r) == len(h):
        body.append(r)
ci = h.index("Instructions Executed")
cs = h.index("Warp Stall Sampling (All Samples)")
ct = h.index("Avg. Threads Executed")
tot = sum(float(r[ci]) for r in body)
print("total warp insts %.0f, %d SASS lines" % (tot, len(body)))
base = int(body[0][0], 16)
for r in body:
    print("%04x %9.0f %5.2f%% st%4s thr%3s | %s" % (int(r[0], 16) - base, float(r[ci]), 100 * float(r[ci]) / tot, r[cs], r[ct][:2], r[1].strip()))
