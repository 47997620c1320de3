"""Regenerates orbslam2commentedbyxcm_b200/csrc/orb_pattern.inc (DATA ONLY: the 256 rBRIEF test pairs) from the
reference tree's bit_pattern_31_ table (src/ORBextractor.cc:176-434).  Run in the authoring container only."""
import re
import sys

ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/src/ORBextractor.cc"
src = open(ref, encoding="utf-8").read().split("\n")
body = "\n".join(src[177:433])
body = re.sub(r"/\*.*?\*/", "", body)
body = re.sub(r"//.*", "", body)
vals = [int(v) for v in re.findall(r"-?\d+", body)]
assert len(vals) == 1024 and max(vals) <= 13 and min(vals) >= -13
hdr = """// rBRIEF sampling pattern: 256 test pairs (x0,y0,x1,y1), i.e. 512 points, 1024 signed bytes.
// DATA ONLY. This is the learned ORB 31x31 pattern that ORB-SLAM2 embeds as
// bit_pattern_31_[256*4] (reference src/ORBextractor.cc:176-434; originally OpenCV's orb.cpp).
// The descriptor definition depends on these exact values, so they cannot be re-derived;
// tools/gen_pattern.py regenerates this file from the reference tree.
// Layout: entry 4*k+0..3 = pair k; bit (k%8) of descriptor byte (k/8) = I(p0) < I(p1).
"""
lines = ["  " + ",".join("%3d" % v for v in vals[i : i + 16]) + "," for i in range(0, 1024, 16)]
open("orbslam2commentedbyxcm_b200/csrc/orb_pattern.inc", "w").write(hdr + "\n".join(lines) + "\n")
