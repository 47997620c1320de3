import sys, os
sys.path.insert(0, '/root/repo')
import numpy as np
import orbslam2commentedbyxcm_b200 as pkg
from orbslam2commentedbyxcm_b200 import synth
img = synth.gen_frame((240, 320), 1)
ex = pkg.ORBextractor(500, 1.2, 8, 20, 7, device=0)
if os.environ.get("NOGRAPH"): ex.set_use_graph(False)
try:
    kps, desc = ex(img)
    print("ok", len(kps))
except Exception as e:
    print("FAIL", str(e)[-200:])
