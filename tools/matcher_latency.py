"""Host-call latency (uploads + kernels + result download, median of 200 calls) of the per-frame matcher entry points on the
seeded cases of the parity tests."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import matcher_cases as mc
import oracle
import orbslam2commentedbyxcm_b200 as pkg

def med(fn, n=200):
    for _ in range(20): fn()
    t = []
    for _ in range(n):
        t0 = time.perf_counter(); fn(); t.append(time.perf_counter() - t0)
    return 1e3 * float(np.median(t))

m = pkg.ORBmatcher(0.9, True)
frame, qq = mc.window_case(oracle, "kitti", 0, "forward")
print("SearchWindow (KITTI frame)        %.3f ms" % med(lambda: m.SearchWindow(frame, qq, 100, 0)))
f1, f2, prev = mc.init_case(oracle, "tum", 6)
print("SearchForInitialization (TUM)     %.3f ms" % med(lambda: m.SearchForInitialization(f1, f2, prev, 100)))
kfa, kfb = mc.bow_case(oracle, 0, kfkf=True)
print("SearchByBoW KF-KF                 %.3f ms" % med(lambda: m.SearchByBoWKF(kfa, kfb)))
kf1, kf2, F12, exx, eyy, sf, sigma2 = mc.triang_case(oracle, 1, 0.5)
print("SearchForTriangulation            %.3f ms" % med(lambda: m.SearchForTriangulation(kf1, kf2, F12, exx, eyy, sf, sigma2, False)))
fr, q2, inv_sigma2 = mc.fuse_case(oracle, "euroc", 1, True)
print("Fuse search                       %.3f ms" % med(lambda: m.FuseSearch(fr, q2, gate=1, inv_level_sigma2=inv_sigma2)))
