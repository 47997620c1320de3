"""Generates tests/golden/*.npz.  Run in the authoring container (needs cv2 4.13 and, for the extractor
fixtures, oracle/_ref/liborbref.so built from /root/reference).  The fixtures pin:

  prim_cv2.npz        cv2 4.13.0 outputs of the five OpenCV primitives on small seeded inputs
                      (resize chain, copyMakeBorder, GaussianBlur, FAST 20/7, fastAtan2)
  sincosf_libm.npz    host glibc 2.39 sinf/cosf on 4096 seeded angles deg*factorPI
  extract_ref_*.npz   keypoints + descriptors produced by the reference's own ORBextractor.cc
                      (compiled unmodified against oracle/shim, canonical tie-break) on seeded synthetic
                      frames of reduced size (fast to re-run) and on the TUM-shaped frame, seed 0.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cv2  # noqa: E402

import oracle  # noqa: E402
from orbslam2commentedbyxcm_b200 import synth  # noqa: E402

G = os.path.join(ROOT, "tests", "golden")
os.makedirs(G, exist_ok=True)


def prim():
    rng = np.random.Generator(np.random.PCG64(7))
    out = {}
    img = synth.gen_frame((83, 97), 3)
    out["img"] = img
    prev = img
    for l in range(1, 5):
        s = np.float32(1.0) / (np.float32(1.2) ** l)
        dw = int(np.rint(np.float32(img.shape[1]) * s))
        dh = int(np.rint(np.float32(img.shape[0]) * s))
        prev = cv2.resize(prev, (dw, dh), interpolation=cv2.INTER_LINEAR)
        out["resize_l%d" % l] = prev
    out["border19"] = cv2.copyMakeBorder(img, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
    out["blur"] = cv2.GaussianBlur(img, (7, 7), 2, None, 2, cv2.BORDER_REFLECT_101)
    for th in (20, 7):
        det = cv2.FastFeatureDetector_create(th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
        kps = det.detect(img)
        out["fast_t%d" % th] = np.array([(int(k.pt[0]), int(k.pt[1]), int(k.response)) for k in kps], np.int32).reshape(-1, 3)
    y = rng.integers(-200000, 200000, 4096).astype(np.float32)
    x = rng.integers(-200000, 200000, 4096).astype(np.float32)
    y[:5] = [0, 0, 1, 0, -1]
    x[:5] = [0, 1, 0, -1, 0]
    out["atan_y"], out["atan_x"] = y, x
    out["atan_deg"] = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], np.float32)
    np.savez_compressed(os.path.join(G, "prim_cv2.npz"), **out)


def sincos():
    rng = np.random.Generator(np.random.PCG64(11))
    deg = (rng.random(4096) * 360).astype(np.float32)
    deg[:4] = [0, 90, 180, 360]
    factor = np.float32(np.float32(np.pi) / np.float32(180.0))
    # factorPI = (float)(CV_PI/180.f): double division then cast
    factor = np.float32(3.1415926535897932384626433832795 / float(np.float32(180.0)))
    ang = (deg * factor).astype(np.float32)
    s = np.zeros_like(ang)
    c = np.zeros_like(ang)
    oracle.lib().orbo_libm_sincosf(ang.ctypes.data, s.ctypes.data, c.ctypes.data, len(ang))
    np.savez_compressed(os.path.join(G, "sincosf_libm.npz"), angle=ang, sin=s, cos=c)


def extract():
    cases = {
        "small_a": ((120, 160), 1, (300, 1.2, 4, 20, 7)),
        "small_b": ((200, 333), 2, (500, 1.2, 6, 20, 7)),
        "tum_s0": (synth.SHAPES["tum"], 0, synth.PRESETS["tum"]),
    }
    for name, (shape, seed, cfg) in cases.items():
        img = synth.gen_frame(shape, seed)
        ex = oracle.Extractor(*cfg, kind="ref")
        kps, desc = ex.extract(img)
        np.savez_compressed(
            os.path.join(G, "extract_ref_%s.npz" % name),
            shape=np.array(shape), seed=np.array(seed), cfg=np.array(cfg, np.float64), kps=kps, desc=desc,
        )
        print(name, len(kps))


UNDISTORT_CASES = [  # (fx, fy, cx, cy, (k1, k2, p1, p2, k3)) : TUM1, TUM2, TUM3-like (no distortion) and a 4-coefficient camera
    (517.306408, 516.469215, 318.643040, 255.313989, (0.262383, -0.953104, -0.005358, 0.002628, 1.163314)),
    (520.908620, 521.007327, 325.141442, 249.701764, (0.231222, -0.784899, -0.003257, -0.000105, 0.917205)),
    (535.4, 539.2, 320.1, 247.6, (0.0, 0.0, 0.0, 0.0, 0.0)),
    (458.654, 457.296, 367.215, 248.375, (-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05)),
]


def undistort():
    """cv2.undistortPoints(pts, K, dist, None, K) -- the call of Frame::UndistortKeyPoints (src/Frame.cc:608-614)."""
    import cv2

    rng = np.random.Generator(np.random.PCG64(11))
    out = {}
    for c, (fx, fy, cx, cy, d) in enumerate(UNDISTORT_CASES):
        K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
        D = np.array(d, np.float32)
        pts = (rng.random((3000, 1, 2)) * [752, 480]).astype(np.float32)
        pts[:4, 0] = [[0, 0], [751, 479], [cx, cy], [19.5, 460.25]]
        if D[0] == 0:
            res = pts.reshape(-1, 2).copy()      # the reference skips the call (src/Frame.cc:589-592)
        else:
            res = cv2.undistortPoints(pts, K, D, None, K).reshape(-1, 2)
        out["pts%d" % c] = pts.reshape(-1, 2)
        out["res%d" % c] = res
        out["cam%d" % c] = np.array([fx, fy, cx, cy], np.float32)
        out["dist%d" % c] = D
    np.savez_compressed(os.path.join(G, "undistort_cv2.npz"), **out)


def matcher():
    """tests/golden/matcher_ref.npz: what the REFERENCE'S OWN matcher-side code (oracle/_ref/libmatchref.so: ORBmatcher.cc,
    Frame.cc, KeyFrame.cc, MapPoint.cc, DBoW2 compiled unmodified) returns on every case of tests/matcher_cases.py.  The
    GPU tests compare the CUDA path with these arrays (tests/test_gpu_matcher_ref.py); the inputs are regenerated from
    seeds on both sides."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import matcher_cases as mc
    from oracle import matchref as R

    out = mc.reference_outputs(oracle, R)
    np.savez_compressed(os.path.join(G, "matcher_ref.npz"), **out)
    print("matcher_ref.npz: %d arrays" % len(out))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "matcher":
        matcher()
        sys.exit(0)
    prim()
    undistort()
    sincos()
    matcher()
    extract()
    print(sorted(os.listdir(G)))
