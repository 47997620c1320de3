"""CPU: the oracle's restated OpenCV primitives against the committed cv2 4.13.0 fixtures (tests/golden/prim_cv2.npz,
made by tools/make_golden.py) and, when cv2 is importable, against cv2 live on extra seeded inputs."""
import ctypes
import os

import numpy as np
import pytest

from orbslam2commentedbyxcm_b200 import synth


def _resize(L, src, dw, dh):
    out = np.empty((dh, dw), np.uint8)
    src = np.ascontiguousarray(src)
    L.orbo_resize_linear_u8(src.ctypes.data, src.shape[1], src.shape[0], src.strides[0], out.ctypes.data, dw, dh, dw)
    return out


def _blur(L, src):
    src = np.ascontiguousarray(src)
    out = np.empty_like(src)
    L.orbo_gaussian_blur_7x7_s2(src.ctypes.data, src.shape[1], src.shape[0], src.strides[0], out.ctypes.data, out.strides[0])
    return out


def _fast(L, img, th):
    img = np.ascontiguousarray(img)
    buf = np.zeros((img.size, 3), np.int32)
    n = L.orbo_fast9_16(img.ctypes.data, img.shape[1], img.shape[0], img.strides[0], th, 1, buf.ctypes.data, img.size)
    return buf[:n].copy()


def _atan(L, y, x):
    out = np.empty_like(y)
    L.orbo_fast_atan2(y.ctypes.data, x.ctypes.data, out.ctypes.data, len(y))
    return out


def test_golden_primitives(oracle_mod, golden_dir):
    L = oracle_mod.lib()
    g = np.load(os.path.join(golden_dir, "prim_cv2.npz"))
    img = g["img"]
    prev = img
    for l in range(1, 5):
        ref = g["resize_l%d" % l]
        got = _resize(L, prev, ref.shape[1], ref.shape[0])
        assert np.array_equal(got, ref), "resize level %d" % l
        prev = ref
    b = np.empty((img.shape[0] + 38, img.shape[1] + 38), np.uint8)
    L.orbo_copy_make_border101(img.ctypes.data, img.shape[1], img.shape[0], b.ctypes.data, 19)
    assert np.array_equal(b, g["border19"])
    assert np.array_equal(_blur(L, img), g["blur"])
    for th in (20, 7):
        assert np.array_equal(_fast(L, img, th), g["fast_t%d" % th])
    got = _atan(L, g["atan_y"], g["atan_x"])
    assert np.array_equal(got.view(np.uint32), g["atan_deg"].view(np.uint32))


def test_golden_undistort(oracle_mod, golden_dir):
    """Frame::UndistortKeyPoints = cv::undistortPoints(pts, K, dist, R=I, P=K): restatement vs cv2 4.13 fixtures, bit for bit"""
    g = np.load(os.path.join(golden_dir, "undistort_cv2.npz"))
    ncases = sum(1 for k in g.files if k.startswith("cam"))
    assert ncases >= 4
    for c in range(ncases):
        fx, fy, cx, cy = (float(v) for v in g["cam%d" % c])
        d = g["dist%d" % c]
        pts, ref = g["pts%d" % c], g["res%d" % c]
        ox, oy = oracle_mod.undistort_points(pts[:, 0], pts[:, 1], fx, fy, cx, cy, np.array(d, np.float32))
        assert np.array_equal(ox.view(np.uint32), ref[:, 0].view(np.uint32)), "case %d x" % c
        assert np.array_equal(oy.view(np.uint32), ref[:, 1].view(np.uint32)), "case %d y" % c


def test_golden_sincosf(oracle_mod, golden_dir):
    """restated glibc sincosf == host libm fixture (and == the libm of the machine running the test)."""
    L = oracle_mod.lib()
    g = np.load(os.path.join(golden_dir, "sincosf_libm.npz"))
    ang = g["angle"]
    s = np.zeros_like(ang)
    c = np.zeros_like(ang)
    L.orbo_sincosf(ang.ctypes.data, s.ctypes.data, c.ctypes.data, len(ang))
    assert np.array_equal(s.view(np.uint32), g["sin"].view(np.uint32))
    assert np.array_equal(c.view(np.uint32), g["cos"].view(np.uint32))
    s2 = np.zeros_like(ang)
    c2 = np.zeros_like(ang)
    L.orbo_libm_sincosf(ang.ctypes.data, s2.ctypes.data, c2.ctypes.data, len(ang))
    # informational pin of the running host's libm (glibc FMA variant expected)
    assert np.array_equal(s2.view(np.uint32), s.view(np.uint32)) and np.array_equal(c2.view(np.uint32), c.view(np.uint32))


def test_sincosf_dense_sample_vs_libm(oracle_mod):
    L = oracle_mod.lib()
    bits = np.arange(0, 0x43B40000 + 1, 4099, dtype=np.uint32)  # every 4099th float in [0, 360]
    deg = bits.view(np.float32)
    factor = np.float32(3.1415926535897932384626433832795 / 180.0)
    ang = (deg * factor).astype(np.float32)
    s, c, s2, c2 = (np.zeros_like(ang) for _ in range(4))
    L.orbo_sincosf(ang.ctypes.data, s.ctypes.data, c.ctypes.data, len(ang))
    L.orbo_libm_sincosf(ang.ctypes.data, s2.ctypes.data, c2.ctypes.data, len(ang))
    assert np.array_equal(s.view(np.uint32), s2.view(np.uint32))
    assert np.array_equal(c.view(np.uint32), c2.view(np.uint32))


def test_cv_round_half_even(oracle_mod):
    L = oracle_mod.lib()
    for v, r in [(0.5, 0), (1.5, 2), (2.5, 2), (-0.5, 0), (-1.5, -2), (2.4999, 2), (1e6 + 0.5, 1000000)]:
        assert L.orbo_cv_round_f(v) == r


cv2 = pytest.importorskip("cv2") if os.environ.get("ORBX_SKIP_CV2") is None else None


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
@pytest.mark.parametrize("shape_name", ["tum", "euroc", "kitti"])
def test_resize_chain_vs_cv2(oracle_mod, shape_name):
    L = oracle_mod.lib()
    h, w = synth.SHAPES[shape_name]
    img = synth.gen_frame((h, w), 5)
    prev = img
    for l in range(1, 8):
        s = np.float32(1.0) / (np.float32(1.2) ** l)
        dw = int(np.rint(np.float32(w) * s))
        dh = int(np.rint(np.float32(h) * s))
        ref = cv2.resize(prev, (dw, dh), interpolation=cv2.INTER_LINEAR)
        assert np.array_equal(_resize(L, prev, dw, dh), ref)
        prev = ref


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
def test_resize_arbitrary_ratios_vs_cv2(oracle_mod):
    L = oracle_mod.lib()
    rng = np.random.default_rng(0)
    for _ in range(40):
        sw, sh = int(rng.integers(8, 300)), int(rng.integers(8, 300))
        dw, dh = int(rng.integers(4, 400)), int(rng.integers(4, 400))
        if sw == 2 * dw and sh == 2 * dh:
            continue  # cv2 switches to INTER_AREA for exact 2x decimation (out of config scope)
        img = rng.integers(0, 256, (sh, sw), dtype=np.uint8)
        ref = cv2.resize(img, (dw, dh), interpolation=cv2.INTER_LINEAR)
        assert np.array_equal(_resize(L, img, dw, dh), ref)


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
def test_blur_vs_cv2(oracle_mod):
    L = oracle_mod.lib()
    rng = np.random.default_rng(1)
    for t in range(12):
        w, h = int(rng.integers(7, 300)), int(rng.integers(7, 300))
        img = rng.integers(0, 256, (h, w), dtype=np.uint8) if t % 2 else synth.gen_frame((h, w), t)
        ref = cv2.GaussianBlur(img, (7, 7), 2, None, 2, cv2.BORDER_REFLECT_101)
        assert np.array_equal(_blur(L, img), ref)


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
def test_fast_vs_cv2(oracle_mod):
    L = oracle_mod.lib()
    rng = np.random.default_rng(2)
    nk = 0
    for t in range(60):
        w, h = int(rng.integers(7, 60)), int(rng.integers(7, 60))
        img = synth.gen_frame((h, w), t, n_rects=10, flat_regions=0) if t % 2 else rng.integers(0, 256, (h, w), dtype=np.uint8)
        for th in (20, 7, 0, 100):
            det = cv2.FastFeatureDetector_create(th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
            ref = np.array([(int(k.pt[0]), int(k.pt[1]), int(k.response)) for k in det.detect(img)], np.int32).reshape(-1, 3)
            got = _fast(L, img, th)
            assert np.array_equal(got, ref)
            nk += len(ref)
    assert nk > 1000


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
def test_undistort_vs_cv2(oracle_mod):
    rng = np.random.default_rng(21)
    for trial in range(6):
        fx, fy = 400 + 300 * rng.random(2)
        cx, cy = 320 + 40 * rng.standard_normal(), 240 + 30 * rng.standard_normal()
        d = np.array([0.3 * rng.standard_normal(), rng.standard_normal(), 0.01 * rng.standard_normal(),
                      0.01 * rng.standard_normal(), rng.standard_normal()][: 4 + trial % 2], np.float32)
        K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
        pts = (rng.random((5000, 1, 2)) * [640, 480]).astype(np.float32)
        ref = cv2.undistortPoints(pts, K, d, None, K).reshape(-1, 2)
        ox, oy = oracle_mod.undistort_points(pts[:, 0, 0], pts[:, 0, 1], float(K[0, 0]), float(K[1, 1]), float(K[0, 2]), float(K[1, 2]), d)
        assert np.array_equal(ox.view(np.uint32), ref[:, 0].view(np.uint32))
        assert np.array_equal(oy.view(np.uint32), ref[:, 1].view(np.uint32))


def test_atan2_vs_cv2(oracle_mod):
    L = oracle_mod.lib()
    rng = np.random.default_rng(3)
    y = rng.integers(-3000000, 3000000, 20000).astype(np.float32)
    x = rng.integers(-3000000, 3000000, 20000).astype(np.float32)
    ref = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], np.float32)
    assert np.array_equal(_atan(L, y, x).view(np.uint32), ref.view(np.uint32))


def test_optimised_fast_equals_plain_definition(oracle_mod):
    """fast9_16 (rolling rows + high-speed test, used by the timed baseline) == the plain transcription."""
    L = oracle_mod.lib()
    rng = np.random.default_rng(4)
    for t in range(40):
        w, h = int(rng.integers(7, 90)), int(rng.integers(7, 90))
        img = synth.gen_frame((h, w), t, n_rects=12, flat_regions=0) if t % 2 else rng.integers(0, 256, (h, w), dtype=np.uint8)
        for th in (20, 7, 0, 60):
            a = _fast(L, img, th)
            buf = np.zeros((img.size, 3), np.int32)
            n = L.orbo_fast9_16_simple(img.ctypes.data, w, h, img.strides[0], th, 1, buf.ctypes.data, img.size)
            assert np.array_equal(a, buf[:n])
