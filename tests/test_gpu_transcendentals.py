"""GPU parity over the whole DOMAIN of the two transcendental device functions the descriptor path depends on:

* csrc/sincosf_dev.cuh `sincosf_glibc` (the rotation of computeOrbDescriptor, src/ORBextractor.cc:109,123-125) against
  oracle/sincosf_glibc.h -- which oracle/_ref/sweep_sincosf pins against the host libm on the same set -- for ALL
  1 135 869 953 float angles deg in [0, 360] (bit patterns 0x00000000 .. 0x43b40000), angle = deg * factorPI;
* `fast_atan2_dev` (IC_Angle, src/ORBextractor.cc:105) against oracle/cvprim.h `fast_atan2` (pinned against
  cv2.fastAtan2) on the dense grid of integer moments [-5000, 5000]^2 (1.0002e8 pairs, every axis / zero special), and on
  1.2e8 pseudo-random pairs of signed 22-bit moments.

Bit-exact: both sides fold each chunk of consecutive values into a position-weighted 64-bit checksum (include/orbx.h);
a differing chunk is then re-run value by value so the failure names the first differing input."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

LAST_DEG_BITS = 0x43B40000  # 360.0f
CHUNK = 1 << 20


@pytest.fixture(scope="module")
def L():
    from orbslam2commentedbyxcm_b200 import _lib

    return _lib


def _gpu_sincosf_checksum(L, first, count, chunk):
    n = (count + chunk - 1) // chunk
    a = np.zeros(n, np.uint64)
    b = np.zeros(n, np.uint64)
    L.check(L.load().orbx_debug_sincosf_checksum(0, first, count, chunk, a.ctypes.data, b.ctypes.data), "sincosf sweep")
    return a, b


def _gpu_atan2_checksum(L, mode, first, count, chunk):
    n = (count + chunk - 1) // chunk
    a = np.zeros(n, np.uint64)
    L.check(L.load().orbx_debug_fast_atan2_checksum(0, mode, first, count, chunk, a.ctypes.data), "fastAtan2 sweep")
    return a


def _first_sincosf_difference(L, oracle_mod, chunk_index):
    first = chunk_index * CHUNK
    n = min(CHUNK, LAST_DEG_BITS + 1 - first)
    s = np.zeros(n, np.float32)
    c = np.zeros(n, np.float32)
    L.check(L.load().orbx_debug_sincosf(0, first, n, s.ctypes.data, c.ctypes.data), "sincosf values")
    os_, oc = oracle_mod.sweep_sincosf_values(first, n, kind="fast")
    bad = np.nonzero((s.view(np.uint32) != os_.view(np.uint32)) | (c.view(np.uint32) != oc.view(np.uint32)))[0]
    return [(hex(first + int(i)), float(s[i]), float(os_[i]), float(c[i]), float(oc[i])) for i in bad[:5]], len(bad)


def test_sincosf_values_small_ranges(L, oracle_mod):
    """Value-by-value on a few ranges (zero, the 2^-12 and pi/4 branch points, quadrant borders, the top of the range),
    strict oracle build."""
    factor_pi = np.float32(np.pi / np.float32(180.0))
    starts = [0, 1, 0x3F800000 - 512, LAST_DEG_BITS - 4095]
    for deg in (45.0, 90.0, 135.0, 180.0, 225.0, 270.0, 315.0, 0.0139, 2.0 ** -12 / float(factor_pi)):
        starts.append(int(np.float32(deg).view(np.uint32)) - 2048)
    for first in starts:
        first = max(0, first)
        n = min(4096, LAST_DEG_BITS + 1 - first)
        s = np.zeros(n, np.float32)
        c = np.zeros(n, np.float32)
        L.check(L.load().orbx_debug_sincosf(0, first, n, s.ctypes.data, c.ctypes.data), "sincosf values")
        os_, oc = oracle_mod.sweep_sincosf_values(first, n, kind="strict")
        assert np.array_equal(s.view(np.uint32), os_.view(np.uint32)), hex(first)
        assert np.array_equal(c.view(np.uint32), oc.view(np.uint32)), hex(first)


def test_sincosf_exhaustive_all_angles(L, oracle_mod):
    count = LAST_DEG_BITS + 1
    assert count == 1135869953
    gs, gc = _gpu_sincosf_checksum(L, 0, count, CHUNK)
    os_, oc = oracle_mod.sincosf_checksum(0, count, CHUNK, kind="fast")
    bad = np.nonzero((gs != os_) | (gc != oc))[0]
    if len(bad):
        detail, nbad = _first_sincosf_difference(L, oracle_mod, int(bad[0]))
        pytest.fail("%d of %d chunks differ; first chunk %d has %d differing angles, e.g. (bits, sin gpu/cpu, cos gpu/cpu) %s"
                    % (len(bad), len(gs), int(bad[0]), nbad, detail))


def test_fast_atan2_values_specials(L, oracle_mod):
    """Axis and zero specials + a dense block around the origin, value by value (strict build)."""
    W = 10001
    rows = [5000 * W + 5000 - 50, 0, W * W - 4096, 4999 * W, 5001 * W]
    for first in rows:
        n = 4096
        a = np.zeros(n, np.float32)
        L.check(L.load().orbx_debug_fast_atan2(0, 0, first, n, a.ctypes.data), "fastAtan2 values")
        o = oracle_mod.sweep_fast_atan2_values(0, first, n, kind="strict")
        assert np.array_equal(a.view(np.uint32), o.view(np.uint32)), first
    # (m01, m10) = (0,0) -> 0; (0,+) -> 0; (+,0) -> 90; (0,-) -> 180; (-,0) -> 270
    centre = 5000 * W + 5000
    got = {}
    for name, idx in (("00", centre), ("0+", centre + 7), ("+0", centre + 7 * W), ("0-", centre - 7), ("-0", centre - 7 * W)):
        a = np.zeros(1, np.float32)
        L.check(L.load().orbx_debug_fast_atan2(0, 0, idx, 1, a.ctypes.data), "fastAtan2 values")
        got[name] = float(a[0])
    assert got == {"00": 0.0, "0+": 0.0, "+0": 90.0, "0-": 180.0, "-0": 270.0}


def test_fast_atan2_dense_grid_and_random_pairs(L, oracle_mod):
    total = 10001 * 10001
    g = _gpu_atan2_checksum(L, 0, 0, total, CHUNK)
    o = oracle_mod.fast_atan2_checksum(0, 0, total, CHUNK, kind="fast")
    assert np.array_equal(g, o), "dense grid: chunks %s differ" % np.nonzero(g != o)[0][:8]
    for mode in (1, 2, 3):
        n = 40_000_000
        g = _gpu_atan2_checksum(L, mode, 0, n, CHUNK)
        o = oracle_mod.fast_atan2_checksum(mode, 0, n, CHUNK, kind="fast")
        assert np.array_equal(g, o), "random pairs mode %d: chunks %s differ" % (mode, np.nonzero(g != o)[0][:8])
