"""CPU side of the transcendental domain sweeps (tests/test_gpu_transcendentals.py): the oracle's checksum functions agree
with a direct evaluation, between the strict and the fast build, and with the host libm / cv2 on samples."""
import numpy as np


def test_checksum_detects_a_single_ulp(oracle_mod):
    """The checksum is position weighted with odd weights: flipping one bit of one value changes it."""
    s, c = oracle_mod.sweep_sincosf_values(0x42000000, 1024, kind="strict")
    w = (2 * np.arange(1024, dtype=np.uint64) + 1)
    ref = (s.view(np.uint32).astype(np.uint64) * w).sum(dtype=np.uint64)
    a, _ = oracle_mod.sincosf_checksum(0x42000000, 1024, 1024, kind="strict")
    assert a[0] == ref
    s2 = s.copy().view(np.uint32)
    s2[517] ^= 1
    assert (s2.astype(np.uint64) * w).sum(dtype=np.uint64) != ref


def test_fast_and_strict_builds_agree(oracle_mod):
    a, b = oracle_mod.sincosf_checksum(0x3F000000, 1 << 18, 1 << 14, kind="strict")
    c, d = oracle_mod.sincosf_checksum(0x3F000000, 1 << 18, 1 << 14, kind="fast")
    assert np.array_equal(a, c) and np.array_equal(b, d)
    for mode in (0, 1):
        first = 5000 * 10001 + 4000 if mode == 0 else 0
        x = oracle_mod.fast_atan2_checksum(mode, first, 1 << 18, 1 << 14, kind="strict")
        y = oracle_mod.fast_atan2_checksum(mode, first, 1 << 18, 1 << 14, kind="fast")
        assert np.array_equal(x, y)


def test_sweep_values_match_libm_and_cv2(oracle_mod):
    first = 0x42B40000 - 3000
    s, c = oracle_mod.sweep_sincosf_values(first, 6000, kind="strict")
    deg = (first + np.arange(6000, dtype=np.uint32)).view(np.float32)
    ang = (deg * np.float32(np.pi / np.float32(180.0))).astype(np.float32)
    ls = np.zeros(6000, np.float32)
    lc = np.zeros(6000, np.float32)
    oracle_mod.lib("strict").orbo_libm_sincosf(ang.ctypes.data, ls.ctypes.data, lc.ctypes.data, 6000)
    assert np.array_equal(s.view(np.uint32), ls.view(np.uint32)) and np.array_equal(c.view(np.uint32), lc.view(np.uint32))
    cv2 = __import__("pytest").importorskip("cv2")
    W = 10001
    first = 5000 * W + 4000
    a = oracle_mod.sweep_fast_atan2_values(0, first, 2000, kind="strict")
    idx = first + np.arange(2000)
    y = (idx // W - 5000).astype(np.float32)
    x = (idx % W - 5000).astype(np.float32)
    ref = np.array([cv2.fastAtan2(float(yy), float(xx)) for yy, xx in zip(y, x)], np.float32)
    assert np.array_equal(a.view(np.uint32), ref.view(np.uint32))
