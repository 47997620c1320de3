"""CPU: the C-ABI library loads and exports every symbol include/orbx.h declares; the Python mirrors fail loudly
(no CPU fallback) when no CUDA device is present; host-only entry points work."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "orbx.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(orb[xm]_[a-z_0-9]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    from orbslam2commentedbyxcm_b200 import _lib

    L = _lib.load()
    names = _declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(L, n), "liborbx.so does not export %s" % n
        assert n in _lib.SIGNATURES, "no ctypes signature for %s" % n
    assert set(_lib.SIGNATURES) == set(names)


def test_host_only_descriptor_distance():
    from orbslam2commentedbyxcm_b200 import ORBmatcher

    a = np.zeros(32, np.uint8)
    b = np.full(32, 0xFF, np.uint8)
    assert ORBmatcher.DescriptorDistance(a, b) == 256
    b[5] = 0x0F
    assert ORBmatcher.DescriptorDistance(a, b) == 252


def test_no_cpu_fallback_without_device():
    import orbslam2commentedbyxcm_b200 as pkg

    L = pkg.load_library()
    if L.orbx_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(pkg.OrbxError):
        pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    with pytest.raises(pkg.OrbxError):
        pkg.ORBmatcher(0.6, True)


def test_product_does_not_import_oracle():
    pkg_dir = os.path.join(ROOT, "orbslam2commentedbyxcm_b200")
    for dp, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cc")):
                src = open(os.path.join(dp, f), errors="replace").read()
                assert "import oracle" not in src and "from oracle" not in src and '#include "../../oracle' not in src, f


def test_frame_grid_helper_matches_oracle():
    """matcher.frame_grid (host glue for orbm_frame_view) == the oracle's restatement of Frame::AssignFeaturesToGrid."""
    import oracle
    from orbslam2commentedbyxcm_b200.matcher import frame_grid

    rng = np.random.default_rng(0)
    x = (rng.random(3000) * 1300 - 20).astype(np.float32)
    y = (rng.random(3000) * 400 - 10).astype(np.float32)
    x[:50] = np.float32(19.390625) * np.arange(50, dtype=np.float32) + np.float32(9.6953125)  # exact .5 cells
    w_inv, h_inv = float(np.float32(64.0) / np.float32(1241.0)), float(np.float32(48.0) / np.float32(376.0))
    s1, i1 = frame_grid(x, y, 0.0, 0.0, w_inv, h_inv)
    s2, i2 = oracle.assign_grid(x, y, 0.0, 0.0, w_inv, h_inv)
    assert np.array_equal(s1, s2) and np.array_equal(i1, i2)
