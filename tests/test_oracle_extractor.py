"""CPU: the restated oracle extractor against (a) the committed outputs of the reference's own ORBextractor.cc
(tests/golden/extract_ref_*.npz) and (b) the reference compiled here (oracle/_ref/liborbref.so) when present."""
import os

import numpy as np
import pytest

from orbslam2commentedbyxcm_b200 import synth


@pytest.mark.parametrize("name", ["small_a", "small_b", "tum_s0"])
def test_oracle_matches_reference_golden(oracle_mod, golden_dir, name):
    g = np.load(os.path.join(golden_dir, "extract_ref_%s.npz" % name))
    shape = tuple(int(v) for v in g["shape"])
    cfg = g["cfg"]
    img = synth.gen_frame(shape, int(g["seed"]))
    ex = oracle_mod.Extractor(int(cfg[0]), float(cfg[1]), int(cfg[2]), int(cfg[3]), int(cfg[4]), kind="strict")
    kps, desc = ex.extract(img)
    assert kps.tobytes() == g["kps"].tobytes()
    assert np.array_equal(desc, g["desc"])


def test_fast_build_identical(oracle_mod):
    img = synth.gen_frame((240, 320), 9)
    a = oracle_mod.Extractor(500, 1.2, 8, 20, 7, kind="strict").extract(img)
    b = oracle_mod.Extractor(500, 1.2, 8, 20, 7, kind="fast").extract(img)
    assert a[0].tobytes() == b[0].tobytes() and np.array_equal(a[1], b[1])


@pytest.mark.parametrize("shape,seed,cfg", [
    ((480, 752), 1, (1200, 1.2, 8, 20, 7)),
    ((376, 1241), 2, (2000, 1.2, 8, 20, 7)),
    ((130, 170), 3, (2000, 1.2, 3, 20, 7)),     # quota never reached: tree runs to exhaustion
    ((240, 320), 4, (50, 1.3, 5, 30, 10)),      # tiny quota, other scale/thresholds
])
def test_oracle_vs_compiled_reference(oracle_mod, shape, seed, cfg):
    if oracle_mod.lib("ref") is None:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    img = synth.gen_frame(shape, seed)
    o = oracle_mod.Extractor(*cfg, kind="strict")
    r = oracle_mod.Extractor(*cfg, kind="ref")
    k1, d1 = o.extract(img)
    k2, d2 = r.extract(img)
    assert len(k1) == len(k2) and k1.tobytes() == k2.tobytes()
    assert np.array_equal(d1, d2)
    for l in range(cfg[2]):
        assert np.array_equal(o.pyramid(l), r.pyramid(l))
    assert np.array_equal(o.scale_factors(), r.scale_factors())
    assert np.array_equal(o.inv_level_sigma2(), r.inv_level_sigma2())


def test_tables(oracle_mod):
    o = oracle_mod.Extractor(1000, 1.2, 8, 20, 7)
    assert o.quotas().tolist() == [217, 181, 151, 126, 105, 87, 73, 60]
    assert o.umax().tolist() == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    assert oracle_mod.Extractor(2000, 1.2, 8, 20, 7).quotas().tolist() == [434, 362, 302, 251, 209, 175, 145, 122]
    img = synth.gen_frame((480, 640), 0)
    o.extract(img)
    assert [o.level_size(l) for l in range(8)] == [(640, 480), (533, 400), (444, 333), (370, 278), (309, 231), (257, 193), (214, 161), (179, 134)]


def test_retry_threshold_fires(oracle_mod):
    """the synthetic frames must exercise the minThFAST retry (some kept candidates have response < iniThFAST)."""
    img = synth.gen_frame((480, 640), 0)
    o = oracle_mod.Extractor(1000, 1.2, 8, 20, 7)
    o.extract(img)
    c = o.candidates(0)
    assert (c[:, 2] < 20).any() and (c[:, 2] >= 20).any()
    assert len(c) > 4 * 217
