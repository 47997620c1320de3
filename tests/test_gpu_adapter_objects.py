"""The C++ drop-in layer EXECUTED against the reference's real classes.

oracle/_ref/libmatchgpu.so = the reference's own src/Frame.cc, KeyFrame.cc, MapPoint.cc, Map.cc, KeyFrameDatabase.cc and
DBoW2, compiled unmodified, linked with the PRODUCT's orbslam2commentedbyxcm_b200/host/ORBmatcher_gpu.cc and
host/ORBextractor.cc (over liborbx.so) in place of src/ORBmatcher.cc and src/ORBextractor.cc -- the swap INTEGRATION.md
describes -- behind the same object-level harness (oracle/matchref_wrap.cc) that pins the oracle to the reference.  Every
ORBmatcher overload is called on real Frame / KeyFrame / MapPoint objects and what it did to them (mvpMapPoints, vpMatched,
vnMatches12, vbPrevMatched, vMatchedPairs, vpReplacePoint, AddObservation / Replace, mvuRight / mvDepth of a stereo Frame
built by the reference's constructor around the adapter extractor) must equal what the reference's own ORBmatcher.cc did
on the same inputs (tests/golden/matcher_ref.npz)."""
import os

import numpy as np
import pytest

import matcher_cases as mc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def adapter_outputs(oracle_mod):
    import orbslam2commentedbyxcm_b200 as pkg
    from oracle import matchref as R

    pkg.load_library()
    if not R.use("gpu"):
        R.use("ref")
        pytest.skip("oracle/_ref/libmatchgpu.so not present (built by `make -C oracle ref` where /root/reference exists)")
    try:
        assert R.lib().orbr_is_gpu_adapter() == 1
        out = mc.reference_outputs(oracle_mod, R)
    finally:
        R.use("ref")
    return out


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "matcher_ref.npz"))


def _keys():
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "matcher_ref.npz")
    return sorted(np.load(path).files)


@pytest.mark.parametrize("key", _keys())
def test_adapter_on_reference_objects_equals_reference(adapter_outputs, gold, key):
    got, ref = np.asarray(adapter_outputs[key]), gold[key]
    assert got.shape == ref.shape and got.dtype == ref.dtype, key
    if ref.dtype.kind == "f":
        u = np.uint32 if ref.dtype.itemsize == 4 else np.uint64
        assert np.array_equal(got.view(u), ref.view(u)), key
    else:
        assert np.array_equal(got, ref), key
