"""The C++ drop-in layer: host/ORBextractor.{h,cc} (same class surface as the reference's include/ORBextractor.h) and
host/ORBmatcher_gpu.cc.  CPU: both compile (against the mini-cv shim and the reference's own headers).  GPU: a C++ program that uses
the adapter class like Frame::ExtractORB does is compared with the reference's own class (oracle/_ref).  The matcher adapter is
EXECUTED on the reference's real objects in tests/test_gpu_adapter_objects.py."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPP = os.path.join(ROOT, "tests", "cpp")


def _make():
    from orbslam2commentedbyxcm_b200 import _build

    _build.build()
    subprocess.check_call(["make", "-s", "-C", CPP])


def test_adapters_compile():
    """host/ORBextractor.cc links into a program; host/ORBmatcher_gpu.cc compiles against the reference's REAL headers
    (include/ORBmatcher.h, Frame.h, KeyFrame.h, MapPoint.h + DBoW2 over the mini-cv shim) wherever the reference tree exists."""
    _make()
    assert os.path.exists(os.path.join(CPP, "build", "test_adapter"))
    if os.path.isdir("/root/reference/include"):
        assert os.path.exists(os.path.join(CPP, "build", "ORBmatcher_gpu.o"))


def test_every_overload_is_defined_by_the_adapter():
    """the adapter replaces src/ORBmatcher.cc as a whole: every member include/ORBmatcher.h declares has a definition."""
    src = open(os.path.join(ROOT, "orbslam2commentedbyxcm_b200", "host",
                            "ORBmatcher_gpu.cc")).read()
    for sig in ["ORBmatcher::ORBmatcher(float nnratio, bool checkOri)", "int ORBmatcher::DescriptorDistance(",
                "int ORBmatcher::SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th)",
                "int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono)",
                "int ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>& sAlreadyFound",
                "int ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw,", "int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F,",
                "int ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2,", "int ORBmatcher::SearchForInitialization(",
                "int ORBmatcher::SearchForTriangulation(", "int ORBmatcher::SearchBySim3(",
                "int ORBmatcher::Fuse(KeyFrame* pKF, const vector<MapPoint*>& vpMapPoints, const float th)",
                "int ORBmatcher::Fuse(KeyFrame* pKF, cv::Mat Scw,", "bool ORBmatcher::CheckDistEpipolarLine(",
                "float ORBmatcher::RadiusByViewingCos(", "void ORBmatcher::ComputeThreeMaxima("]:
        assert sig in src, sig


def test_adapter_header_keeps_reference_surface():
    """every public name of the reference's ORBextractor class is declared by the drop-in header."""
    hdr = open(os.path.join(ROOT, "orbslam2commentedbyxcm_b200", "host", "ORBextractor.h")).read()
    for name in ["namespace ORB_SLAM2", "class ExtractorNode", "class ORBextractor", "HARRIS_SCORE=0".replace("=", " = "),
                 "ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST)",
                 "void operator()(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint>& keypoints",
                 "GetLevels()", "GetScaleFactor()", "GetScaleFactors()", "GetInverseScaleFactors()", "GetScaleSigmaSquares()",
                 "GetInverseScaleSigmaSquares()", "std::vector<cv::Mat> mvImagePyramid"]:
        assert name in hdr, name


@pytest.mark.gpu
def test_adapter_equals_reference_class_on_gpu():
    ref = os.path.join(ROOT, "oracle", "_ref", "liborbref.so")
    if not os.path.exists(ref):
        pytest.skip("oracle/_ref/liborbref.so not present")
    _make()
    out = subprocess.run([os.path.join(CPP, "build", "test_adapter"), ref], capture_output=True, text=True, cwd=ROOT)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "0 mismatches" in out.stdout
