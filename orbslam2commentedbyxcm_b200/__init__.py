"""orbslam2commentedbyxcm_b200 -- B200-native (sm_100a) ORB-SLAM2 feature front-end.

Hand-written CUDA kernels behind a C-ABI (include/orbx.h, built into liborbx.so) with host-side mirrors of the
reference's ORBextractor / ORBmatcher interfaces:

  C++ : host/ORBextractor.{h,cc} (drop-in class, replaces include/ORBextractor.h + src/ORBextractor.cc) and
        host/ORBmatcher_gpu.cc (every member of the reference's own include/ORBmatcher.h, replaces src/ORBmatcher.cc)
  Py  : extractor.ORBextractor, matcher.ORBmatcher (used by tests/ and bench.py)

There is no CPU fallback: compute entry points raise OrbxError when liborbx.so or a CUDA device is missing.
"""
from ._lib import KP_DTYPE, OrbxError, load as load_library  # noqa: F401
from .extractor import ORBextractor  # noqa: F401
from .matcher import ORBmatcher  # noqa: F401
