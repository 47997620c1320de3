"""Python mirror of ORB_SLAM2::ORBextractor (reference include/ORBextractor.h:76-220) over the C-ABI.

Same constructor arguments, call operator, accessors and ``mvImagePyramid`` as the reference class; used by the
parity tests and bench.py.  The drop-in for C++ callers is host/ORBextractor.h.  All compute happens in liborbx.so
on the GPU; nothing here falls back to the CPU.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, NUM_STAGES, STAGE_NAMES, OrbxError, check


class ORBextractor:
    HARRIS_SCORE = 0
    FAST_SCORE = 1

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device=0):
        self._L = _lib.load()
        self.nfeatures, self.scaleFactor, self.nlevels = int(nfeatures), float(scaleFactor), int(nlevels)
        self.iniThFAST, self.minThFAST = int(iniThFAST), int(minThFAST)
        cfg = _lib.Config(self.nfeatures, self.scaleFactor, self.nlevels, self.iniThFAST, self.minThFAST)
        h = C.c_void_p()
        check(self._L.orbx_create(C.byref(cfg), int(device), C.byref(h)), "orbx_create")
        self._h = h
        nl = self.nlevels
        self._scale = np.zeros(nl, np.float32)
        self._inv_scale = np.zeros(nl, np.float32)
        self._sigma2 = np.zeros(nl, np.float32)
        self._inv_sigma2 = np.zeros(nl, np.float32)
        self._quota = np.zeros(nl, np.int32)
        check(self._L.orbx_get_tables(h, self._scale.ctypes.data, self._inv_scale.ctypes.data, self._sigma2.ctypes.data,
                                      self._inv_sigma2.ctypes.data, self._quota.ctypes.data), "orbx_get_tables")
        self._last_n = 0
        self._pinned = []

    def close(self):
        if getattr(self, "_h", None):
            self._L.orbx_destroy(self._h)
            self._h = None
            for p in self._pinned:
                self._L.orbx_free_pinned(p)
            self._pinned = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- accessors (include/ORBextractor.h:119-159) ----
    def GetLevels(self):
        return self.nlevels

    def GetScaleFactor(self):
        return self.scaleFactor

    def GetScaleFactors(self):
        return self._scale.copy()

    def GetInverseScaleFactors(self):
        return self._inv_scale.copy()

    def GetScaleSigmaSquares(self):
        return self._sigma2.copy()

    def GetInverseScaleSigmaSquares(self):
        return self._inv_sigma2.copy()

    @property
    def mnFeaturesPerLevel(self):
        return self._quota.copy()

    def max_keypoints(self, rows, cols):
        r = self._L.orbx_max_keypoints(self._h, int(rows), int(cols))
        if r < 0:
            check(r, "orbx_max_keypoints")
        return r

    # ---- operator() ----
    def __call__(self, image, mask=None):
        """(keypoints, descriptors) for one CV_8UC1 image; an empty image returns empty outputs like the reference's
        silent return (src/ORBextractor.cc:1517-1518).  `mask` is ignored, as in the reference."""
        if image is None or image.size == 0:
            return np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        assert image.dtype == np.uint8 and image.ndim == 2, "image must be CV_8UC1"
        k, d, c = self.extract_batch(image[None])
        return k[0, : c[0]].copy(), d[0, : c[0]].copy()

    def alloc_outputs(self, n, rows, cols, pinned=False):
        """Output arrays for extract_batch.  pinned=True puts them in page-locked memory (orbx_alloc_pinned) so that results
        are DMA'd straight into them; the arrays stay valid until this extractor object is closed."""
        cap = self.max_keypoints(rows, cols)
        if not pinned:
            return np.zeros((n, cap), KP_DTYPE), np.zeros((n, cap, 32), np.uint8), np.zeros(n, np.int32)
        return (self._pinned_array((n, cap), KP_DTYPE), self._pinned_array((n, cap, 32), np.uint8), self._pinned_array((n,), np.int32))

    def _pinned_array(self, shape, dtype):
        nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = self._L.orbx_alloc_pinned(nbytes)
        if not p:
            check(-3, "orbx_alloc_pinned")
        self._pinned.append(p)
        buf = (C.c_uint8 * nbytes).from_address(p)
        a = np.frombuffer(buf, dtype=dtype).reshape(shape)
        a[...] = np.zeros((), dtype)
        return a

    def extract_batch(self, images, out=None):
        """images: (n, rows, cols) uint8 host array.  Returns (kps[n, cap], desc[n, cap, 32], counts[n]); pass
        out=alloc_outputs(...) to reuse host buffers across calls."""
        images = np.ascontiguousarray(images, np.uint8)
        n, rows, cols = images.shape
        kps, desc, counts = out if out is not None else self.alloc_outputs(n, rows, cols)
        cap = kps.shape[1]
        check(self._L.orbx_extract_batch(self._h, images.ctypes.data, n, rows, cols, cols, rows * cols, kps.ctypes.data,
                                         desc.ctypes.data, cap, counts.ctypes.data), "orbx_extract_batch")
        self._last_n = n
        return kps, desc, counts

    def extract_batch_device(self, d_imgs_ptr, n, rows, cols, d_kps_ptr, d_desc_ptr, d_counts_ptr, capacity,
                             step=None, frame_stride=None):
        """Device-resident variant (raw device pointers as ints), asynchronous on self.stream."""
        step = cols if step is None else step
        frame_stride = rows * cols if frame_stride is None else frame_stride
        check(self._L.orbx_extract_batch_device(self._h, d_imgs_ptr, n, rows, cols, step, frame_stride, d_kps_ptr,
                                                d_desc_ptr, capacity, d_counts_ptr), "orbx_extract_batch_device")
        self._last_n = n

    def sync(self):
        check(self._L.orbx_sync(self._h), "orbx_sync")

    @property
    def stream(self):
        return self._L.orbx_stream(self._h)

    # ---- mvImagePyramid (include/ORBextractor.h:162) ----
    def level_size(self, level):
        r, c = C.c_int(), C.c_int()
        check(self._L.orbx_get_level_size(self._h, level, C.byref(r), C.byref(c)), "orbx_get_level_size")
        return r.value, c.value

    def pyramid_level_padded(self, level, frame=0):
        rows, cols = self.level_size(level)
        out = np.zeros((rows + 38, cols + 38), np.uint8)
        check(self._L.orbx_get_pyramid_level(self._h, frame, level, out.ctypes.data, out.strides[0]), "orbx_get_pyramid_level")
        return out

    @property
    def mvImagePyramid(self):
        """List of level images of the last extracted frame 0: views at (19,19) of padded buffers, like the reference."""
        return [self.pyramid_level_padded(l)[19:-19, 19:-19] for l in range(self.nlevels)]

    # ---- instrumentation ----
    def set_profiling(self, on):
        check(self._L.orbx_set_profiling(self._h, int(bool(on))), "orbx_set_profiling")

    def set_use_graph(self, on):
        check(self._L.orbx_set_use_graph(self._h, int(bool(on))), "orbx_set_use_graph")

    def stage_times(self):
        ms = np.zeros(NUM_STAGES, np.float32)
        check(self._L.orbx_stage_times(self._h, ms.ctypes.data), "orbx_stage_times")
        return dict(zip(STAGE_NAMES, ms.tolist()))

    def last_launch_count(self):
        return self._L.orbx_last_launch_count(self._h)

    def debug_candidates(self, level, frame=0):
        n = C.c_int()
        check(self._L.orbx_debug_get_candidates(self._h, frame, level, None, 0, C.byref(n)), "orbx_debug_get_candidates")
        out = np.zeros((max(n.value, 1), 3), np.int32)
        check(self._L.orbx_debug_get_candidates(self._h, frame, level, out.ctypes.data, n.value, C.byref(n)), "orbx_debug_get_candidates")
        return out[: n.value]

    def debug_blurred(self, level, frame=0):
        rows, cols = self.level_size(level)
        out = np.zeros((rows, cols), np.uint8)
        check(self._L.orbx_debug_get_blurred(self._h, frame, level, out.ctypes.data), "orbx_debug_get_blurred")
        return out
