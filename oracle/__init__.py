"""oracle/ -- TEST INFRASTRUCTURE (CPU oracle for the ORB front-end hot path).

ctypes loader over ``liborboracle.so`` (strict build, defines results), ``liborboracle_fast.so``
(-O3, timed CPU baseline only) and, when present, ``_ref/liborbref.so`` (the reference's own
``src/ORBextractor.cc`` compiled unmodified against ``oracle/shim``).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may import this
package.  The product package (``orbslam2commentedbyxcm_b200``) never does.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

KP_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4")]
)

_c = ctypes
_vp, _i, _f, _sz = _c.c_void_p, _c.c_int, _c.c_float, _c.c_size_t


def build(ref=True):
    """(Re)build the oracle libraries with oracle/Makefile. `ref` also builds _ref/ when /root/reference exists."""
    subprocess.check_call(["make", "-s", "-C", _HERE, "all"])
    if ref and os.path.exists("/root/reference/src/ORBextractor.cc"):
        subprocess.check_call(["make", "-s", "-C", _HERE, "ref"])


def _load(name):
    path = os.path.join(_HERE, name)
    if not os.path.exists(path):
        if name.startswith("_ref"):
            return None
        build(ref=False)
    return _c.CDLL(path)


def _bind_extractor(lib, prefix):
    g = lambda n: getattr(lib, prefix + n)
    g("create").restype = _vp
    g("create").argtypes = [_i, _f, _i, _i, _i]
    g("destroy").argtypes = [_vp]
    g("extract").restype = _i
    g("extract").argtypes = [_vp, _vp, _i, _i, _sz]
    g("get_keypoints").argtypes = [_vp, _vp]
    g("get_descriptors").argtypes = [_vp, _vp]
    g("get_levels").restype = _i
    g("get_levels").argtypes = [_vp]
    g("get_scale_factors").argtypes = [_vp, _vp]
    g("get_inv_level_sigma2").argtypes = [_vp, _vp]
    g("get_level_size").argtypes = [_vp, _i, _vp, _vp]
    g("get_pyramid").argtypes = [_vp, _i, _vp]


class _Libs:
    strict = None
    fast = None
    ref = None
    ref_fast = None


def _bind_ref(L):
    _bind_extractor(L, "orbref_")
    L.orbref_extract_batch.restype = _c.c_long
    L.orbref_extract_batch.argtypes = [_i, _f, _i, _i, _i, _vp, _i, _i, _i, _i, _vp]
    L.orbref_extract_batch_full.restype = _c.c_long
    L.orbref_extract_batch_full.argtypes = [_i, _f, _i, _i, _i, _vp, _i, _i, _i, _i, _i, _vp, _vp, _vp]


def lib(kind="strict"):
    if kind == "strict":
        if _Libs.strict is None:
            _Libs.strict = _load("liborboracle.so")
            _bind_all(_Libs.strict)
        return _Libs.strict
    if kind == "fast":
        if _Libs.fast is None:
            _Libs.fast = _load("liborboracle_fast.so")
            _bind_all(_Libs.fast)
        return _Libs.fast
    if kind == "ref":
        if _Libs.ref is None:
            _Libs.ref = _load("_ref/liborbref.so")
            if _Libs.ref is not None:
                _bind_ref(_Libs.ref)
        return _Libs.ref
    if kind == "ref_fast":
        if _Libs.ref_fast is None:
            _Libs.ref_fast = _load("_ref/liborbref_fast.so")
            if _Libs.ref_fast is not None:
                _bind_ref(_Libs.ref_fast)
        return _Libs.ref_fast
    raise ValueError(kind)


def _bind_all(L):
    _bind_extractor(L, "orbo_")
    L.orbo_get_inv_scale_factors.argtypes = [_vp, _vp]
    L.orbo_get_level_sigma2.argtypes = [_vp, _vp]
    L.orbo_get_quotas.argtypes = [_vp, _vp]
    L.orbo_get_umax.argtypes = [_vp, _vp]
    L.orbo_get_blurred.restype = _i
    L.orbo_get_blurred.argtypes = [_vp, _i, _vp]
    L.orbo_get_num_candidates.restype = _i
    L.orbo_get_num_candidates.argtypes = [_vp, _i]
    L.orbo_get_candidates.argtypes = [_vp, _i, _vp]
    L.orbo_get_num_level_keypoints.restype = _i
    L.orbo_get_num_level_keypoints.argtypes = [_vp, _i]
    L.orbo_distribute.restype = _i
    L.orbo_distribute.argtypes = [_vp, _i, _i, _i, _i, _i, _i, _vp]
    L.orbo_sincosf.argtypes = [_vp, _vp, _vp, _i]
    L.orbo_libm_sincosf.argtypes = [_vp, _vp, _vp, _i]
    L.orbo_extract_batch.restype = _c.c_long
    L.orbo_extract_batch.argtypes = [_i, _f, _i, _i, _i, _vp, _i, _i, _i, _i, _vp]
    # primitives
    L.orbo_resize_linear_u8.argtypes = [_vp, _i, _i, _sz, _vp, _i, _i, _sz]
    L.orbo_copy_make_border101.argtypes = [_vp, _i, _i, _vp, _i]
    L.orbo_fast9_16.restype = _i
    L.orbo_fast9_16.argtypes = [_vp, _i, _i, _sz, _i, _i, _vp, _i]
    L.orbo_fast9_16_simple.restype = _i
    L.orbo_fast9_16_simple.argtypes = [_vp, _i, _i, _sz, _i, _i, _vp, _i]
    L.orbo_gaussian_blur_7x7_s2.argtypes = [_vp, _i, _i, _sz, _vp, _sz]
    L.orbo_fast_atan2.argtypes = [_vp, _vp, _vp, _i]
    _u64 = _c.c_uint64
    L.orbo_sincosf_checksum.argtypes = [_c.c_uint32, _u64, _u64, _vp, _vp, _i]
    L.orbo_fast_atan2_checksum.argtypes = [_i, _u64, _u64, _u64, _vp, _i]
    L.orbo_sweep_sincosf_values.argtypes = [_c.c_uint32, _i, _vp, _vp]
    L.orbo_sweep_fast_atan2_values.argtypes = [_i, _u64, _i, _vp]
    L.orbo_cv_round_f.restype = _i
    L.orbo_cv_round_f.argtypes = [_f]
    # matcher
    L.orbo_descriptor_distance.restype = _i
    L.orbo_descriptor_distance.argtypes = [_vp, _vp]
    L.orbo_bruteforce.argtypes = [_vp, _i, _vp, _i, _vp, _vp, _vp, _i]
    L.orbo_assign_grid.argtypes = [_vp, _vp, _i, _f, _f, _f, _f, _vp, _vp]
    L.orbo_search_by_projection.restype = _i
    L.orbo_search_by_projection.argtypes = (
        [_i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _f, _f, _f, _f, _vp]
        + [_i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]
        + [_f, _f, _vp]
    )
    L.orbo_stereo_rows.argtypes = [_i, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _i, _f, _f, _vp, _vp]
    L.orbo_get_level_ptr.restype = _vp
    L.orbo_get_level_ptr.argtypes = [_vp, _i, _vp]
    L.orbo_stereo_match.restype = _i
    L.orbo_stereo_match.argtypes = [_i, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _f, _f, _i, _vp, _vp]
    L.orbo_search_window.restype = _i
    L.orbo_search_window.argtypes = ([_i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _f, _f, _f, _f, _vp]
                                     + [_i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp] + [_i, _i, _f, _i, _vp])
    L.orbo_distinctive_descriptors.argtypes = [_i, _vp, _vp, _vp]
    L.orbo_search_best.restype = None
    L.orbo_search_best.argtypes = ([_i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _f, _f, _f, _f]
                                   + [_i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _i, _vp, _vp])
    L.orbo_sim3_cross_check.restype = _i
    L.orbo_sim3_cross_check.argtypes = [_i, _vp, _i, _vp, _vp]
    L.orbo_search_for_initialization.restype = _i
    L.orbo_search_for_initialization.argtypes = ([_i, _vp, _vp, _vp, _vp, _vp] + [_i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _f, _f, _f, _f]
                                                 + [_i, _f, _i, _vp])
    L.orbo_search_for_triangulation.restype = _i
    L.orbo_search_for_triangulation.argtypes = ([_i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp]
                                                + [_i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp]
                                                + [_vp, _f, _f, _vp, _vp, _i, _i, _vp])
    L.orbo_search_by_bow_kf.restype = _i
    L.orbo_search_by_bow_kf.argtypes = [_i, _vp, _vp, _vp, _i, _vp, _vp, _vp] * 2 + [_f, _i, _vp]
    L.orbo_bow_transform.argtypes = [_i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp]
    L.orbo_search_by_bow.restype = _i
    L.orbo_search_by_bow.argtypes = (
        [_i, _vp, _vp, _vp, _i, _vp, _vp, _vp] + [_i, _vp, _vp, _i, _vp, _vp, _vp] + [_f, _i, _vp]
    )


def _ptr(a):
    return a.ctypes.data if a is not None else None


class Extractor:
    """CPU oracle extractor. kind: 'strict' | 'fast' (restatement) or 'ref' (reference source + shim)."""

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, kind="strict"):
        self.kind = kind
        self.L = lib(kind)
        if self.L is None:
            raise RuntimeError("oracle library for kind=%r is not built" % kind)
        self.p = "orbref_" if kind.startswith("ref") else "orbo_"
        self.h = self._f("create")(nfeatures, scale_factor, nlevels, ini_th, min_th)
        self.nlevels = nlevels
        self.n = 0

    def _f(self, name):
        return getattr(self.L, self.p + name)

    def close(self):
        if self.h:
            self._f("destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def extract(self, img):
        img = np.ascontiguousarray(img, dtype=np.uint8)
        self.n = self._f("extract")(self.h, img.ctypes.data, img.shape[0], img.shape[1], img.strides[0])
        kps = np.zeros(self.n, KP_DTYPE)
        desc = np.zeros((self.n, 32), np.uint8)
        if self.n:
            self._f("get_keypoints")(self.h, kps.ctypes.data)
            self._f("get_descriptors")(self.h, desc.ctypes.data)
        return kps, desc

    def scale_factors(self):
        out = np.zeros(self.nlevels, np.float32)
        self._f("get_scale_factors")(self.h, out.ctypes.data)
        return out

    def inv_level_sigma2(self):
        out = np.zeros(self.nlevels, np.float32)
        self._f("get_inv_level_sigma2")(self.h, out.ctypes.data)
        return out

    def level_size(self, level):
        w, h = _c.c_int(), _c.c_int()
        self._f("get_level_size")(self.h, level, _c.byref(w), _c.byref(h))
        return w.value, h.value

    def pyramid(self, level):
        w, h = self.level_size(level)
        out = np.zeros((h + 38, w + 38), np.uint8)
        self._f("get_pyramid")(self.h, level, out.ctypes.data)
        return out

    # --- restatement-only introspection ---
    def quotas(self):
        out = np.zeros(self.nlevels, np.int32)
        self.L.orbo_get_quotas(self.h, out.ctypes.data)
        return out

    def umax(self):
        out = np.zeros(16, np.int32)
        self.L.orbo_get_umax(self.h, out.ctypes.data)
        return out

    def blurred(self, level):
        w, h = self.level_size(level)
        out = np.zeros((h, w), np.uint8)
        ok = self.L.orbo_get_blurred(self.h, level, out.ctypes.data)
        return out if ok else None

    def candidates(self, level):
        n = self.L.orbo_get_num_candidates(self.h, level)
        out = np.zeros((n, 3), np.int32)
        if n:
            self.L.orbo_get_candidates(self.h, level, out.ctypes.data)
        return out

    def num_level_keypoints(self, level):
        return self.L.orbo_get_num_level_keypoints(self.h, level)


def distribute(xyr, min_x, max_x, min_y, max_y, n_wanted):
    """Quad-tree selection on candidates (n,3) int32 (x, y, response) relative to the grid origin -> picked indices."""
    L = lib()
    xyr = np.ascontiguousarray(xyr, np.int32)
    out = np.zeros(max(len(xyr), 1), np.int32)
    k = L.orbo_distribute(xyr.ctypes.data, len(xyr), min_x, max_x, min_y, max_y, n_wanted, out.ctypes.data)
    return out[:k].copy()


def extract_batch(imgs, nfeatures, scale_factor, nlevels, ini_th, min_th, threads, kind="fast"):
    """Frame-parallel CPU extraction (timed baseline).  kind: 'fast'/'strict' = restated oracle, 'ref_fast'/'ref' =
    the reference's own ORBextractor.cc (oracle/_ref)."""
    L = lib(kind)
    if L is None:
        raise RuntimeError("oracle library for kind=%r is not built" % kind)
    imgs = np.ascontiguousarray(imgs, np.uint8)
    n, rows, cols = imgs.shape
    counts = np.zeros(n, np.int32)
    fn = L.orbref_extract_batch if kind.startswith("ref") else L.orbo_extract_batch
    tot = fn(nfeatures, scale_factor, nlevels, ini_th, min_th, imgs.ctypes.data, n, rows, cols, threads, counts.ctypes.data)
    return tot, counts


def ref_extract_batch_full(imgs, nfeatures, scale_factor, nlevels, ini_th, min_th, cap, threads=0, kind="ref_fast"):
    """The reference's own ORBextractor.cc (oracle/_ref) over a batch, frame-parallel: (kps[n, cap] KP_DTYPE,
    desc[n, cap, 32] u8, counts[n]).  Returns None when oracle/_ref is not built."""
    L = lib(kind)
    if L is None:
        return None
    imgs = np.ascontiguousarray(imgs, np.uint8)
    n, rows, cols = imgs.shape
    kps = np.zeros((n, cap), KP_DTYPE)
    desc = np.zeros((n, cap, 32), np.uint8)
    counts = np.zeros(n, np.int32)
    L.orbref_extract_batch_full(nfeatures, scale_factor, nlevels, ini_th, min_th, imgs.ctypes.data, n, rows, cols,
                                threads or (os.cpu_count() or 1), cap, kps.ctypes.data, desc.ctypes.data, counts.ctypes.data)
    return kps, desc, counts


def bruteforce(q, db, threads=1, kind="strict"):
    L = lib(kind)
    q = np.ascontiguousarray(q, np.uint8)
    db = np.ascontiguousarray(db, np.uint8)
    nq, ndb = len(q), len(db)
    best = np.zeros(nq, np.int32)
    idx = np.zeros(nq, np.int32)
    second = np.zeros(nq, np.int32)
    L.orbo_bruteforce(q.ctypes.data, nq, db.ctypes.data, ndb, best.ctypes.data, idx.ctypes.data, second.ctypes.data, threads)
    return best, idx, second


def assign_grid(kx, ky, min_x, min_y, w_inv, h_inv):
    L = lib()
    kx = np.ascontiguousarray(kx, np.float32)
    ky = np.ascontiguousarray(ky, np.float32)
    start = np.zeros(64 * 48 + 1, np.int32)
    items = np.zeros(max(len(kx), 1), np.int32)
    L.orbo_assign_grid(kx.ctypes.data, ky.ctypes.data, len(kx), min_x, min_y, w_inv, h_inv, start.ctypes.data, items.ctypes.data)
    return start, items[: start[-1]].copy()


def undistort_points(x, y, fx, fy, cx, cy, dist):
    """Frame::UndistortKeyPoints (src/Frame.cc:586-628) = cv::undistortPoints(pts, K, dist, R=I, P=K), restated."""
    L = lib()
    x = np.ascontiguousarray(x, np.float32)
    y = np.ascontiguousarray(y, np.float32)
    dist = np.ascontiguousarray(dist, np.float32)
    ox, oy = np.empty_like(x), np.empty_like(y)
    L.orbo_undistort_points.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_float,
                                        ctypes.c_float, ctypes.c_float, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                        ctypes.c_void_p]
    L.orbo_undistort_points.restype = None
    L.orbo_undistort_points(x.ctypes.data, y.ctypes.data, len(x), fx, fy, cx, cy, dist.ctypes.data, len(dist),
                            ox.ctypes.data, oy.ctypes.data)
    return ox, oy


def search_by_projection(frame, mps, th, nnratio, kind="strict"):
    """frame / mps: dicts of numpy arrays (see orbslam2commentedbyxcm_b200.matcher for the field names)."""
    L = lib(kind)
    nkp = len(frame["x"])
    out = np.full(nkp, -1, np.int32)
    ur = frame.get("u_right")
    n = L.orbo_search_by_projection(
        nkp, _ptr(frame["x"]), _ptr(frame["y"]), _ptr(frame["octave"]), _ptr(ur), _ptr(frame["desc"]),
        _ptr(frame["occupied"]), _ptr(frame["cell_start"]), _ptr(frame["cell_items"]),
        frame["min_x"], frame["min_y"], frame["w_inv"], frame["h_inv"], _ptr(frame["scale_factors"]),
        len(mps["x"]), _ptr(mps["x"]), _ptr(mps["y"]), _ptr(mps["xr"]), _ptr(mps["level"]), _ptr(mps["view_cos"]),
        _ptr(mps["valid"]), _ptr(mps["obs"]), _ptr(mps["desc"]),
        th, nnratio, out.ctypes.data,
    )
    return n, out


def stereo_rows(left, right, scale_factors, n_rows, min_d, max_d):
    L = lib()
    nl = len(left["x"])
    best_r = np.zeros(nl, np.int32)
    best_d = np.zeros(nl, np.int32)
    L.orbo_stereo_rows(
        nl, _ptr(left["x"]), _ptr(left["y"]), _ptr(left["octave"]), _ptr(left["desc"]),
        len(right["x"]), _ptr(right["x"]), _ptr(right["y"]), _ptr(right["octave"]), _ptr(right["desc"]),
        _ptr(scale_factors), n_rows, min_d, max_d, best_r.ctypes.data, best_d.ctypes.data,
    )
    return best_r, best_d


def search_by_bow(kf, fr, nnratio, check_orientation):
    L = lib()
    nf = len(fr["angle"])
    out = np.full(nf, -1, np.int32)
    n = L.orbo_search_by_bow(
        len(kf["angle"]), _ptr(kf["desc"]), _ptr(kf["angle"]), _ptr(kf["has_mp"]),
        len(kf["node_id"]), _ptr(kf["node_id"]), _ptr(kf["node_start"]), _ptr(kf["node_feat"]),
        nf, _ptr(fr["desc"]), _ptr(fr["angle"]),
        len(fr["node_id"]), _ptr(fr["node_id"]), _ptr(fr["node_start"]), _ptr(fr["node_feat"]),
        nnratio, int(check_orientation), out.ctypes.data,
    )
    return n, out


def stereo_match(exL, exR, left, right, n_rows, mbf, mb, upstream_outliers=False):
    """Full ComputeStereoMatches on two oracle Extractor objects that have just extracted the left / right image.
    upstream_outliers=False: the SAD outlier block inside the keypoint loop as in this reference (src/Frame.cc:868-884);
    True: once after the loop as in upstream ORB-SLAM2."""
    L = lib()
    nl, nlev = len(left["x"]), exL.nlevels
    sf = exL.scale_factors()
    inv = np.zeros(nlev, np.float32)
    L.orbo_get_inv_scale_factors(exL.h, inv.ctypes.data)
    lp = (_c.c_void_p * nlev)()
    rp = (_c.c_void_p * nlev)()
    ls = np.zeros(nlev, np.int32)
    rs = np.zeros(nlev, np.int32)
    rc = np.zeros(nlev, np.int32)
    for l in range(nlev):
        st = _c.c_int()
        lp[l] = L.orbo_get_level_ptr(exL.h, l, _c.byref(st)); ls[l] = st.value
        rp[l] = L.orbo_get_level_ptr(exR.h, l, _c.byref(st)); rs[l] = st.value
        rc[l] = exR.level_size(l)[0]
    ur = np.zeros(max(nl, 1), np.float32)
    dp = np.zeros(max(nl, 1), np.float32)
    n = L.orbo_stereo_match(
        nl, _ptr(left["x"]), _ptr(left["y"]), _ptr(left["octave"]), _ptr(left["desc"]),
        len(right["x"]), _ptr(right["x"]), _ptr(right["y"]), _ptr(right["octave"]), _ptr(right["desc"]),
        nlev, sf.ctypes.data, inv.ctypes.data, lp, rp, ls.ctypes.data, rs.ctypes.data, rc.ctypes.data,
        n_rows, mbf, mb, 1 if upstream_outliers else 0, ur.ctypes.data, dp.ctypes.data)
    return n, ur[:nl], dp[:nl]


def bow_transform(vocab, desc, levelsup=4):
    """vocab: dict(L, child_start, child_ids, node_desc, word_id, weight) -> (word, node, weight) per descriptor."""
    L = lib()
    desc = np.ascontiguousarray(desc, np.uint8)
    n = len(desc)
    w = np.zeros(max(n, 1), np.int32)
    nd = np.zeros(max(n, 1), np.int32)
    wt = np.zeros(max(n, 1), np.float64)
    L.orbo_bow_transform(len(vocab["word_id"]), vocab["L"], _ptr(vocab["child_start"]), _ptr(vocab["child_ids"]),
                         _ptr(vocab["node_desc"]), _ptr(vocab["word_id"]), _ptr(vocab["weight"]), desc.ctypes.data, n, levelsup,
                         w.ctypes.data, nd.ctypes.data, wt.ctypes.data)
    return w[:n], nd[:n], wt[:n]


def search_window(frame, q, max_dist, ratio_mode, nnratio, check_orientation):
    """Generic windowed search (SearchByProjection(Cur, Last) / (Cur, KF, found) after projection); see match_oracle.cc."""
    L = lib()
    nkp = len(frame["x"])
    out = np.full(max(nkp, 1), -1, np.int32)
    n = L.orbo_search_window(
        nkp, _ptr(frame["x"]), _ptr(frame["y"]), _ptr(frame["octave"]), _ptr(frame.get("u_right")), _ptr(frame["desc"]),
        _ptr(frame["occupied"]), _ptr(frame["cell_start"]), _ptr(frame["cell_items"]),
        frame["min_x"], frame["min_y"], frame["w_inv"], frame["h_inv"], _ptr(frame.get("angle")),
        len(q["x"]), _ptr(q["x"]), _ptr(q["y"]), _ptr(q["radius"]), _ptr(q["min_level"]), _ptr(q["max_level"]), _ptr(q["xr"]),
        _ptr(q.get("angle")), _ptr(q["valid"]), _ptr(q["blocks"]), _ptr(q["desc"]),
        int(max_dist), int(ratio_mode), float(nnratio), int(check_orientation), out.ctypes.data)
    return n, out[:nkp]


def distinctive_descriptors(set_start, desc):
    L = lib()
    set_start = np.ascontiguousarray(set_start, np.int32)
    desc = np.ascontiguousarray(desc, np.uint8)
    out = np.zeros(max(len(set_start) - 1, 1), np.int32)
    L.orbo_distinctive_descriptors(len(set_start) - 1, set_start.ctypes.data, desc.ctypes.data, out.ctypes.data)
    return out[: len(set_start) - 1]


def search_by_bow_kf(kf1, kf2, nnratio, check_orientation):
    L = lib()
    n1 = len(kf1["angle"])
    out = np.full(max(n1, 1), -1, np.int32)
    args = []
    for kf in (kf1, kf2):
        args += [len(kf["angle"]), _ptr(kf["desc"]), _ptr(kf["angle"]), _ptr(kf["has_mp"]), len(kf["node_id"]), _ptr(kf["node_id"]),
                 _ptr(kf["node_start"]), _ptr(kf["node_feat"])]
    n = L.orbo_search_by_bow_kf(*args, nnratio, int(check_orientation), out.ctypes.data)
    return n, out[:n1]


def _frame_args(frame):
    return [len(frame["x"]), _ptr(frame["x"]), _ptr(frame["y"]), _ptr(frame["octave"]), _ptr(frame.get("u_right")), _ptr(frame["desc"]),
            _ptr(frame["cell_start"]), _ptr(frame["cell_items"]), frame["min_x"], frame["min_y"], frame["w_inv"], frame["h_inv"]]


def search_best(kf, q, gate, inv_level_sigma2, max_dist):
    """Claim-free windowed search (Fuse x2, SearchBySim3 passes); see match_oracle.cc."""
    L = lib()
    nq = len(q["x"])
    bi = np.full(max(nq, 1), -1, np.int32)
    bd = np.full(max(nq, 1), 256, np.int32)
    sig = np.ascontiguousarray(inv_level_sigma2, np.float32) if inv_level_sigma2 is not None else None
    L.orbo_search_best(*_frame_args(kf), nq, _ptr(q["x"]), _ptr(q["y"]), _ptr(q["radius"]), _ptr(q["min_level"]),
                       _ptr(q["max_level"]), _ptr(q.get("xr")), _ptr(q["valid"]), _ptr(q["desc"]), int(gate), _ptr(sig),
                       int(max_dist), bi.ctypes.data, bd.ctypes.data)
    return bi[:nq], bd[:nq]


def search_by_sim3(kf1, q2_in_1, kf2, q1_in_2):
    L = lib()
    m1, _ = search_best(kf2, q1_in_2, 0, None, 100)
    m2, _ = search_best(kf1, q2_in_1, 0, None, 100)
    m1 = np.ascontiguousarray(m1); m2 = np.ascontiguousarray(m2)
    out = np.full(max(len(m1), 1), -1, np.int32)
    n = L.orbo_sim3_cross_check(len(m1), m1.ctypes.data, len(m2), m2.ctypes.data, out.ctypes.data)
    return n, out[: len(m1)]


def search_for_initialization(f1, f2, prev_matched, window_size, nnratio, check_orientation):
    L = lib()
    n1 = len(f1["octave"])
    pm = np.ascontiguousarray(prev_matched, np.float32).reshape(-1, 2)
    px, py = np.ascontiguousarray(pm[:, 0]), np.ascontiguousarray(pm[:, 1])
    out = np.full(max(n1, 1), -1, np.int32)
    n = L.orbo_search_for_initialization(
        n1, _ptr(f1["octave"]), _ptr(f1["angle"]), _ptr(f1["desc"]), px.ctypes.data, py.ctypes.data,
        len(f2["x"]), _ptr(f2["x"]), _ptr(f2["y"]), _ptr(f2["octave"]), _ptr(f2["angle"]), _ptr(f2["desc"]),
        _ptr(f2["cell_start"]), _ptr(f2["cell_items"]), f2["min_x"], f2["min_y"], f2["w_inv"], f2["h_inv"],
        int(window_size), float(nnratio), int(check_orientation), out.ctypes.data)
    return n, out[:n1]


def search_for_triangulation(kf1, kf2, F12, ex, ey, scale_factors2, level_sigma2_2, only_stereo, check_orientation):
    L = lib()
    n1 = len(kf1["x"])
    out = np.full(max(n1, 1), -1, np.int32)
    F = np.ascontiguousarray(np.asarray(F12, np.float32).reshape(9))
    sf = np.ascontiguousarray(scale_factors2, np.float32)
    sg = np.ascontiguousarray(level_sigma2_2, np.float32)
    n = L.orbo_search_for_triangulation(
        n1, _ptr(kf1["x"]), _ptr(kf1["y"]), _ptr(kf1["angle"]), _ptr(kf1["u_right"]), _ptr(kf1["desc"]), _ptr(kf1["has_mp"]),
        len(kf1["node_id"]), _ptr(kf1["node_id"]), _ptr(kf1["node_start"]), _ptr(kf1["node_feat"]),
        len(kf2["x"]), _ptr(kf2["x"]), _ptr(kf2["y"]), _ptr(kf2["octave"]), _ptr(kf2["angle"]), _ptr(kf2["u_right"]), _ptr(kf2["desc"]),
        _ptr(kf2["has_mp"]), len(kf2["node_id"]), _ptr(kf2["node_id"]), _ptr(kf2["node_start"]), _ptr(kf2["node_feat"]),
        F.ctypes.data, float(ex), float(ey), sf.ctypes.data, sg.ctypes.data, int(only_stereo), int(check_orientation), out.ctypes.data)
    return n, out[:n1]


def sincosf_checksum(first_bits, count, chunk, threads=0, kind="fast"):
    """Checker side of orbx_debug_sincosf_checksum: per-chunk position-weighted checksums of glibc sinf / cosf of
    deg * factorPI over the float bit patterns first_bits .. first_bits + count - 1."""
    n = (count + chunk - 1) // chunk
    a = np.zeros(n, np.uint64)
    b = np.zeros(n, np.uint64)
    lib(kind).orbo_sincosf_checksum(first_bits, count, chunk, _ptr(a), _ptr(b), threads or (os.cpu_count() or 1))
    return a, b


def fast_atan2_checksum(mode, first, count, chunk, threads=0, kind="fast"):
    n = (count + chunk - 1) // chunk
    a = np.zeros(n, np.uint64)
    lib(kind).orbo_fast_atan2_checksum(mode, first, count, chunk, _ptr(a), threads or (os.cpu_count() or 1))
    return a


def sweep_sincosf_values(first_bits, n, kind="strict"):
    s = np.zeros(n, np.float32)
    c = np.zeros(n, np.float32)
    lib(kind).orbo_sweep_sincosf_values(first_bits, n, _ptr(s), _ptr(c))
    return s, c


def sweep_fast_atan2_values(mode, first, n, kind="strict"):
    a = np.zeros(n, np.float32)
    lib(kind).orbo_sweep_fast_atan2_values(mode, first, n, _ptr(a))
    return a
