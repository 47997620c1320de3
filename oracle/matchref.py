"""oracle/matchref.py -- TEST INFRASTRUCTURE.

ctypes loader over ``oracle/_ref/libmatchref.so``: the reference's OWN ``src/ORBmatcher.cc``, ``src/Frame.cc``,
``src/KeyFrame.cc``, ``src/MapPoint.cc``, ``src/Map.cc`` and the vendored DBoW2, compiled unmodified (oracle/Makefile
target ``ref``) behind ``oracle/matchref_wrap.cc``.  The functions take the same flat dicts as the oracle wrappers in
``oracle/__init__.py`` and the GPU mirror in ``orbslam2commentedbyxcm_b200/matcher.py`` so that tests can call all three
with one set of inputs.

Only ``tests/`` and ``tools/make_golden.py`` import this module.  ``available()`` is False on the GPU box unless the built
library travelled with the snapshot (it does: oracle/_ref is git-ignored but not gpurun-ignored).
"""
import ctypes as C
import os
import tempfile

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATHS = {"ref": os.path.join(_HERE, "_ref", "libmatchref.so"),     # the reference's ORBmatcher.cc / ORBextractor.cc
          "gpu": os.path.join(_HERE, "_ref", "libmatchgpu.so"),     # the product's host/ORBmatcher_gpu.cc / ORBextractor.cc instead
          "ref_fast": os.path.join(_HERE, "_ref", "libmatchref_fast.so")}   # "ref" at -O3 -march=x86-64-v3: timed baselines (bench.py)
_PATH = _PATHS["ref"]
_lib = None
_libs = {}


def use(variant):
    """Select which build of the object-level harness the functions below call: "ref" (default) or "gpu" (the product's C++
    drop-in layer linked with the reference's own Frame / KeyFrame / MapPoint / Map / DBoW2; needs a CUDA device)."""
    global _PATH, _lib
    _PATH = _PATHS[variant]
    _lib = _libs.get(variant)
    return available()


_vp, _i, _f = C.c_void_p, C.c_int, C.c_float

# inverse depths the harness accepts (powers of two keep the reference's projection arithmetic exact)
INVZ_CHOICES = np.array([0.125, 0.25, 0.5, 1.0, 2.0], np.float32)


class FrameIn(C.Structure):
    _fields_ = [("n", C.c_int32), ("x", _vp), ("y", _vp), ("octave", _vp), ("angle", _vp), ("u_right", _vp), ("desc", _vp),
                ("occupied", _vp), ("min_x", _f), ("min_y", _f), ("max_x", _f), ("max_y", _f), ("w_inv", _f), ("h_inv", _f),
                ("scale_factors", _vp), ("nlevels", C.c_int32), ("mbf", _f), ("n_nodes", C.c_int32), ("node_id", _vp),
                ("node_start", _vp), ("node_feat", _vp), ("has_mp", _vp)]


class QueryIn(C.Structure):
    _fields_ = [("n", C.c_int32), ("x", _vp), ("y", _vp), ("invz", _vp), ("level", _vp), ("angle", _vp), ("valid", _vp),
                ("blocks", _vp), ("desc", _vp)]


def available():
    return os.path.exists(_PATH)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("%s is not built (make -C oracle ref; needs /root/reference)" % _PATH)
        L = C.CDLL(_PATH)
        P = C.POINTER
        L.orbr_assign_grid.argtypes = [P(FrameIn), _vp, _vp]
        L.orbr_features_in_area.restype = _i
        L.orbr_features_in_area.argtypes = [P(FrameIn), _i, _f, _f, _f, _i, _i, _vp]
        L.orbr_descriptor_distance.restype = _i
        L.orbr_descriptor_distance.argtypes = [_vp, _vp]
        L.orbr_search_by_projection.restype = _i
        L.orbr_search_by_projection.argtypes = [P(FrameIn), _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _f, _f, _vp]
        L.orbr_search_by_bow.restype = _i
        L.orbr_search_by_bow.argtypes = [P(FrameIn), P(FrameIn), _f, _i, _vp]
        L.orbr_search_by_bow_kf.restype = _i
        L.orbr_search_by_bow_kf.argtypes = [P(FrameIn), P(FrameIn), _f, _i, _vp]
        L.orbr_search_for_initialization.restype = _i
        L.orbr_search_for_initialization.argtypes = [P(FrameIn), _vp, _vp, P(FrameIn), _i, _f, _i, _vp, _vp]
        L.orbr_search_for_triangulation.restype = _i
        L.orbr_search_for_triangulation.argtypes = [P(FrameIn), P(FrameIn), _vp, _f, _f, _i, _i, _f, _vp]
        L.orbr_search_by_projection_last.restype = _i
        L.orbr_search_by_projection_last.argtypes = [P(FrameIn), P(QueryIn), _f, _i, _i, _f, _i, _vp]
        L.orbr_search_by_projection_kf.restype = _i
        L.orbr_search_by_projection_kf.argtypes = [P(FrameIn), P(QueryIn), _f, _i, _f, _i, _vp]
        L.orbr_search_by_projection_scw.restype = _i
        L.orbr_search_by_projection_scw.argtypes = [P(FrameIn), P(QueryIn), _i, _vp]
        L.orbr_fuse.restype = _i
        L.orbr_fuse.argtypes = [P(FrameIn), P(QueryIn), _f, _vp]
        L.orbr_fuse_scw.restype = _i
        L.orbr_fuse_scw.argtypes = [P(FrameIn), P(QueryIn), _f, _vp]
        L.orbr_search_by_sim3.restype = _i
        L.orbr_search_by_sim3.argtypes = [P(FrameIn), P(QueryIn), P(FrameIn), P(QueryIn), _f, _vp]
        L.orbr_distinctive_descriptors.argtypes = [_i, _vp, _vp, _vp, P(FrameIn)]
        L.orbr_vocab_load.restype = _vp
        L.orbr_vocab_load.argtypes = [C.c_char_p]
        L.orbr_vocab_free.argtypes = [_vp]
        L.orbr_bow_transform.argtypes = [_vp, _vp, _i, _i, _vp, _vp, _vp]
        L.orbr_compute_bow.restype = _i
        L.orbr_compute_bow.argtypes = [_vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp]
        L.orbr_stereo_frame.restype = _vp
        L.orbr_stereo_frame.argtypes = [_vp, _vp, _i, _i, _i, _f, _i, _i, _i, _f, _f, _f, _f, _vp, _f]
        L.orbr_mono_frame.restype = _vp
        L.orbr_mono_frame.argtypes = [_vp, _i, _i, _i, _f, _i, _i, _i, _f, _f, _f, _f, _vp, _i, _f]
        L.orbr_stereo_n.restype = _i
        L.orbr_stereo_n.argtypes = [_vp, _i]
        L.orbr_stereo_get.argtypes = [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]
        L.orbr_stereo_free.argtypes = [_vp]
        L.orbr_bruteforce_scan.restype = C.c_double
        L.orbr_bruteforce_scan.argtypes = [_vp, _i, _vp, _i, _i, _i, _vp, _vp, _vp]
        L.orbr_last_ms.restype = C.c_double
        L.orbr_last_ms.argtypes = [_i]
        L.orbr_set_timing.argtypes = [_i]
        _lib = L
        _libs[[k for k, v in _PATHS.items() if v == _PATH][0]] = L
    return _lib


def _p(a):
    return a.ctypes.data if a is not None else None


class _Keep(list):
    """keeps converted arrays alive while a ctypes struct points at them"""

    def arr(self, a, dtype):
        if a is None:
            return None
        a = np.ascontiguousarray(a, dtype)
        self.append(a)
        return a.ctypes.data


def frame_in(fr, keep, mbf=40.0, max_x=None, max_y=None):
    """FrameIn from the flat frame / key-frame dicts the tests use (x, y, octave, [angle], [u_right], desc, [occupied],
    min_x, min_y, w_inv, h_inv, scale_factors, [node_id/node_start/node_feat], [has_mp])."""
    n = len(fr["desc"])
    zeros = np.zeros(max(n, 1), np.float32)
    f = FrameIn()
    f.n = n
    f.x = keep.arr(fr.get("x", zeros[:n]) if fr.get("x") is not None else zeros[:n], np.float32)
    f.y = keep.arr(fr.get("y", zeros[:n]) if fr.get("y") is not None else zeros[:n], np.float32)
    f.octave = keep.arr(fr["octave"] if fr.get("octave") is not None else np.zeros(n, np.int32), np.int32)
    f.angle = keep.arr(fr.get("angle"), np.float32)
    f.u_right = keep.arr(fr.get("u_right"), np.float32)
    f.desc = keep.arr(fr["desc"], np.uint8)
    f.occupied = keep.arr(fr.get("occupied"), np.uint8)
    f.min_x, f.min_y = fr.get("min_x", 0.0), fr.get("min_y", 0.0)
    w_inv, h_inv = fr.get("w_inv"), fr.get("h_inv")
    f.max_x = max_x if max_x is not None else fr.get("max_x", f.min_x + 64.0 / w_inv if w_inv else 1e9)
    f.max_y = max_y if max_y is not None else fr.get("max_y", f.min_y + 48.0 / h_inv if h_inv else 1e9)
    f.w_inv, f.h_inv = (w_inv or 0.1), (h_inv or 0.1)
    sf = fr.get("scale_factors")
    if sf is None:
        sf = (np.float32(1.2) ** np.arange(8)).astype(np.float32)
    f.scale_factors = keep.arr(sf, np.float32)
    f.nlevels = len(sf)
    f.mbf = mbf
    if fr.get("node_id") is not None:
        f.n_nodes = len(fr["node_id"])
        f.node_id = keep.arr(fr["node_id"], np.int32)
        f.node_start = keep.arr(fr["node_start"], np.int32)
        f.node_feat = keep.arr(fr["node_feat"], np.int32)
    f.has_mp = keep.arr(fr.get("has_mp"), np.uint8)
    return f


def query_in(q, keep, level_key="level"):
    qi = QueryIn()
    qi.n = len(q["desc"])
    qi.x = keep.arr(q["x"], np.float32)
    qi.y = keep.arr(q["y"], np.float32)
    qi.invz = keep.arr(q.get("invz"), np.float32)
    qi.level = keep.arr(q[level_key], np.int32)
    qi.angle = keep.arr(q.get("angle"), np.float32)
    qi.valid = keep.arr(q.get("valid"), np.uint8)
    qi.blocks = keep.arr(q.get("blocks"), np.uint8)
    qi.desc = keep.arr(q["desc"], np.uint8)
    return qi


def assign_grid(fr):
    keep = _Keep()
    f = frame_in(fr, keep)
    cs = np.zeros(64 * 48 + 1, np.int32)
    ci = np.zeros(max(f.n, 1), np.int32)
    lib().orbr_assign_grid(C.byref(f), cs.ctypes.data, ci.ctypes.data)
    return cs, ci[: cs[-1]]


def features_in_area(fr, x, y, r, min_level=-1, max_level=-1, keyframe=False):
    keep = _Keep()
    f = frame_in(fr, keep)
    out = np.zeros(max(f.n, 1), np.int32)
    k = lib().orbr_features_in_area(C.byref(f), int(keyframe), x, y, r, min_level, max_level, out.ctypes.data)
    return out[:k].copy()


def descriptor_distance(a, b):
    a = np.ascontiguousarray(a, np.uint8)
    b = np.ascontiguousarray(b, np.uint8)
    return lib().orbr_descriptor_distance(a.ctypes.data, b.ctypes.data)


def search_by_projection(frame, mps, th, nnratio):
    keep = _Keep()
    f = frame_in(frame, keep)
    out = np.full(max(f.n, 1), -1, np.int32)
    n = lib().orbr_search_by_projection(
        C.byref(f), len(mps["x"]), keep.arr(mps["x"], np.float32), keep.arr(mps["y"], np.float32), keep.arr(mps["xr"], np.float32),
        keep.arr(mps["level"], np.int32), keep.arr(mps["view_cos"], np.float32), keep.arr(mps["valid"], np.uint8),
        keep.arr(mps["obs"], np.uint8), keep.arr(mps["desc"], np.uint8), th, nnratio, out.ctypes.data)
    return n, out[: f.n]


def _bow_side(d):
    fr = dict(d)
    fr.setdefault("octave", None)
    return fr


def search_by_bow(kf, fr, nnratio, check_orientation):
    keep = _Keep()
    a = frame_in(_bow_side(kf), keep)
    b = frame_in(_bow_side(fr), keep)
    out = np.full(max(b.n, 1), -1, np.int32)
    n = lib().orbr_search_by_bow(C.byref(a), C.byref(b), nnratio, int(check_orientation), out.ctypes.data)
    return n, out[: b.n]


def search_by_bow_kf(kf1, kf2, nnratio, check_orientation):
    keep = _Keep()
    a = frame_in(_bow_side(kf1), keep)
    b = frame_in(_bow_side(kf2), keep)
    out = np.full(max(a.n, 1), -1, np.int32)
    n = lib().orbr_search_by_bow_kf(C.byref(a), C.byref(b), nnratio, int(check_orientation), out.ctypes.data)
    return n, out[: a.n]


def search_for_initialization(f1, f2, prev_matched, window_size, nnratio, check_orientation):
    keep = _Keep()
    a = frame_in(_bow_side(f1), keep)
    b = frame_in(f2, keep)
    pm = np.ascontiguousarray(prev_matched, np.float32).reshape(-1, 2)
    out = np.full(max(a.n, 1), -1, np.int32)
    prev_out = np.zeros((max(a.n, 1), 2), np.float32)
    n = lib().orbr_search_for_initialization(C.byref(a), keep.arr(pm[:, 0], np.float32), keep.arr(pm[:, 1], np.float32), C.byref(b),
                                             int(window_size), float(nnratio), int(check_orientation), out.ctypes.data,
                                             prev_out.ctypes.data)
    return n, out[: a.n], prev_out[: a.n]


def search_for_triangulation(kf1, kf2, F12, ex, ey, only_stereo, check_orientation, nnratio=0.6):
    keep = _Keep()
    a = frame_in(_bow_side(kf1), keep)
    b = frame_in(kf2, keep)
    F = np.ascontiguousarray(np.asarray(F12, np.float32).reshape(9))
    out = np.full(max(a.n, 1), -1, np.int32)
    n = lib().orbr_search_for_triangulation(C.byref(a), C.byref(b), F.ctypes.data, float(ex), float(ey), int(only_stereo),
                                            int(check_orientation), nnratio, out.ctypes.data)
    return n, out[: a.n]


def search_by_projection_last(cur, q, th, motion, mono, nnratio, check_orientation, mbf=40.0):
    """motion: 0 band / 1 forward / 2 backward (how far the last frame sits along z, against Cur.mb = mbf / fx = mbf)."""
    keep = _Keep()
    f = frame_in(cur, keep, mbf=mbf)
    qi = query_in(q, keep)
    out = np.full(max(f.n, 1), -1, np.int32)
    n = lib().orbr_search_by_projection_last(C.byref(f), C.byref(qi), th, motion, int(mono), nnratio, int(check_orientation),
                                             out.ctypes.data)
    return n, out[: f.n]


def search_by_projection_kf(cur, q, th, orb_dist, nnratio, check_orientation):
    keep = _Keep()
    f = frame_in(cur, keep)
    qi = query_in(q, keep)
    out = np.full(max(f.n, 1), -1, np.int32)
    n = lib().orbr_search_by_projection_kf(C.byref(f), C.byref(qi), th, int(orb_dist), nnratio, int(check_orientation), out.ctypes.data)
    return n, out[: f.n]


def search_by_projection_scw(kf, q, th):
    keep = _Keep()
    f = frame_in(kf, keep)
    qi = query_in(q, keep)
    out = np.full(max(f.n, 1), -1, np.int32)
    n = lib().orbr_search_by_projection_scw(C.byref(f), C.byref(qi), int(th), out.ctypes.data)
    return n, out[: f.n]


def fuse(kf, q, th, mbf=40.0, scw=False):
    keep = _Keep()
    f = frame_in(kf, keep, mbf=mbf)
    qi = query_in(q, keep)
    out = np.full(max(qi.n, 1), -1, np.int32)
    fn = lib().orbr_fuse_scw if scw else lib().orbr_fuse
    n = fn(C.byref(f), C.byref(qi), th, out.ctypes.data)
    return n, out[: qi.n]


def search_by_sim3(kf1, q2_in_1, kf2, q1_in_2, th):
    keep = _Keep()
    a = frame_in(kf1, keep)
    b = frame_in(kf2, keep)
    qa = query_in(q2_in_1, keep)
    qb = query_in(q1_in_2, keep)
    out = np.full(max(a.n, 1), -1, np.int32)
    n = lib().orbr_search_by_sim3(C.byref(a), C.byref(qa), C.byref(b), C.byref(qb), th, out.ctypes.data)
    return n, out[: a.n]


def distinctive_descriptors(set_start, desc):
    """-> (n_sets, 32) u8: the descriptor MapPoint::ComputeDistinctiveDescriptors keeps for every observation set."""
    keep = _Keep()
    set_start = np.ascontiguousarray(set_start, np.int32)
    desc = np.ascontiguousarray(desc, np.uint8)
    like = frame_in(dict(desc=np.zeros((1, 32), np.uint8), w_inv=0.1, h_inv=0.1), keep)
    out = np.zeros((max(len(set_start) - 1, 1), 32), np.uint8)
    lib().orbr_distinctive_descriptors(len(set_start) - 1, set_start.ctypes.data, desc.ctypes.data, out.ctypes.data, C.byref(like))
    return out[: len(set_start) - 1]


def write_vocabulary_text(vocab, path):
    """Write a flat synthetic vocabulary (vocabulary.synthetic_vocabulary) in the ORBvoc.txt format that the reference's
    TemplatedVocabulary::loadFromTextFile reads: header `k L scoring weighting`, then one line per node in id order:
    `parent is_leaf d0 .. d31 weight`.  Node ids must be 1..n in file order (the loader assigns them sequentially)."""
    cs, ci = vocab["child_start"], vocab["child_ids"]
    n = len(cs) - 1
    parent = np.zeros(n, np.int64)
    for p in range(n):
        parent[ci[cs[p]:cs[p + 1]]] = p
    with open(path, "w") as f:
        f.write("%d %d %d %d\n" % (int(cs[1] - cs[0]), vocab["L"], 0, 0))  # L1_NORM scoring, TF_IDF weighting
        lines = []
        for nid in range(1, n):
            leaf = cs[nid + 1] == cs[nid]
            lines.append("%d %d %s %r" % (parent[nid], int(leaf), " ".join(str(int(b)) for b in vocab["node_desc"][nid]),
                                          float(vocab["weight"][nid])))
        f.write("\n".join(lines))  # no trailing newline: the loader would read an empty line as one more node


class Vocabulary:
    """the reference's ORBVocabulary loaded from a text file written by write_vocabulary_text."""

    def __init__(self, vocab):
        fd, path = tempfile.mkstemp(suffix=".txt")
        os.close(fd)
        try:
            write_vocabulary_text(vocab, path)
            self.h = lib().orbr_vocab_load(path.encode())
        finally:
            os.unlink(path)
        if not self.h:
            raise RuntimeError("the reference's loadFromTextFile rejected the vocabulary")

    def close(self):
        if self.h:
            lib().orbr_vocab_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def transform_features(self, desc, levelsup=4):
        desc = np.ascontiguousarray(desc, np.uint8)
        n = len(desc)
        w = np.zeros(max(n, 1), np.int32)
        nd = np.zeros(max(n, 1), np.int32)
        wt = np.zeros(max(n, 1), np.float64)
        lib().orbr_bow_transform(self.h, desc.ctypes.data, n, levelsup, w.ctypes.data, nd.ctypes.data, wt.ctypes.data)
        return w[:n], nd[:n], wt[:n]

    def compute_bow(self, desc):
        """Frame::ComputeBoW (levelsup = 4): ((word ids, values), FeatureVector CSR dict)."""
        desc = np.ascontiguousarray(desc, np.uint8)
        n = len(desc)
        ids = np.zeros(max(n, 1), np.int32)
        vals = np.zeros(max(n, 1), np.float64)
        nn = C.c_int()
        nid = np.zeros(max(n, 1), np.int32)
        nst = np.zeros(max(n, 1) + 1, np.int32)
        nft = np.zeros(max(n, 1), np.int32)
        k = lib().orbr_compute_bow(self.h, desc.ctypes.data, n, ids.ctypes.data, vals.ctypes.data, C.byref(nn), nid.ctypes.data,
                                   nst.ctypes.data, nft.ctypes.data)
        m = nn.value
        return (ids[:k], vals[:k]), dict(node_id=nid[:m].copy(), node_start=nst[: m + 1].copy(), node_feat=nft[: nst[m]].copy())


class RefFrame:
    """The reference's Frame built by its own stereo / monocular constructor on images (extraction included)."""

    def __init__(self, h):
        self.h = h
        L = lib()
        self.n = L.orbr_stereo_n(h, 0)
        self.n_right = L.orbr_stereo_n(h, 1)
        n = max(self.n, 1)
        self.u_right = np.zeros(n, np.float32)
        self.depth = np.zeros(n, np.float32)
        self.keys = np.zeros((n, 2), np.float32)
        self.keys_un = np.zeros((n, 2), np.float32)
        self.cell_start = np.zeros(64 * 48 + 1, np.int32)
        ci = np.zeros(n, np.int32)
        self.bounds = np.zeros(4, np.float32)
        stereo = self.n_right > 0
        L.orbr_stereo_get(h, self.u_right.ctypes.data if stereo else None, self.depth.ctypes.data if stereo else None,
                          self.keys.ctypes.data, self.keys_un.ctypes.data, self.cell_start.ctypes.data, ci.ctypes.data,
                          self.bounds.ctypes.data)
        self.u_right, self.depth = self.u_right[: self.n], self.depth[: self.n]
        self.keys, self.keys_un = self.keys[: self.n], self.keys_un[: self.n]
        self.cell_items = ci[: self.cell_start[-1]]
        L.orbr_stereo_free(h)
        self.h = None


def bruteforce_scan(q, db, threads=1, variant=0):
    """Best / second-best scan on the host: variant 0 = the reference's own ORBmatcher::DescriptorDistance per pair,
    1 = __builtin_popcountll.  Returns (best, idx, second, ms of the scan alone)."""
    q = np.ascontiguousarray(q, np.uint8)
    db = np.ascontiguousarray(db, np.uint8)
    nq = len(q)
    best, idx, second = (np.zeros(nq, np.int32) for _ in range(3))
    ms = lib().orbr_bruteforce_scan(q.ctypes.data, nq, db.ctypes.data, len(db), int(threads), int(variant), best.ctypes.data,
                                    idx.ctypes.data, second.ctypes.data)
    return best, idx, second, float(ms)


def last_ms(which):
    """Wall time (ms) of the reference call alone inside the last wrapper invocation: 0 = SearchByProjection(Frame, MPs, th),
    1 = stereo Frame constructor, 2 = Frame::ComputeStereoMatches re-run (after set_timing(True))."""
    return float(lib().orbr_last_ms(which))


def set_timing(on):
    lib().orbr_set_timing(1 if on else 0)


def stereo_frame(left, right, cfg, fx, fy, cx, cy, mbf, dist4=None):
    left = np.ascontiguousarray(left, np.uint8)
    right = np.ascontiguousarray(right, np.uint8)
    d = np.ascontiguousarray(dist4, np.float32) if dist4 is not None else None
    h = lib().orbr_stereo_frame(left.ctypes.data, right.ctypes.data, left.shape[0], left.shape[1], cfg[0], cfg[1], cfg[2], cfg[3],
                                cfg[4], fx, fy, cx, cy, _p(d), mbf)
    return RefFrame(h)


def mono_frame(img, cfg, fx, fy, cx, cy, dist, mbf=40.0):
    img = np.ascontiguousarray(img, np.uint8)
    d = np.ascontiguousarray(dist, np.float32)
    h = lib().orbr_mono_frame(img.ctypes.data, img.shape[0], img.shape[1], cfg[0], cfg[1], cfg[2], cfg[3], cfg[4], fx, fy, cx, cy,
                              d.ctypes.data, len(d), mbf)
    return RefFrame(h)
