"""Python mirror of the Hamming side of ORB_SLAM2::ORBmatcher (reference include/ORBmatcher.h:48-264) over the C-ABI.

Frames, key frames and map points are passed as dicts of flat numpy arrays (the same flat views the C++ adapter
host/ORBmatcher.h gathers from Frame / KeyFrame / MapPoint):

  frame:  x, y (float32, mvKeysUn), octave (int32), u_right (float32 or None), desc (n,32 uint8), occupied (uint8),
          cell_start (64*48+1 int32), cell_items (int32), min_x, min_y, w_inv, h_inv (float), scale_factors (float32)
  mps:    x, y, xr (float32 mTrackProjX/Y/XR), level (int32), view_cos (float32), valid, obs (uint8), desc (n,32)
  featvec: node_id, node_start, node_feat (int32 CSR of DBoW2::FeatureVector)
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check


def _p(a):
    return a.ctypes.data if a is not None else None


def frame_grid(x, y, min_x, min_y, w_inv, h_inv):
    """Host-side Frame::AssignFeaturesToGrid / PosInGrid (reference src/Frame.cc:351-370, 558-567): 64x48 grid, cell
    (ix, iy) flattened as ix*48+iy, CSR (cell_start, cell_items) with items in keypoint order.  Caller-side glue (the
    reference keeps it in Frame); shipped so Python callers can build an orbm_frame_view."""
    x = np.asarray(x, np.float32)
    y = np.asarray(y, np.float32)
    fx = (x - np.float32(min_x)) * np.float32(w_inv)
    fy = (y - np.float32(min_y)) * np.float32(h_inv)
    # C round(): half away from zero
    px = np.where(fx >= 0, np.floor(fx + np.float32(0.5)), np.ceil(fx - np.float32(0.5))).astype(np.int64)
    py = np.where(fy >= 0, np.floor(fy + np.float32(0.5)), np.ceil(fy - np.float32(0.5))).astype(np.int64)
    ok = (px >= 0) & (px < 64) & (py >= 0) & (py < 48)
    idx = np.nonzero(ok)[0]
    cell = px[idx] * 48 + py[idx]
    order = np.argsort(cell, kind="stable")
    items = idx[order].astype(np.int32)
    counts = np.bincount(cell, minlength=64 * 48)
    start = np.zeros(64 * 48 + 1, np.int32)
    start[1:] = np.cumsum(counts)
    return start, items


class PreparedDatabase:
    """orbm_db handle: a database expanded to +-1 bytes with its TMA descriptor (see csrc/orbx_match_tc.cuh)."""

    def __init__(self, L, h, n):
        self._L, self._h, self.n = L, h, n

    def close(self):
        if getattr(self, "_h", None):
            self._L.orbm_db_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ORBmatcher:
    TH_LOW = 50
    TH_HIGH = 100
    HISTO_LENGTH = 30

    def __init__(self, nnratio=0.6, checkOri=True, device=0):
        self._L = _lib.load()
        self.mfNNratio = float(nnratio)
        self.mbCheckOrientation = bool(checkOri)
        m = C.c_void_p()
        check(self._L.orbm_create(int(device), C.byref(m)), "orbm_create")
        self._m = m

    def close(self):
        if getattr(self, "_m", None):
            self._L.orbm_destroy(self._m)
            self._m = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @staticmethod
    def DescriptorDistance(a, b):
        a = np.ascontiguousarray(a, np.uint8).reshape(-1)
        b = np.ascontiguousarray(b, np.uint8).reshape(-1)
        assert a.size == 32 and b.size == 32
        return _lib.load().orbm_descriptor_distance(a.ctypes.data, b.ctypes.data)

    @property
    def stream(self):
        return self._L.orbm_stream(self._m)

    def sync(self):
        check(self._L.orbm_sync(self._m), "orbm_sync")

    # ---- brute force (north_star item 6 / config 5) ----
    def BruteForce(self, q, db, idx_offset=0):
        q = np.ascontiguousarray(q, np.uint8)
        db = np.ascontiguousarray(db, np.uint8)
        nq = len(q)
        best, idx, second = (np.zeros(nq, np.int32) for _ in range(3))
        check(self._L.orbm_bruteforce(self._m, q.ctypes.data, nq, _p(db) if len(db) else None, len(db), idx_offset,
                                      best.ctypes.data, idx.ctypes.data, second.ctypes.data, 0), "orbm_bruteforce")
        return best, idx, second

    def bruteforce_device(self, dq, nq, ddb, ndb, idx_offset, d_best, d_idx, d_second):
        check(self._L.orbm_bruteforce(self._m, dq, nq, ddb, ndb, idx_offset, d_best, d_idx, d_second, 1), "orbm_bruteforce")

    # ---- brute force on the tensor cores (tcgen05 int8 GEMM + best / second epilogue), database prepared once ----
    def PrepareDatabase(self, db=None, d_db=None, ndb=None):
        """Expands a descriptor database for the tensor-core search.  db: (n, 32) uint8 host array, or d_db / ndb: raw device
        pointer and row count.  Returns a PreparedDatabase (close() frees the 256 B / row expansion)."""
        h = C.c_void_p()
        if db is not None:
            db = np.ascontiguousarray(db, np.uint8)
            check(self._L.orbm_db_create(self._m, db.ctypes.data, len(db), 0, C.byref(h)), "orbm_db_create")
            return PreparedDatabase(self._L, h, len(db))
        check(self._L.orbm_db_create(self._m, d_db, int(ndb), 1, C.byref(h)), "orbm_db_create")
        return PreparedDatabase(self._L, h, int(ndb))

    def BruteForceTC(self, q, prepared, idx_offset=0):
        q = np.ascontiguousarray(q, np.uint8)
        nq = len(q)
        best, idx, second = (np.zeros(nq, np.int32) for _ in range(3))
        check(self._L.orbm_bruteforce_db(self._m, prepared._h, q.ctypes.data, nq, idx_offset, best.ctypes.data, idx.ctypes.data,
                                         second.ctypes.data, 0), "orbm_bruteforce_db")
        return best, idx, second

    def bruteforce_tc_device(self, prepared, dq, nq, idx_offset, d_best, d_idx, d_second):
        check(self._L.orbm_bruteforce_db(self._m, prepared._h, dq, nq, idx_offset, d_best, d_idx, d_second, 1), "orbm_bruteforce_db")

    def bruteforce_sharded_tc_device(self, prepared, dq, nq, idx_offset, d_best, d_idx, d_second):
        check(self._L.orbm_bruteforce_sharded_db(self._m, prepared._h, dq, nq, idx_offset, d_best, d_idx, d_second),
              "orbm_bruteforce_sharded_db")

    def merge_partials_device(self, d_best, d_idx, d_second, nparts, nq, o_best, o_idx, o_second):
        check(self._L.orbm_merge_partials(self._m, d_best, d_idx, d_second, nparts, nq, o_best, o_idx, o_second),
              "orbm_merge_partials")

    # ---- frame preparation: Frame::UndistortKeyPoints / AssignFeaturesToGrid on the device ----
    def UndistortKeyPoints(self, x, y, fx, fy, cx, cy, dist):
        """mvKeysUn coordinates from mvKeys (src/Frame.cc:586-628); numpy in / out."""
        x = np.ascontiguousarray(x, np.float32)
        y = np.ascontiguousarray(y, np.float32)
        dist = np.ascontiguousarray(dist, np.float32)
        xu, yu = np.empty_like(x), np.empty_like(y)
        check(self._L.orbm_undistort_keypoints(self._m, _p(x), _p(y), len(x), fx, fy, cx, cy, _p(dist), len(dist), _p(xu), _p(yu), 0),
              "orbm_undistort_keypoints")
        return xu, yu

    def AssignFeaturesToGrid(self, x, y, min_x, min_y, w_inv, h_inv):
        """mGrid as CSR (cell_start[64*48+1], cell_items) from undistorted keypoints (src/Frame.cc:351-370)."""
        x = np.ascontiguousarray(x, np.float32)
        y = np.ascontiguousarray(y, np.float32)
        start = np.zeros(64 * 48 + 1, np.int32)
        items = np.zeros(max(len(x), 1), np.int32)
        check(self._L.orbm_assign_grid(self._m, _p(x), _p(y), len(x), min_x, min_y, w_inv, h_inv, _p(start), _p(items), 0),
              "orbm_assign_grid")
        return start, items[: start[-1]].copy()

    def undistort_device(self, dx, dy, n, fx, fy, cx, cy, dist, dxu, dyu):
        dist = np.ascontiguousarray(dist, np.float32)
        check(self._L.orbm_undistort_keypoints(self._m, dx, dy, n, fx, fy, cx, cy, _p(dist), len(dist), dxu, dyu, 1),
              "orbm_undistort_keypoints")

    def assign_grid_device(self, dx, dy, n, min_x, min_y, w_inv, h_inv, d_start, d_items):
        check(self._L.orbm_assign_grid(self._m, dx, dy, n, min_x, min_y, w_inv, h_inv, d_start, d_items, 1), "orbm_assign_grid")

    # ---- sharded brute force with the fused merge + exchange over peer memory (include/orbx.h) ----
    def exchange_create(self, world, rank, nq_max):
        """Allocates this rank's exchange buffer; returns (ipc_handle_bytes, local_device_pointer)."""
        import ctypes

        handle = (ctypes.c_uint8 * 64)()
        ptr = ctypes.c_void_p()
        check(self._L.orbm_exchange_create(self._m, world, rank, nq_max, ctypes.addressof(handle), ctypes.addressof(ptr)),
              "orbm_exchange_create")
        return bytes(handle), ptr.value

    def exchange_connect(self, handles):
        """handles: list of `world` 64-byte IPC handles in rank order (other processes on the same node)."""
        blob = b"".join(handles)
        buf = np.frombuffer(blob, np.uint8).copy()
        check(self._L.orbm_exchange_connect(self._m, buf.ctypes.data), "orbm_exchange_connect")

    def exchange_connect_local(self, ptrs):
        """ptrs: list of `world` device pointers (contexts of the same process)."""
        arr = np.asarray([p or 0 for p in ptrs], np.uint64)
        check(self._L.orbm_exchange_connect_local(self._m, arr.ctypes.data), "orbm_exchange_connect_local")

    def exchange_connect_group(self, world, rank, nq_max, group=None):
        """Convenience for one process per GPU under torch.distributed: create, all-gather the handles, connect."""
        import torch.distributed as dist

        handle, _ = self.exchange_create(world, rank, nq_max)
        handles = [None] * world
        dist.all_gather_object(handles, handle, group=group)
        if world > 1:
            self.exchange_connect(handles)

    def exchange_destroy(self):
        self._L.orbm_exchange_destroy(self._m)

    def bruteforce_sharded_device(self, dq, nq, ddb, ndb, idx_offset, d_best, d_idx, d_second, phases=3):
        """Asynchronous on the context's stream; every rank ends up with the merged result of all shards.
        phases: 3 = fused call, 1 = search + publish only, 2 = collect only (see include/orbx.h)."""
        check(self._L.orbm_bruteforce_sharded_phases(self._m, dq, nq, ddb, ndb, idx_offset, d_best, d_idx, d_second, phases),
              "orbm_bruteforce_sharded")

    def exchange_timed_out(self):
        return bool(self._L.orbm_exchange_timed_out(self._m))

    def debug_exchange_snapshot(self, world, nq_max):
        """(slot words [2, world, nq_max] uint64, searches published so far) of this rank's exchange buffer (test hook)."""
        words = np.zeros((2, world, nq_max), np.uint64)
        ep = C.c_uint32()
        check(self._L.orbm_debug_exchange_snapshot(self._m, words.ctypes.data, words.size, C.byref(ep)), "orbm_debug_exchange_snapshot")
        return words, ep.value

    # ---- SearchByProjection(Frame&, vector<MapPoint*>&, th) ----
    def SearchByProjection(self, frame, mps, th=3.0):
        keep = []

        def arr(v, dt):
            a = np.ascontiguousarray(v, dt)
            keep.append(a)
            return a

        n = len(frame["x"])
        fv = _lib.FrameView()
        fv.n = n
        fv.x = _p(arr(frame["x"], np.float32)); fv.y = _p(arr(frame["y"], np.float32))
        fv.octave = _p(arr(frame["octave"], np.int32))
        fv.u_right = _p(arr(frame["u_right"], np.float32)) if frame.get("u_right") is not None else None
        fv.desc = _p(arr(frame["desc"], np.uint8)); fv.occupied = _p(arr(frame["occupied"], np.uint8))
        fv.cell_start = _p(arr(frame["cell_start"], np.int32)); fv.cell_items = _p(arr(frame["cell_items"], np.int32))
        fv.min_x, fv.min_y, fv.grid_w_inv, fv.grid_h_inv = frame["min_x"], frame["min_y"], frame["w_inv"], frame["h_inv"]
        sf = arr(frame["scale_factors"], np.float32)
        fv.scale_factors = _p(sf); fv.nlevels = len(sf)
        mv = _lib.MapPointView()
        mv.n = len(mps["x"])
        mv.proj_x = _p(arr(mps["x"], np.float32)); mv.proj_y = _p(arr(mps["y"], np.float32))
        mv.proj_xr = _p(arr(mps["xr"], np.float32)); mv.level = _p(arr(mps["level"], np.int32))
        mv.view_cos = _p(arr(mps["view_cos"], np.float32)); mv.valid = _p(arr(mps["valid"], np.uint8))
        mv.obs_pos = _p(arr(mps["obs"], np.uint8)); mv.desc = _p(arr(mps["desc"], np.uint8))
        out = np.full(max(n, 1), -1, np.int32)
        nm = C.c_int()
        check(self._L.orbm_search_by_projection(self._m, C.byref(fv), C.byref(mv), float(th), self.mfNNratio,
                                                out.ctypes.data, C.byref(nm)), "orbm_search_by_projection")
        return nm.value, out[:n]

    # ---- windowed search shared by SearchByProjection(Cur, Last, th, bMono) and (Cur, KF, found, th, ORBdist) ----
    def SearchWindow(self, frame, q, max_dist=100, ratio_mode=0):
        """frame: frame dict (+ "angle" = mvKeysUn angles when mbCheckOrientation); q: dict with x, y, radius, min_level,
        max_level, xr (or None), angle (or None), valid, blocks, desc -- the source features after projection.
        Returns (nmatches, out) with out[i] = -1 untouched, -2 set to NULL by the rotation filter, else query index."""
        keep = []

        def arr(v, dt):
            a = np.ascontiguousarray(v, dt)
            keep.append(a)
            return a

        n = len(frame["x"])
        fv = _lib.FrameView()
        fv.n = n
        fv.x = _p(arr(frame["x"], np.float32)); fv.y = _p(arr(frame["y"], np.float32))
        fv.octave = _p(arr(frame["octave"], np.int32))
        fv.u_right = _p(arr(frame["u_right"], np.float32)) if frame.get("u_right") is not None else None
        fv.desc = _p(arr(frame["desc"], np.uint8)); fv.occupied = _p(arr(frame["occupied"], np.uint8))
        fv.cell_start = _p(arr(frame["cell_start"], np.int32)); fv.cell_items = _p(arr(frame["cell_items"], np.int32))
        fv.min_x, fv.min_y, fv.grid_w_inv, fv.grid_h_inv = frame["min_x"], frame["min_y"], frame["w_inv"], frame["h_inv"]
        sf = arr(frame["scale_factors"], np.float32)
        fv.scale_factors = _p(sf); fv.nlevels = len(sf)
        f_angle = arr(frame["angle"], np.float32) if self.mbCheckOrientation else None
        qv = _lib.WindowQueryView()
        qv.n = len(q["x"])
        qv.x = _p(arr(q["x"], np.float32)); qv.y = _p(arr(q["y"], np.float32)); qv.radius = _p(arr(q["radius"], np.float32))
        qv.min_level = _p(arr(q["min_level"], np.int32)); qv.max_level = _p(arr(q["max_level"], np.int32))
        qv.xr = _p(arr(q["xr"], np.float32)) if q.get("xr") is not None else None
        qv.angle = _p(arr(q["angle"], np.float32)) if self.mbCheckOrientation else None
        qv.valid = _p(arr(q["valid"], np.uint8)); qv.claims_block = _p(arr(q["blocks"], np.uint8))
        qv.desc = _p(arr(q["desc"], np.uint8))
        out = np.full(max(n, 1), -1, np.int32)
        nm = C.c_int()
        check(self._L.orbm_search_window(self._m, C.byref(fv), _p(f_angle), C.byref(qv), int(max_dist), int(ratio_mode),
                                         self.mfNNratio, int(self.mbCheckOrientation), out.ctypes.data, C.byref(nm)),
              "orbm_search_window")
        return nm.value, out[:n]

    # ---- stereo row search (Frame::ComputeStereoMatches coarse part) ----
    def StereoRows(self, left, right, scale_factors, n_rows, min_d, max_d):
        keep = []

        def arr(v, dt):
            a = np.ascontiguousarray(v, dt)
            keep.append(a)
            return a

        nl, nr = len(left["x"]), len(right["x"])
        sf = arr(scale_factors, np.float32)
        best_r = np.full(max(nl, 1), -1, np.int32)
        best_d = np.zeros(max(nl, 1), np.int32)
        check(self._L.orbm_stereo_rows(
            self._m, nl, _p(arr(left["x"], np.float32)), _p(arr(left["y"], np.float32)), _p(arr(left["octave"], np.int32)),
            _p(arr(left["desc"], np.uint8)), nr, _p(arr(right["x"], np.float32)), _p(arr(right["y"], np.float32)),
            _p(arr(right["octave"], np.int32)), _p(arr(right["desc"], np.uint8)), _p(sf), len(sf), int(n_rows),
            float(min_d), float(max_d), best_r.ctypes.data, best_d.ctypes.data), "orbm_stereo_rows")
        return best_r[:nl], best_d[:nl]

    # ---- Frame::ComputeStereoMatches, complete (coarse search + SAD refinement + outlier removal) ----
    def ComputeStereoMatches(self, ex_left, ex_right, left, right, mbf, mb, frame_left=0, frame_right=0, upstream_outliers=False):
        """ex_left / ex_right: ORBextractor objects that have just extracted the left / right image (their device
        pyramids are read in place).  Returns (mvuRight, mvDepth).  upstream_outliers=True runs the SAD outlier block once
        after the loop (upstream ORB-SLAM2) instead of inside it (this reference, src/Frame.cc:868-884)."""
        keep = []
        check(self._L.orbm_set_stereo_outlier_mode(self._m, 1 if upstream_outliers else 0), "orbm_set_stereo_outlier_mode")

        def arr(v, dt):
            a = np.ascontiguousarray(v, dt)
            keep.append(a)
            return a

        nl, nr = len(left["x"]), len(right["x"])
        ur = np.full(max(nl, 1), -1, np.float32)
        dp = np.full(max(nl, 1), -1, np.float32)
        nm = C.c_int()
        check(self._L.orbm_stereo_match(
            self._m, ex_left._h, frame_left, ex_right._h, frame_right, nl, _p(arr(left["x"], np.float32)),
            _p(arr(left["y"], np.float32)), _p(arr(left["octave"], np.int32)), _p(arr(left["desc"], np.uint8)), nr,
            _p(arr(right["x"], np.float32)), _p(arr(right["y"], np.float32)), _p(arr(right["octave"], np.int32)),
            _p(arr(right["desc"], np.uint8)), float(mbf), float(mb), ur.ctypes.data, dp.ctypes.data, C.byref(nm)), "orbm_stereo_match")
        return ur[:nl], dp[:nl]

    # ---- SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12) ----
    def SearchByBoWKF(self, kf1, kf2):
        keep = []

        def arr(v, dt):
            a = np.ascontiguousarray(v, dt)
            keep.append(a)
            return a

        def fvview(d):
            v = _lib.FeatVecView()
            v.n_nodes = len(d["node_id"])
            v.node_id = _p(arr(d["node_id"], np.int32)); v.node_start = _p(arr(d["node_start"], np.int32))
            v.node_feat = _p(arr(d["node_feat"], np.int32))
            return v

        n1, n2 = len(kf1["angle"]), len(kf2["angle"])
        v1, v2 = fvview(kf1), fvview(kf2)
        out = np.full(max(n1, 1), -1, np.int32)
        nm = C.c_int()
        check(self._L.orbm_search_by_bow_kf(
            self._m, n1, _p(arr(kf1["desc"], np.uint8)), _p(arr(kf1["angle"], np.float32)), _p(arr(kf1["has_mp"], np.uint8)), C.byref(v1),
            n2, _p(arr(kf2["desc"], np.uint8)), _p(arr(kf2["angle"], np.float32)), _p(arr(kf2["has_mp"], np.uint8)), C.byref(v2),
            self.mfNNratio, int(self.mbCheckOrientation), out.ctypes.data, C.byref(nm)), "orbm_search_by_bow_kf")
        return nm.value, out[:n1]

    # ---- MapPoint::ComputeDistinctiveDescriptors, batched ----
    def ComputeDistinctiveDescriptors(self, set_start, desc):
        set_start = np.ascontiguousarray(set_start, np.int32)
        desc = np.ascontiguousarray(desc, np.uint8)
        n = len(set_start) - 1
        out = np.zeros(max(n, 1), np.int32)
        check(self._L.orbm_distinctive_descriptors(self._m, n, set_start.ctypes.data, desc.ctypes.data, out.ctypes.data),
              "orbm_distinctive_descriptors")
        return out[:n]

    # ---- SearchByBoW(KeyFrame*, Frame&, matches) ----
    def SearchByBoW(self, kf, fr):
        keep = []

        def arr(v, dt):
            a = np.ascontiguousarray(v, dt)
            keep.append(a)
            return a

        def fvview(d):
            v = _lib.FeatVecView()
            v.n_nodes = len(d["node_id"])
            v.node_id = _p(arr(d["node_id"], np.int32)); v.node_start = _p(arr(d["node_start"], np.int32))
            v.node_feat = _p(arr(d["node_feat"], np.int32))
            return v

        nkf, nf = len(kf["angle"]), len(fr["angle"])
        kv, fv = fvview(kf), fvview(fr)
        out = np.full(max(nf, 1), -1, np.int32)
        nm = C.c_int()
        check(self._L.orbm_search_by_bow(
            self._m, nkf, _p(arr(kf["desc"], np.uint8)), _p(arr(kf["angle"], np.float32)), _p(arr(kf["has_mp"], np.uint8)),
            C.byref(kv), nf, _p(arr(fr["desc"], np.uint8)), _p(arr(fr["angle"], np.float32)), C.byref(fv),
            self.mfNNratio, int(self.mbCheckOrientation), out.ctypes.data, C.byref(nm)), "orbm_search_by_bow")
        return nm.value, out[:nf]

    # ---- shared view builders for the remaining overloads ----
    @staticmethod
    def _frame_view(frame, keep):
        def arr(v, dt):
            a = np.ascontiguousarray(v, dt)
            keep.append(a)
            return a

        fv = _lib.FrameView()
        fv.n = len(frame["x"])
        fv.x = _p(arr(frame["x"], np.float32)); fv.y = _p(arr(frame["y"], np.float32))
        fv.octave = _p(arr(frame["octave"], np.int32))
        fv.u_right = _p(arr(frame["u_right"], np.float32)) if frame.get("u_right") is not None else None
        fv.desc = _p(arr(frame["desc"], np.uint8))
        fv.occupied = _p(arr(frame["occupied"], np.uint8)) if frame.get("occupied") is not None else None
        fv.cell_start = _p(arr(frame["cell_start"], np.int32)); fv.cell_items = _p(arr(frame["cell_items"], np.int32))
        fv.min_x, fv.min_y, fv.grid_w_inv, fv.grid_h_inv = frame["min_x"], frame["min_y"], frame["w_inv"], frame["h_inv"]
        sf = arr(frame["scale_factors"], np.float32)
        fv.scale_factors = _p(sf); fv.nlevels = len(sf)
        return fv

    @staticmethod
    def _query_view(q, keep):
        def arr(v, dt):
            a = np.ascontiguousarray(v, dt)
            keep.append(a)
            return a

        qv = _lib.WindowQueryView()
        qv.n = len(q["x"])
        qv.x = _p(arr(q["x"], np.float32)); qv.y = _p(arr(q["y"], np.float32)); qv.radius = _p(arr(q["radius"], np.float32))
        qv.min_level = _p(arr(q["min_level"], np.int32)); qv.max_level = _p(arr(q["max_level"], np.int32))
        qv.xr = _p(arr(q["xr"], np.float32)) if q.get("xr") is not None else None
        qv.angle = None
        qv.valid = _p(arr(q["valid"], np.uint8)); qv.claims_block = None
        qv.desc = _p(arr(q["desc"], np.uint8))
        return qv

    # ---- Fuse (both overloads): the claim-free search; the map mutation stays with the caller ----
    def FuseSearch(self, kf, q, gate=1, inv_level_sigma2=None, max_dist=None):
        """kf: frame dict of the KeyFrame; q: projected map points (x, y, xr, radius, min_level, max_level, valid, desc).
        gate=1: Fuse(pKF, vpMapPoints, th) (chi2 gate), gate=0: Fuse(pKF, Scw, ...).  Returns (best_idx, best_dist)."""
        keep = []
        fv, qv = self._frame_view(kf, keep), self._query_view(q, keep)
        sig = np.ascontiguousarray(inv_level_sigma2, np.float32) if inv_level_sigma2 is not None else None
        nq = qv.n
        bi = np.full(max(nq, 1), -1, np.int32)
        bd = np.full(max(nq, 1), 256, np.int32)
        check(self._L.orbm_search_best(self._m, C.byref(fv), C.byref(qv), int(gate), _p(sig),
                                       self.TH_LOW if max_dist is None else int(max_dist), bi.ctypes.data, bd.ctypes.data),
              "orbm_search_best")
        return bi[:nq], bd[:nq]

    def SearchBySim3(self, kf1, q2_in_1, kf2, q1_in_2):
        """kf1/kf2: frame dicts; q1_in_2: KF1's map points projected into KF2 (one per KF1 feature), q2_in_1 the reverse.
        Returns (nFound, match12)."""
        keep = []
        f1, f2 = self._frame_view(kf1, keep), self._frame_view(kf2, keep)
        qa, qb = self._query_view(q2_in_1, keep), self._query_view(q1_in_2, keep)
        out = np.full(max(f1.n, 1), -1, np.int32)
        nf = C.c_int()
        check(self._L.orbm_search_by_sim3(self._m, C.byref(f1), C.byref(qa), C.byref(f2), C.byref(qb), out.ctypes.data,
                                          C.byref(nf)), "orbm_search_by_sim3")
        return nf.value, out[: f1.n]

    def SearchForInitialization(self, f1, f2, prev_matched, windowSize=10):
        """f1: dict octave, angle, desc; f2: frame dict (+ "angle"); prev_matched: (n1, 2) float32 = vbPrevMatched.
        Returns (nmatches, vnMatches12, updated vbPrevMatched)."""
        keep = []
        fv = self._frame_view(f2, keep)
        n1 = len(f1["octave"])
        oct1 = np.ascontiguousarray(f1["octave"], np.int32)
        ang1 = np.ascontiguousarray(f1["angle"], np.float32)
        d1 = np.ascontiguousarray(f1["desc"], np.uint8)
        pm = np.ascontiguousarray(prev_matched, np.float32).reshape(-1, 2)
        px, py = np.ascontiguousarray(pm[:, 0]), np.ascontiguousarray(pm[:, 1])
        ang2 = np.ascontiguousarray(f2["angle"], np.float32)
        out = np.full(max(n1, 1), -1, np.int32)
        nm = C.c_int()
        check(self._L.orbm_search_for_initialization(
            self._m, n1, _p(oct1), _p(ang1), _p(d1), _p(px), _p(py), C.byref(fv), _p(ang2), int(windowSize), self.mfNNratio,
            int(self.mbCheckOrientation), out.ctypes.data, C.byref(nm)), "orbm_search_for_initialization")
        out = out[:n1]
        new_prev = pm.copy()
        ok = out >= 0
        x2 = np.asarray(f2["x"], np.float32); y2 = np.asarray(f2["y"], np.float32)
        new_prev[ok, 0] = x2[out[ok]]
        new_prev[ok, 1] = y2[out[ok]]
        return nm.value, out, new_prev

    def SearchForTriangulation(self, kf1, kf2, F12, ex, ey, scale_factors2, level_sigma2_2, bOnlyStereo=False):
        """kf1/kf2: dicts x, y, octave, angle, u_right, desc, has_mp, node_id, node_start, node_feat.
        Returns (nmatches, match12); vMatchedPairs = [(i, match12[i]) for i where match12[i] >= 0]."""
        keep = []

        def arr(v, dt):
            a = np.ascontiguousarray(v, dt)
            keep.append(a)
            return a

        def kfview(d):
            fv = _lib.FeatVecView()
            fv.n_nodes = len(d["node_id"])
            fv.node_id = _p(arr(d["node_id"], np.int32)); fv.node_start = _p(arr(d["node_start"], np.int32))
            fv.node_feat = _p(arr(d["node_feat"], np.int32))
            keep.append(fv)
            v = _lib.KFView()
            v.n = len(d["x"])
            v.x = _p(arr(d["x"], np.float32)); v.y = _p(arr(d["y"], np.float32)); v.octave = _p(arr(d["octave"], np.int32))
            v.angle = _p(arr(d["angle"], np.float32)); v.u_right = _p(arr(d["u_right"], np.float32))
            v.desc = _p(arr(d["desc"], np.uint8)); v.has_mp = _p(arr(d["has_mp"], np.uint8))
            v.fv = C.pointer(fv)
            return v

        v1, v2 = kfview(kf1), kfview(kf2)
        F = arr(np.asarray(F12, np.float32).reshape(9), np.float32)
        sf = arr(scale_factors2, np.float32); sg = arr(level_sigma2_2, np.float32)
        out = np.full(max(v1.n, 1), -1, np.int32)
        nm = C.c_int()
        check(self._L.orbm_search_for_triangulation(
            self._m, C.byref(v1), C.byref(v2), _p(F), float(ex), float(ey), _p(sf), _p(sg), len(sf), int(bool(bOnlyStereo)),
            int(self.mbCheckOrientation), out.ctypes.data, C.byref(nm)), "orbm_search_for_triangulation")
        return nm.value, out[: v1.n]
