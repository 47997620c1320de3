"""Deterministic matcher test cases shared by three consumers:

* tests/test_oracle_matcher.py  (CPU): oracle/match_oracle.cc == the compiled reference (oracle/_ref/libmatchref.so);
* tools/make_golden.py          : freezes the reference's outputs for every case into tests/golden/matcher_ref.npz;
* tests/test_gpu_matcher_ref.py (GPU): the CUDA path == those frozen reference outputs (and the live reference library
  when it travelled to the box).

Every case is a pure function of its parameters (numpy PCG64 + the pinned CPU extractor), so all three see the same bytes.
Inputs are "after projection" (u, v, ur, radius, level range) as the C-ABI takes them, plus what the reference harness needs
to reproduce that projection exactly: the inverse depth `invz` (a power of two, ur = u - mbf * invz in float32) and the
predicted level `level` (see oracle/matchref_wrap.cc)."""
import numpy as np

import helpers
from orbslam2commentedbyxcm_b200 import synth

MBF = 40.0
INVZ = np.array([0.125, 0.25, 0.5, 1.0, 2.0], np.float32)

_cache = {}


def extracted(oracle_mod, shape_name, seed):
    key = (shape_name, seed)
    if key not in _cache:
        img, cfg, oex, kps, desc = helpers.extract_with_oracle(oracle_mod, shape_name, seed)
        _cache[key] = (kps, desc, np.asarray(oex.scale_factors(), np.float32).copy())
    return _cache[key]


def flip(desc, rng, rates=(0.03, 0.1, 0.3)):
    n = len(desc)
    flips = rng.random((n, 256)) < rng.choice(list(rates), n)[:, None]
    return np.ascontiguousarray(desc ^ np.packbits(flips, axis=1, bitorder="little"))


def frame_view(oracle_mod, shape_name, seed, occupied_p=0.1, stereo=False):
    """Frame / key-frame view of an extracted image.  Stereo keypoints get u_right = x - MBF * invz_kp + jitter so that
    queries reusing invz_kp pass the right-coordinate gates."""
    kps, desc, sf = extracted(oracle_mod, shape_name, seed)
    rows, cols = synth.SHAPES[shape_name]
    fr = helpers.make_frame_view(oracle_mod, kps, desc, rows, cols, sf, occupied_p=occupied_p, seed=seed, stereo=False)
    rng = np.random.Generator(np.random.PCG64(seed + 4242))
    n = len(kps)
    fr["kp_invz"] = rng.choice(INVZ, n).astype(np.float32)
    if stereo:
        ur = (fr["x"] - np.float32(MBF) * fr["kp_invz"]).astype(np.float32) + rng.normal(0, 1.0, n).astype(np.float32)
        fr["u_right"] = np.where(rng.random(n) < 0.6, np.maximum(ur, np.float32(0.5)), np.float32(-1)).astype(np.float32)
    else:
        fr["u_right"] = None
    fr["angle"] = kps["angle"].astype(np.float32).copy()
    fr["max_x"], fr["max_y"] = float(cols), float(rows)
    return fr, kps


def _pick(rng, n, nq):
    pick = rng.integers(0, n, nq)
    pick[nq * 3 // 4:] = pick[rng.integers(0, nq * 3 // 4, nq - nq * 3 // 4)]  # several sources per keypoint
    return pick


def _positions(frame, pick, rng, sigma):
    nq = len(pick)
    x = np.clip(frame["x"][pick] + rng.normal(0, sigma, nq), 1.0, frame["max_x"] - 2.0).astype(np.float32)
    y = np.clip(frame["y"][pick] + rng.normal(0, sigma, nq), 1.0, frame["max_y"] - 2.0).astype(np.float32)
    invz = np.where(rng.random(nq) < 0.8, frame["kp_invz"][pick], rng.choice(INVZ, nq)).astype(np.float32)
    xr = (x - np.float32(MBF) * invz).astype(np.float32)
    return x, y, invz, xr


def window_queries(frame, kps, nq, seed, mode, nlevels=8):
    """Source features 'after projection' for SearchByProjection(Cur, Last, th, bMono): mode 'forward' / 'backward' /
    'band' = the three GetFeaturesInArea level patterns (src/ORBmatcher.cc:1684-1690); mode 'kf' = the
    (Cur, KF, sAlreadyFound, th, ORBdist) overload (:1792-1924: band [pred-1, pred+1], every non-NULL entry blocks, no
    right-coordinate gate)."""
    rng = np.random.Generator(np.random.PCG64(seed + 7))
    n = len(frame["x"])
    pick = _pick(rng, n, nq)
    octv = np.clip(frame["octave"][pick] + rng.integers(-1, 2, nq), 0, nlevels - 1).astype(np.int32)
    th = 7.0 if mode != "kf" else 10.0
    radius = (np.float32(th) * frame["scale_factors"][octv]).astype(np.float32)
    if mode == "forward":
        lmin, lmax = octv.copy(), np.full(nq, -1, np.int32)
    elif mode == "backward":
        lmin, lmax = np.zeros(nq, np.int32), octv.copy()
    else:
        lmin, lmax = octv - 1, octv + 1
    x, y, invz, xr = _positions(frame, pick, rng, 3.0)
    rot = np.where(rng.random(nq) < 0.75, 12.0, rng.random(nq) * 360)
    return dict(x=x, y=y, radius=radius, min_level=lmin.astype(np.int32), max_level=lmax.astype(np.int32),
                xr=None if mode == "kf" else xr, invz=invz, level=octv, th=th,
                angle=np.mod(kps["angle"][pick] + rot, 360).astype(np.float32), valid=(rng.random(nq) < 0.9).astype(np.uint8),
                blocks=np.ones(nq, np.uint8) if mode == "kf" else (rng.random(nq) < 0.8).astype(np.uint8),
                desc=flip(frame["desc"][pick], rng))


def projected_points(frame, nq, seed, th, nlevels=8, stereo=False, sigma=1.5):
    """Map points 'after projection' into a key frame: levels [pred-1, pred], radius = th * scale[pred]
    (SearchByProjection(KF, Scw), Fuse x2, the SearchBySim3 passes)."""
    rng = np.random.Generator(np.random.PCG64(seed + 17))
    n = len(frame["x"])
    pick = _pick(rng, n, nq)
    pred = np.clip(frame["octave"][pick] + rng.integers(-1, 2, nq), 0, nlevels - 1).astype(np.int32)
    x, y, invz, xr = _positions(frame, pick, rng, sigma)
    return dict(x=x, y=y, xr=xr if stereo else None, invz=invz, level=pred, th=th,
                radius=(np.float32(th) * frame["scale_factors"][pred]).astype(np.float32),
                min_level=(pred - 1).astype(np.int32), max_level=pred.copy(), valid=(rng.random(nq) < 0.9).astype(np.uint8),
                blocks=np.ones(nq, np.uint8), angle=np.zeros(nq, np.float32), desc=flip(frame["desc"][pick], rng))


def second_view(oracle_mod, frame, rng, jitter=1.0):
    """A second key frame seeing the same features: permuted, jittered, descriptors perturbed."""
    n = len(frame["x"])
    perm = rng.permutation(n)
    x = np.clip(frame["x"][perm] + rng.normal(0, jitter, n), 0, frame["max_x"] - 1).astype(np.float32)
    y = np.clip(frame["y"][perm] + rng.normal(0, jitter, n), 0, frame["max_y"] - 1).astype(np.float32)
    cs, ci = oracle_mod.assign_grid(x, y, frame["min_x"], frame["min_y"], frame["w_inv"], frame["h_inv"])
    f2 = dict(frame)
    f2.update(x=x, y=y, octave=frame["octave"][perm].copy(), desc=flip(frame["desc"][perm], rng, (0.03, 0.08)),
              cell_start=cs, cell_items=ci, u_right=None, occupied=np.zeros(n, np.uint8), kp_invz=frame["kp_invz"][perm].copy(),
              angle=frame["angle"][perm].copy())
    return f2, perm


def featvec_from_nodes(node):
    ids = np.unique(node)
    start, feat = [0], []
    for nid in ids:
        f = np.nonzero(node == nid)[0]
        feat.extend(f.tolist())
        start.append(len(feat))
    return dict(node_id=ids.astype(np.int32), node_start=np.array(start, np.int32), node_feat=np.array(feat, np.int32))


# ------------------------------------------------------------------------------------------------------------------
# cases
# ------------------------------------------------------------------------------------------------------------------
SBP_CASES = [("kitti", 0, 4000, 1.0, False), ("kitti", 1, 4000, 3.0, True), ("tum", 2, 1500, 5.0, False), ("euroc", 3, 50, 1.0, True)]


def sbp_case(oracle_mod, shape_name, seed, nmp, th, stereo):
    frame, kps = frame_view(oracle_mod, shape_name, seed, stereo=stereo)
    mps = helpers.make_mappoints(frame, nmp, seed=seed, stereo=stereo)
    return frame, mps


def sbp_chain_case(oracle_mod):
    """a long chain of map points that all want the same keypoint: worst case for the order-dependent claims."""
    frame, kps = frame_view(oracle_mod, "tum", 4, occupied_p=0.0)
    tgt = int(np.argmax((kps["octave"] == 0)))
    nmp = 64
    rng = np.random.default_rng(0)
    mps = dict(x=np.full(nmp, frame["x"][tgt], np.float32), y=np.full(nmp, frame["y"][tgt], np.float32),
               xr=np.full(nmp, -1, np.float32), level=np.zeros(nmp, np.int32), view_cos=np.full(nmp, 0.9, np.float32),
               valid=np.ones(nmp, np.uint8), obs=(rng.random(nmp) < 0.7).astype(np.uint8),
               desc=np.ascontiguousarray(np.repeat(frame["desc"][tgt][None], nmp, 0)))
    return frame, mps


BOW_CASES = [(0, True), (1, False), (2, True)]


def bow_case(oracle_mod, seed, kfkf=False):
    kps, desc, sf = extracted(oracle_mod, "tum", seed + (10 if kfkf else 0))
    rng = np.random.default_rng(seed)
    n = len(kps)
    perm = rng.permutation(n)
    flips = rng.random((n, 256)) < 0.06
    d1 = np.ascontiguousarray(desc[perm] ^ np.packbits(flips, axis=1, bitorder="little"))
    rot = np.where(rng.random(n) < 0.8, 40.0 if kfkf else 24.0, rng.random(n) * 360).astype(np.float32)
    a1 = np.mod(kps["angle"][perm] + rot, 360).astype(np.float32)
    nn = 300 if kfkf else 400
    fv2, node_of = helpers.make_featvec(n, nn, seed)
    node1 = np.where(rng.random(n) < 0.9, node_of[perm], rng.integers(0, nn, n))
    kf1 = dict(desc=d1, angle=a1, has_mp=(rng.random(n) < 0.8).astype(np.uint8), **featvec_from_nodes(node1))
    side2 = dict(desc=np.ascontiguousarray(desc), angle=kps["angle"].astype(np.float32).copy(), **fv2)
    if kfkf:
        side2["has_mp"] = (rng.random(n) < 0.85).astype(np.uint8)
    return kf1, side2


INIT_CASES = [("tum", 6, True, 100), ("tum", 7, False, 100), ("kitti", 8, True, 40), ("euroc", 9, True, 100)]


def init_case(oracle_mod, shape_name, seed):
    base, kps = frame_view(oracle_mod, shape_name, seed)
    rng = np.random.default_rng(seed)
    f2, perm = second_view(oracle_mod, base, rng, jitter=3.0)
    rot = np.where(rng.random(len(perm)) < 0.8, 25.0, rng.random(len(perm)) * 360)
    f2["angle"] = np.mod(kps["angle"][perm] + rot, 360).astype(np.float32)
    lvl0 = np.nonzero(base["octave"] == 0)[0]
    extra = np.repeat(lvl0, rng.integers(0, 4, len(lvl0)))
    order = rng.permutation(np.concatenate([np.arange(len(kps)), extra]))
    f1 = dict(octave=base["octave"][order].copy(), angle=kps["angle"][order].astype(np.float32),
              desc=flip(base["desc"][order], rng, (0.0, 0.02, 0.05)), scale_factors=base["scale_factors"],
              min_x=base["min_x"], min_y=base["min_y"], w_inv=base["w_inv"], h_inv=base["h_inv"], max_x=base["max_x"],
              max_y=base["max_y"])
    prev = np.stack([base["x"][order], base["y"][order]], axis=1).astype(np.float32)
    return f1, f2, prev


TRIANG_CASES = [(0, True, False, 0.0), (1, True, False, 0.5), (2, False, True, 0.6), (3, True, False, 1.0)]


def triang_case(oracle_mod, seed, stereo_p):
    """rectified pair (F12 of a pure x translation: epipolar lines are rows), few vocabulary nodes so that nodes hold
    many features and the vbMatched2 claims and the last-of-equal-distances rule are exercised."""
    base, kps = frame_view(oracle_mod, "tum", seed + 20)
    kps, desc, sf = extracted(oracle_mod, "tum", seed + 20)
    rng = np.random.default_rng(seed)
    n = len(kps)
    octave = kps["octave"].astype(np.int32)
    node2 = rng.integers(0, 60, n)
    common = dict(scale_factors=sf, min_x=base["min_x"], min_y=base["min_y"], w_inv=base["w_inv"], h_inv=base["h_inv"],
                  max_x=base["max_x"], max_y=base["max_y"])

    def kf(x, y, octv, angle, ur, d, has_mp, node):
        return dict(x=x.astype(np.float32), y=y.astype(np.float32), octave=octv.astype(np.int32), angle=angle.astype(np.float32),
                    u_right=ur.astype(np.float32), desc=np.ascontiguousarray(d), has_mp=has_mp.astype(np.uint8),
                    **featvec_from_nodes(node), **common)

    kf2 = kf(kps["x"], kps["y"], octave, kps["angle"], np.where(rng.random(n) < stereo_p, kps["x"] - 5, -1), desc,
             rng.random(n) < 0.3, node2)
    perm = rng.permutation(n)
    disp = rng.uniform(2, 40, n)
    d1 = flip(desc[perm], rng, (0.0, 0.03, 0.08))
    d1[::7] = desc[perm][::7]                           # exact copies: ties between equal distances
    rot = np.where(rng.random(n) < 0.8, 15.0, rng.random(n) * 360)
    kf1 = kf(kps["x"][perm] + disp, kps["y"][perm] + rng.normal(0, 1.2, n) * sf[octave[perm]], octave[perm],
             np.mod(kps["angle"][perm] + rot, 360), np.where(rng.random(n) < stereo_p, 1.0, -1.0), d1, rng.random(n) < 0.3,
             np.where(rng.random(n) < 0.9, node2[perm], rng.integers(0, 60, n)))
    F12 = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32)
    ex, ey = 320.0, 240.0                              # epipole inside the image: the monocular distance gate fires
    return kf1, kf2, F12, ex, ey, sf, (sf * sf).astype(np.float32)


WINDOW_CASES = [("kitti", 0, "forward", True, 100), ("kitti", 1, "backward", True, 100), ("tum", 2, "band", True, 100),
                ("euroc", 3, "band", False, 100), ("tum", 4, "kf", True, 64), ("kitti", 5, "kf", False, 100)]


def window_case(oracle_mod, shape_name, seed, mode):
    frame, kps = frame_view(oracle_mod, shape_name, seed, occupied_p=0.15, stereo=(mode != "kf"))
    q = window_queries(frame, kps, 1500, seed, mode)
    return frame, q


SCW_CASES = [("tum", 0), ("kitti", 1)]


def scw_case(oracle_mod, shape_name, seed):
    frame, kps = frame_view(oracle_mod, shape_name, seed, occupied_p=0.3)
    return frame, projected_points(frame, 2500, seed, th=10.0)


# Key frames of a DISTORTED camera: the image bounds are non-integral and negative (src/Frame.cc:630-668), so the KeyFrame's
# int bounds (include/KeyFrame.h:331-332, used by GetFeaturesInArea's window) differ from the Frame's float bounds its grid
# was built with (src/KeyFrame.cc:53-59).  Object-level harness only (tests/test_gpu_adapter_objects.py): (shape, seed, kind).
SHIFTED_BOUNDS = (-13.4, -7.7, 11.3, 9.6)   # min_x, min_y, max_x - cols, max_y - rows
SHIFTED_CASES = [("tum", 6, "scw"), ("tum", 7, "fuse")]


def shifted_case(oracle_mod, shape_name, seed, kind):
    frame, kps = frame_view(oracle_mod, shape_name, seed, occupied_p=0.3 if kind == "scw" else 0.1)
    rows, cols = synth.SHAPES[shape_name]
    mnx, mny = np.float32(SHIFTED_BOUNDS[0]), np.float32(SHIFTED_BOUNDS[1])
    mxx, mxy = np.float32(cols + SHIFTED_BOUNDS[2]), np.float32(rows + SHIFTED_BOUNDS[3])
    frame["min_x"], frame["min_y"], frame["max_x"], frame["max_y"] = float(mnx), float(mny), float(mxx), float(mxy)
    frame["w_inv"] = float(np.float32(64.0) / (mxx - mnx))
    frame["h_inv"] = float(np.float32(48.0) / (mxy - mny))
    frame["cell_start"], frame["cell_items"] = oracle_mod.assign_grid(frame["x"], frame["y"], frame["min_x"], frame["min_y"],
                                                                      frame["w_inv"], frame["h_inv"])
    q = projected_points(frame, 2500, seed, th=10.0 if kind == "scw" else 3.0)
    # a share of the projections right at the low bounds, where float and truncated bounds pick different cell windows
    rng = np.random.Generator(np.random.PCG64(seed + 31))
    edge = rng.random(len(q["x"])) < 0.15
    q["x"] = np.where(edge, frame["x"][rng.integers(0, len(frame["x"]), len(edge))] * np.float32(0.05), q["x"]).astype(np.float32)
    return frame, q


FUSE_CASES = [("tum", 0, 1, False), ("euroc", 1, 1, True), ("kitti", 2, 0, False), ("tum", 3, 0, True)]


def fuse_case(oracle_mod, shape_name, seed, stereo):
    frame, kps = frame_view(oracle_mod, shape_name, seed, stereo=stereo)
    q = projected_points(frame, 3000, seed, th=3.0, stereo=stereo)
    sf = frame["scale_factors"]
    inv_sigma2 = (np.float32(1.0) / (sf * sf)).astype(np.float32)
    return frame, q, inv_sigma2


SIM3_CASES = [("tum", 4), ("kitti", 5)]


def sim3_case(oracle_mod, shape_name, seed):
    kf1, kps = frame_view(oracle_mod, shape_name, seed)
    rng = np.random.default_rng(seed)
    kf2, perm = second_view(oracle_mod, kf1, rng)
    n = len(kf1["x"])
    inv = np.empty(n, np.int64)
    inv[perm] = np.arange(n)          # KF1 feature i is KF2 feature inv[i]
    sf = kf1["scale_factors"]

    def queries(src, dst, dst_of_src):
        pred = np.clip(src["octave"] + rng.integers(-1, 2, n), 0, 7).astype(np.int32)
        x = np.clip(dst["x"][dst_of_src] + rng.normal(0, 2.0, n), 1.0, dst["max_x"] - 2.0).astype(np.float32)
        y = np.clip(dst["y"][dst_of_src] + rng.normal(0, 2.0, n), 1.0, dst["max_y"] - 2.0).astype(np.float32)
        return dict(x=x, y=y, xr=None, invz=rng.choice(INVZ, n).astype(np.float32), level=pred,
                    radius=(np.float32(7.5) * sf[pred]).astype(np.float32), min_level=pred - 1, max_level=pred,
                    valid=(rng.random(n) < 0.85).astype(np.uint8), desc=src["desc"])

    q1_in_2 = queries(kf1, kf2, inv)
    q2_in_1 = queries(kf2, kf1, perm)
    return kf1, q2_in_1, kf2, q1_in_2


def distinctive_case():
    """set sizes 0..70 incl. duplicates (median ties -> first index)."""
    rng = np.random.default_rng(3)
    sizes = np.concatenate([[0, 1, 2, 3, 33, 64, 70], rng.integers(1, 40, 500)])
    start = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    base = synth.gen_descriptors(len(sizes), seed=8)
    desc = np.zeros((start[-1], 32), np.uint8)
    for s, n in enumerate(sizes):
        flips = rng.random((n, 256)) < 0.1
        desc[start[s]:start[s + 1]] = base[s] ^ np.packbits(flips, axis=1, bitorder="little")
        if n > 4:
            desc[start[s] + 3] = desc[start[s] + 1]
    return start, desc


VOCAB_CASES = [(10, 3, 2), (10, 4, 4), (7, 5, 4), (3, 2, 4)]


def vocab_case(k, L):
    from orbslam2commentedbyxcm_b200 import vocabulary

    vocab = vocabulary.synthetic_vocabulary(k, L, seed=k * 10 + L)
    # duplicate child descriptors so that the "first minimum in child order" rule is exercised
    vocab["node_desc"][2] = vocab["node_desc"][1]
    vocab["node_desc"][k + 3] = vocab["node_desc"][k + 2]
    desc = synth.gen_descriptors(1500, seed=5)
    desc[:50] = vocab["node_desc"][1]
    return vocab, desc


STEREO_SEEDS = [0, 1, 2]
STEREO_CAM = dict(fx=435.2, fy=435.2, cx=367.2, cy=252.2, bf=47.9)


def stereo_case(seed):
    return synth.gen_stereo_pair(synth.SHAPES["euroc"], seed)


UNDISTORT_CAMS = [
    ("tum", 0, dict(fx=517.3, fy=516.5, cx=318.6, cy=255.3), [0.2624, -0.9531, -0.0054, 0.0026, 1.1633]),
    ("euroc", 1, dict(fx=458.654, fy=457.296, cx=367.215, cy=248.375), [-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05]),
    ("kitti", 2, dict(fx=718.856, fy=718.856, cx=607.19, cy=185.2), [0.0, 0.0, 0.0, 0.0]),
]


# ------------------------------------------------------------------------------------------------------------------
# one pass over every case: the same keys from the compiled reference (tools/make_golden.py -> tests/golden/
# matcher_ref.npz, tests/test_oracle_matcher.py::test_golden_is_current) and from the CUDA path (tests/test_gpu_matcher_ref.py)
# ------------------------------------------------------------------------------------------------------------------
def _norm(out):
    return np.where(out == -2, -1, out).astype(np.int32)


def reference_outputs(oracle_mod, R):
    """name -> int32/float32 array, produced by oracle/_ref/libmatchref.so (the reference's own code)."""
    out = {}
    for c in SBP_CASES:
        frame, mps = sbp_case(oracle_mod, *c)
        n, m = R.search_by_projection(frame, mps, c[3], 0.8)
        out["sbp_%s_%d" % (c[0], c[1])] = np.concatenate([[n], m]).astype(np.int32)
    frame, mps = sbp_chain_case(oracle_mod)
    n, m = R.search_by_projection(frame, mps, 15.0, 0.8)
    out["sbp_chain"] = np.concatenate([[n], m]).astype(np.int32)
    for seed, ori in BOW_CASES:
        kf, fr = bow_case(oracle_mod, seed)
        n, m = R.search_by_bow(kf, fr, 0.7, ori)
        out["bow_%d" % seed] = np.concatenate([[n], m]).astype(np.int32)
        kf1, kf2 = bow_case(oracle_mod, seed, kfkf=True)
        n, m = R.search_by_bow_kf(kf1, kf2, 0.8, ori)
        out["bowkf_%d" % seed] = np.concatenate([[n], m]).astype(np.int32)
    for shape_name, seed, ori, window in INIT_CASES:
        f1, f2, prev = init_case(oracle_mod, shape_name, seed)
        n, m, newprev = R.search_for_initialization(f1, f2, prev, window, 0.9, ori)
        out["init_%s_%d" % (shape_name, seed)] = np.concatenate([[n], m]).astype(np.int32)
        out["init_prev_%s_%d" % (shape_name, seed)] = newprev.astype(np.float32)
    for seed, ori, only_stereo, stereo_p in TRIANG_CASES:
        kf1, kf2, F12, ex, ey, sf, sigma2 = triang_case(oracle_mod, seed, stereo_p)
        n, m = R.search_for_triangulation(kf1, kf2, F12, ex, ey, only_stereo, ori)
        out["triang_%d" % seed] = np.concatenate([[n], m]).astype(np.int32)
    for shape_name, seed, mode, ori, max_dist in WINDOW_CASES:
        frame, q = window_case(oracle_mod, shape_name, seed, mode)
        if mode == "kf":
            n, m = R.search_by_projection_kf(frame, q, q["th"], max_dist, 0.9, ori)
        else:
            n, m = R.search_by_projection_last(frame, q, q["th"], {"band": 0, "forward": 1, "backward": 2}[mode], False, 0.9, ori, mbf=MBF)
        out["window_%s_%d_%s" % (shape_name, seed, mode)] = np.concatenate([[n], m]).astype(np.int32)
    for shape_name, seed in SCW_CASES:
        frame, q = scw_case(oracle_mod, shape_name, seed)
        n, m = R.search_by_projection_scw(frame, q, int(q["th"]))
        out["scw_%s_%d" % (shape_name, seed)] = np.concatenate([[n], m]).astype(np.int32)
    for shape_name, seed, gate, stereo in FUSE_CASES:
        frame, q, inv_sigma2 = fuse_case(oracle_mod, shape_name, seed, stereo)
        n, bi = R.fuse(frame, q, q["th"], mbf=MBF, scw=(gate == 0))
        out["fuse_%s_%d" % (shape_name, seed)] = np.concatenate([[n], bi]).astype(np.int32)
    for shape_name, seed in SIM3_CASES:
        kf1, q21, kf2, q12 = sim3_case(oracle_mod, shape_name, seed)
        n, m = R.search_by_sim3(kf1, q21, kf2, q12, 7.5)
        out["sim3_%s_%d" % (shape_name, seed)] = np.concatenate([[n], m]).astype(np.int32)
    for shape_name, seed, kind in SHIFTED_CASES:
        frame, q = shifted_case(oracle_mod, shape_name, seed, kind)
        if kind == "scw":
            n, m = R.search_by_projection_scw(frame, q, int(q["th"]))
        else:
            n, m = R.fuse(frame, q, q["th"], mbf=MBF, scw=False)
        out["shifted_%s_%s_%d" % (kind, shape_name, seed)] = np.concatenate([[n], m]).astype(np.int32)
    start, desc = distinctive_case()
    out["distinctive"] = R.distinctive_descriptors(start, desc)
    for k, L, levelsup in VOCAB_CASES:
        vocab, desc = vocab_case(k, L)
        voc = R.Vocabulary(vocab)
        w, nd, wt = voc.transform_features(desc, levelsup)
        out["vocab_%d_%d_word" % (k, L)] = w.astype(np.int32)
        out["vocab_%d_%d_node" % (k, L)] = nd.astype(np.int32)
        out["vocab_%d_%d_weight" % (k, L)] = wt.astype(np.float64)
        voc.close()
    cfg = synth.PRESETS["euroc"]
    cam = STEREO_CAM
    for seed in STEREO_SEEDS:
        left, right = stereo_case(seed)
        ref = R.stereo_frame(left, right, cfg, cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["bf"])
        out["stereo_%d_uright" % seed] = ref.u_right.astype(np.float32)
        out["stereo_%d_depth" % seed] = ref.depth.astype(np.float32)
        out["stereo_%d_cell_start" % seed] = ref.cell_start
        out["stereo_%d_cell_items" % seed] = ref.cell_items.astype(np.int32)
    for shape_name, seed, camd, dist in UNDISTORT_CAMS:
        img = synth.gen_frame(synth.SHAPES[shape_name], seed)
        ref = R.mono_frame(img, synth.PRESETS[shape_name], camd["fx"], camd["fy"], camd["cx"], camd["cy"], dist)
        out["mono_%s_keys_un" % shape_name] = ref.keys_un.astype(np.float32)
        out["mono_%s_bounds" % shape_name] = ref.bounds
        out["mono_%s_cell_start" % shape_name] = ref.cell_start
        out["mono_%s_cell_items" % shape_name] = ref.cell_items.astype(np.int32)
    return out
