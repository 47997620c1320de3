"""GPU parity of the remaining ORBmatcher overloads against the CPU oracle, through the C-ABI:
SearchByProjection(KF, Scw, ...), Fuse x2, SearchBySim3, SearchForInitialization, SearchForTriangulation.  Bit-exact."""
import numpy as np
import pytest

import helpers
from orbslam2commentedbyxcm_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    import orbslam2commentedbyxcm_b200 as pkg
    return pkg


def _flip(desc, rng, rates=(0.03, 0.1, 0.3)):
    n = len(desc)
    flips = rng.random((n, 256)) < rng.choice(list(rates), n)[:, None]
    return np.ascontiguousarray(desc ^ np.packbits(flips, axis=1, bitorder="little"))


def _projected_points(frame, nq, seed, th, nlevels=8, stereo=False):
    """map points 'after projection' into a key frame: levels [pred-1, pred], radius = th * scale[pred]."""
    rng = np.random.Generator(np.random.PCG64(seed + 17))
    n = len(frame["x"])
    pick = rng.integers(0, n, nq)
    pick[nq * 3 // 4:] = pick[rng.integers(0, nq * 3 // 4, nq - nq * 3 // 4)]
    pred = np.clip(frame["octave"][pick] + rng.integers(-1, 2, nq), 0, nlevels - 1).astype(np.int32)
    x = (frame["x"][pick] + rng.normal(0, 1.5, nq)).astype(np.float32)
    y = (frame["y"][pick] + rng.normal(0, 1.5, nq)).astype(np.float32)
    xr = None
    if stereo:
        ur = frame["u_right"][pick]
        xr = np.where(ur >= 0, ur + rng.normal(0, 1.5, nq), x - 10).astype(np.float32)
    return dict(x=x, y=y, xr=xr, radius=(np.float32(th) * frame["scale_factors"][pred]).astype(np.float32),
                min_level=(pred - 1).astype(np.int32), max_level=pred.copy(), valid=(rng.random(nq) < 0.9).astype(np.uint8),
                blocks=np.ones(nq, np.uint8), angle=np.zeros(nq, np.float32), desc=_flip(frame["desc"][pick], rng))


@pytest.mark.parametrize("shape_name,seed", [("tum", 0), ("kitti", 1)])
def test_search_by_projection_kf_scw_vs_oracle(gpu, oracle_mod, shape_name, seed):
    """SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) src/ORBmatcher.cc:398-520 = orbm_search_window with
    occupied = (vpMatched[i] != NULL), every claim blocking, TH_LOW, no ratio test, no orientation check."""
    img, cfg, oex, kps, desc = helpers.extract_with_oracle(oracle_mod, shape_name, seed)
    rows, cols = synth.SHAPES[shape_name]
    frame = helpers.make_frame_view(oracle_mod, kps, desc, rows, cols, oex.scale_factors(), seed=seed, occupied_p=0.3)
    q = _projected_points(frame, 2500, seed, th=10.0)
    m = gpu.ORBmatcher(0.75, False)
    n_gpu, out_gpu = m.SearchWindow(frame, q, 50, 0)
    n_ref, out_ref = oracle_mod.search_window(frame, q, 50, 0, 0.75, False)
    assert n_ref > 200
    assert n_gpu == n_ref and np.array_equal(out_gpu, out_ref)
    m.close()


@pytest.mark.parametrize("shape_name,seed,gate,stereo", [("tum", 0, 1, False), ("euroc", 1, 1, True), ("kitti", 2, 0, False),
                                                          ("tum", 3, 0, True)])
def test_fuse_search_vs_oracle(gpu, oracle_mod, shape_name, seed, gate, stereo):
    """candidate loop of Fuse(pKF, vpMapPoints, th) (gate 1: chi2 7.8 / 5.99) and Fuse(pKF, Scw, ...) (gate 0)."""
    img, cfg, oex, kps, desc = helpers.extract_with_oracle(oracle_mod, shape_name, seed)
    rows, cols = synth.SHAPES[shape_name]
    frame = helpers.make_frame_view(oracle_mod, kps, desc, rows, cols, oex.scale_factors(), seed=seed, stereo=stereo)
    q = _projected_points(frame, 3000, seed, th=3.0, stereo=stereo)
    sf = frame["scale_factors"]
    inv_sigma2 = (np.float32(1.0) / (sf * sf)).astype(np.float32)
    m = gpu.ORBmatcher()
    bi, bd = m.FuseSearch(frame, q, gate=gate, inv_level_sigma2=inv_sigma2)
    rbi, rbd = oracle_mod.search_best(frame, q, gate, inv_sigma2, 50)
    assert (rbi >= 0).sum() > 300
    if gate == 1:   # the gate must actually reject something
        nbi, _ = oracle_mod.search_best(frame, q, 0, inv_sigma2, 50)
        assert not np.array_equal(nbi, rbi)
    assert np.array_equal(bi, rbi) and np.array_equal(bd, rbd)
    m.close()


def _second_view(oracle_mod, frame, kps, rows, cols, rng, jitter=1.0):
    """a second key frame seeing the same features: permuted, jittered, descriptors perturbed."""
    n = len(frame["x"])
    perm = rng.permutation(n)
    x = np.clip(frame["x"][perm] + rng.normal(0, jitter, n), 0, cols - 1).astype(np.float32)
    y = np.clip(frame["y"][perm] + rng.normal(0, jitter, n), 0, rows - 1).astype(np.float32)
    cs, ci = oracle_mod.assign_grid(x, y, frame["min_x"], frame["min_y"], frame["w_inv"], frame["h_inv"])
    f2 = dict(frame)
    f2.update(x=x, y=y, octave=frame["octave"][perm].copy(), desc=_flip(frame["desc"][perm], rng, (0.03, 0.08)),
              cell_start=cs, cell_items=ci, u_right=None, occupied=np.zeros(n, np.uint8))
    return f2, perm


@pytest.mark.parametrize("shape_name,seed", [("tum", 4), ("kitti", 5)])
def test_search_by_sim3_vs_oracle(gpu, oracle_mod, shape_name, seed):
    img, cfg, oex, kps, desc = helpers.extract_with_oracle(oracle_mod, shape_name, seed)
    rows, cols = synth.SHAPES[shape_name]
    rng = np.random.default_rng(seed)
    kf1 = helpers.make_frame_view(oracle_mod, kps, desc, rows, cols, oex.scale_factors(), seed=seed)
    kf1["u_right"] = None
    kf2, perm = _second_view(oracle_mod, kf1, kps, rows, cols, rng)
    n = len(kf1["x"])
    inv = np.empty(n, np.int64)
    inv[perm] = np.arange(n)          # KF1 feature i is KF2 feature inv[i]
    sf = kf1["scale_factors"]

    def queries(src, dst, dst_of_src):
        pred = np.clip(src["octave"] + rng.integers(-1, 2, n), 0, 7).astype(np.int32)
        return dict(x=(dst["x"][dst_of_src] + rng.normal(0, 2.0, n)).astype(np.float32),
                    y=(dst["y"][dst_of_src] + rng.normal(0, 2.0, n)).astype(np.float32), xr=None,
                    radius=(np.float32(7.5) * sf[pred]).astype(np.float32), min_level=pred - 1, max_level=pred,
                    valid=(rng.random(n) < 0.85).astype(np.uint8), desc=src["desc"])

    q1_in_2 = queries(kf1, kf2, inv)
    q2_in_1 = queries(kf2, kf1, perm)
    m = gpu.ORBmatcher()
    n_gpu, out_gpu = m.SearchBySim3(kf1, q2_in_1, kf2, q1_in_2)
    n_ref, out_ref = oracle_mod.search_by_sim3(kf1, q2_in_1, kf2, q1_in_2)
    assert n_ref > 200
    assert n_gpu == n_ref and np.array_equal(out_gpu, out_ref)
    m.close()


@pytest.mark.parametrize("shape_name,seed,check_ori,window", [("tum", 6, True, 100), ("tum", 7, False, 100), ("kitti", 8, True, 40),
                                                              ("euroc", 9, True, 100)])
def test_search_for_initialization_vs_oracle(gpu, oracle_mod, shape_name, seed, check_ori, window):
    """F1 holds near-duplicate features (same place, few flipped bits) so that later features steal F2 features from
    earlier ones (vMatchedDistance / vnMatches21) and the fixed-point sweeps of k_init_search have work to do."""
    img, cfg, oex, kps, desc = helpers.extract_with_oracle(oracle_mod, shape_name, seed)
    rows, cols = synth.SHAPES[shape_name]
    rng = np.random.default_rng(seed)
    base = helpers.make_frame_view(oracle_mod, kps, desc, rows, cols, oex.scale_factors(), seed=seed)
    f2, perm = _second_view(oracle_mod, base, kps, rows, cols, rng, jitter=3.0)
    rot = np.where(rng.random(len(perm)) < 0.8, 25.0, rng.random(len(perm)) * 360)
    f2["angle"] = np.mod(kps["angle"][perm] + rot, 360).astype(np.float32)
    # F1 = the extracted features, each level-0 one followed by 0-3 near-duplicates
    lvl0 = np.nonzero(base["octave"] == 0)[0]
    extra = np.repeat(lvl0, rng.integers(0, 4, len(lvl0)))
    order = rng.permutation(np.concatenate([np.arange(len(kps)), extra]))
    f1 = dict(octave=base["octave"][order].copy(), angle=kps["angle"][order].astype(np.float32),
              desc=_flip(base["desc"][order], rng, (0.0, 0.02, 0.05)))
    prev = np.stack([base["x"][order], base["y"][order]], axis=1).astype(np.float32)
    m = gpu.ORBmatcher(0.9, check_ori)
    n_gpu, out_gpu, new_prev = m.SearchForInitialization(f1, f2, prev, window)
    n_ref, out_ref = oracle_mod.search_for_initialization(f1, f2, prev, window, 0.9, check_ori)
    assert n_ref > 50
    assert n_gpu == n_ref and np.array_equal(out_gpu, out_ref)
    ok = out_ref >= 0
    assert len(set(out_ref[ok].tolist())) == ok.sum()          # an F2 feature is held by one F1 feature at most
    assert np.array_equal(new_prev[ok, 0], f2["x"][out_ref[ok]]) and np.array_equal(new_prev[~ok], prev[~ok])
    m.close()


def _kf_with_featvec(x, y, octave, angle, ur, desc, has_mp, node):
    ids = np.unique(node)
    start, feat = [0], []
    for nid in ids:
        f = np.nonzero(node == nid)[0]
        feat.extend(f.tolist())
        start.append(len(feat))
    return dict(x=x.astype(np.float32), y=y.astype(np.float32), octave=octave.astype(np.int32), angle=angle.astype(np.float32),
                u_right=ur.astype(np.float32), desc=np.ascontiguousarray(desc), has_mp=has_mp.astype(np.uint8),
                node_id=ids.astype(np.int32), node_start=np.array(start, np.int32), node_feat=np.array(feat, np.int32))


@pytest.mark.parametrize("seed,check_ori,only_stereo,stereo_p", [(0, True, False, 0.0), (1, True, False, 0.5), (2, False, True, 0.6),
                                                                 (3, True, False, 1.0)])
def test_search_for_triangulation_vs_oracle(gpu, oracle_mod, seed, check_ori, only_stereo, stereo_p):
    """rectified pair (F12 of a pure x translation: epipolar lines are rows), few vocabulary nodes so that nodes hold
    many features and the vbMatched2 claims and the last-of-equal-distances rule are exercised."""
    img, cfg, oex, kps, desc = helpers.extract_with_oracle(oracle_mod, "tum", seed + 20)
    rows, cols = synth.SHAPES["tum"]
    rng = np.random.default_rng(seed)
    n = len(kps)
    sf = np.asarray(oex.scale_factors(), np.float32)
    sigma2 = (sf * sf).astype(np.float32)
    octave = kps["octave"].astype(np.int32)
    node2 = rng.integers(0, 60, n)
    kf2 = _kf_with_featvec(kps["x"], kps["y"], octave, kps["angle"], np.where(rng.random(n) < stereo_p, kps["x"] - 5, -1),
                           desc, rng.random(n) < 0.3, node2)
    perm = rng.permutation(n)
    disp = rng.uniform(2, 40, n)
    d1 = _flip(desc[perm], rng, (0.0, 0.03, 0.08))
    d1[::7] = desc[perm][::7]                           # exact copies: ties between equal distances
    rot = np.where(rng.random(n) < 0.8, 15.0, rng.random(n) * 360)
    kf1 = _kf_with_featvec(kps["x"][perm] + disp, kps["y"][perm] + rng.normal(0, 1.2, n) * sf[octave[perm]], octave[perm],
                           np.mod(kps["angle"][perm] + rot, 360), np.where(rng.random(n) < stereo_p, 1.0, -1.0), d1,
                           rng.random(n) < 0.3, np.where(rng.random(n) < 0.9, node2[perm], rng.integers(0, 60, n)))
    F12 = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32)
    ex, ey = 320.0, 240.0                              # epipole inside the image: the monocular distance gate fires
    m = gpu.ORBmatcher(0.6, check_ori)
    n_gpu, out_gpu = m.SearchForTriangulation(kf1, kf2, F12, ex, ey, sf, sigma2, only_stereo)
    n_ref, out_ref = oracle_mod.search_for_triangulation(kf1, kf2, F12, ex, ey, sf, sigma2, only_stereo, check_ori)
    assert n_ref > 30
    assert n_gpu == n_ref and np.array_equal(out_gpu, out_ref)
    ok = out_ref >= 0
    assert len(set(out_ref[ok].tolist())) == ok.sum()
    m.close()
