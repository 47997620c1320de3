"""Pins oracle/match_oracle.cc (the flat-array restatement the GPU parity tests compare with) to the REFERENCE'S OWN matcher
code: src/ORBmatcher.cc, src/Frame.cc (grid, undistortion, ComputeStereoMatches), src/KeyFrame.cc, src/MapPoint.cc and the
vendored DBoW2, compiled unmodified into oracle/_ref/libmatchref.so (oracle/Makefile target `ref`, harness
oracle/matchref_wrap.cc).  Every entry point, on the C2 / C3-sized inputs the GPU tests use (tests/matcher_cases.py).

Skipped when the library is absent (it is built wherever /root/reference exists and travels to the GPU box)."""
import numpy as np
import pytest

import matcher_cases as mc
from oracle import matchref as R
from orbslam2commentedbyxcm_b200 import synth, vocabulary

pytestmark = pytest.mark.skipif(not R.available(), reason="oracle/_ref/libmatchref.so not built (needs /root/reference)")


def _norm(out):
    """the oracle marks entries cleared by the orientation filter with -2; the reference leaves NULL"""
    return np.where(out == -2, -1, out)


def test_descriptor_distance(oracle_mod):
    db = synth.gen_descriptors(300)
    L = oracle_mod.lib()
    for i in range(0, 299):
        assert R.descriptor_distance(db[i], db[i + 1]) == L.orbo_descriptor_distance(db[i].ctypes.data, db[i + 1].ctypes.data)
    assert R.descriptor_distance(db[0], db[0]) == 0
    assert R.descriptor_distance(np.zeros(32, np.uint8), np.full(32, 255, np.uint8)) == 256


def _area_numpy(frame, x, y, r, lo, hi):
    """Frame::GetFeaturesInArea (src/Frame.cc:488-548) in float32 numpy over the CSR grid, reference visiting order."""
    f = np.float32
    x, y, r = f(x), f(y), f(r)
    c0 = max(0, int(np.floor((x - f(frame["min_x"]) - r) * f(frame["w_inv"]))))
    c1 = min(63, int(np.ceil((x - f(frame["min_x"]) + r) * f(frame["w_inv"]))))
    r0 = max(0, int(np.floor((y - f(frame["min_y"]) - r) * f(frame["h_inv"]))))
    r1 = min(47, int(np.ceil((y - f(frame["min_y"]) + r) * f(frame["h_inv"]))))
    if c0 >= 64 or c1 < 0 or r0 >= 48 or r1 < 0:
        return []
    check = lo > 0 or hi >= 0
    out = []
    for ix in range(c0, c1 + 1):
        for iy in range(r0, r1 + 1):
            c = ix * 48 + iy
            for i in frame["cell_items"][frame["cell_start"][c]:frame["cell_start"][c + 1]]:
                o = frame["octave"][i]
                if check and (o < lo or (hi >= 0 and o > hi)):
                    continue
                if abs(frame["x"][i] - x) < r and abs(frame["y"][i] - y) < r:
                    out.append(int(i))
    return out


@pytest.mark.parametrize("shape_name", ["tum", "euroc", "kitti"])
def test_assign_grid_and_features_in_area(oracle_mod, shape_name):
    frame, kps = mc.frame_view(oracle_mod, shape_name, 0)
    cs, ci = R.assign_grid(frame)
    assert np.array_equal(cs, frame["cell_start"]) and np.array_equal(ci, frame["cell_items"])
    rng = np.random.default_rng(1)
    for _ in range(60):
        x, y = float(rng.uniform(-30, frame["max_x"] + 30)), float(rng.uniform(-30, frame["max_y"] + 30))
        r = float(rng.choice([3.0, 10.0, 40.0]))
        lo, hi = int(rng.integers(-1, 4)), int(rng.integers(-1, 8))
        assert R.features_in_area(frame, x, y, r, lo, hi).tolist() == _area_numpy(frame, x, y, r, lo, hi)
        assert R.features_in_area(frame, x, y, r, keyframe=True).tolist() == _area_numpy(frame, x, y, r, -1, -1)


@pytest.mark.parametrize("shape_name,seed,nmp,th,stereo", mc.SBP_CASES)
def test_search_by_projection(oracle_mod, shape_name, seed, nmp, th, stereo):
    frame, mps = mc.sbp_case(oracle_mod, shape_name, seed, nmp, th, stereo)
    n, out = oracle_mod.search_by_projection(frame, mps, th, 0.8)
    rn, rout = R.search_by_projection(frame, mps, th, 0.8)
    assert rn > 0.1 * nmp
    assert n == rn and np.array_equal(out, rout)


def test_search_by_projection_claim_chain(oracle_mod):
    frame, mps = mc.sbp_chain_case(oracle_mod)
    n, out = oracle_mod.search_by_projection(frame, mps, 15.0, 0.8)
    rn, rout = R.search_by_projection(frame, mps, 15.0, 0.8)
    assert n == rn and np.array_equal(out, rout)


@pytest.mark.parametrize("seed,check_ori", mc.BOW_CASES)
def test_search_by_bow_kf_frame(oracle_mod, seed, check_ori):
    kf, fr = mc.bow_case(oracle_mod, seed)
    n, out = oracle_mod.search_by_bow(kf, fr, 0.7, check_ori)
    rn, rout = R.search_by_bow(kf, fr, 0.7, check_ori)
    assert rn > 50
    assert n == rn and np.array_equal(out, rout)


@pytest.mark.parametrize("seed,check_ori", mc.BOW_CASES)
def test_search_by_bow_kf_kf(oracle_mod, seed, check_ori):
    kf1, kf2 = mc.bow_case(oracle_mod, seed, kfkf=True)
    n, out = oracle_mod.search_by_bow_kf(kf1, kf2, 0.8, check_ori)
    rn, rout = R.search_by_bow_kf(kf1, kf2, 0.8, check_ori)
    assert rn > 50
    assert n == rn and np.array_equal(out, rout)


@pytest.mark.parametrize("shape_name,seed,check_ori,window", mc.INIT_CASES)
def test_search_for_initialization(oracle_mod, shape_name, seed, check_ori, window):
    f1, f2, prev = mc.init_case(oracle_mod, shape_name, seed)
    n, out = oracle_mod.search_for_initialization(f1, f2, prev, window, 0.9, check_ori)
    rn, rout, rprev = R.search_for_initialization(f1, f2, prev, window, 0.9, check_ori)
    assert rn > 50
    assert n == rn and np.array_equal(out, rout)
    ok = rout >= 0   # vbPrevMatched is updated to the matched F2 position (src/ORBmatcher.cc:678-680)
    assert np.array_equal(rprev[ok, 0], f2["x"][rout[ok]]) and np.array_equal(rprev[ok, 1], f2["y"][rout[ok]])
    assert np.array_equal(rprev[~ok], prev[~ok])


@pytest.mark.parametrize("seed,check_ori,only_stereo,stereo_p", mc.TRIANG_CASES)
def test_search_for_triangulation(oracle_mod, seed, check_ori, only_stereo, stereo_p):
    kf1, kf2, F12, ex, ey, sf, sigma2 = mc.triang_case(oracle_mod, seed, stereo_p)
    n, out = oracle_mod.search_for_triangulation(kf1, kf2, F12, ex, ey, sf, sigma2, only_stereo, check_ori)
    rn, rout = R.search_for_triangulation(kf1, kf2, F12, ex, ey, only_stereo, check_ori)
    assert rn > 30
    assert n == rn and np.array_equal(out, rout)


@pytest.mark.parametrize("shape_name,seed,mode,check_ori,max_dist", mc.WINDOW_CASES)
def test_search_by_projection_last_frame_and_keyframe(oracle_mod, shape_name, seed, mode, check_ori, max_dist):
    """SearchByProjection(Cur, Last, th, bMono) in its three level patterns and SearchByProjection(Cur, KF, found, th,
    ORBdist), against the oracle's shared windowed search."""
    frame, q = mc.window_case(oracle_mod, shape_name, seed, mode)
    n, out = oracle_mod.search_window(frame, q, max_dist, 0, 0.9, check_ori)
    if mode == "kf":
        rn, rout = R.search_by_projection_kf(frame, q, q["th"], max_dist, 0.9, check_ori)
    else:
        motion = {"band": 0, "forward": 1, "backward": 2}[mode]
        rn, rout = R.search_by_projection_last(frame, q, q["th"], motion, False, 0.9, check_ori, mbf=mc.MBF)
    assert rn > 100
    assert n == rn and np.array_equal(_norm(out), rout)


def test_search_by_projection_last_frame_monocular(oracle_mod):
    """bMono = true: no forward / backward level pattern whatever the motion, no right-coordinate gate."""
    frame, kps = mc.frame_view(oracle_mod, "tum", 11, occupied_p=0.1, stereo=False)
    q = mc.window_queries(frame, kps, 1200, 11, "band")
    q["xr"] = None
    n, out = oracle_mod.search_window(frame, q, 100, 0, 0.9, True)
    rn, rout = R.search_by_projection_last(frame, q, q["th"], 1, True, 0.9, True, mbf=mc.MBF)
    assert rn > 100 and n == rn and np.array_equal(_norm(out), rout)


@pytest.mark.parametrize("shape_name,seed", mc.SCW_CASES)
def test_search_by_projection_kf_scw(oracle_mod, shape_name, seed):
    frame, q = mc.scw_case(oracle_mod, shape_name, seed)
    n, out = oracle_mod.search_window(frame, q, 50, 0, 0.75, False)
    rn, rout = R.search_by_projection_scw(frame, q, int(q["th"]))
    assert rn > 200
    assert n == rn and np.array_equal(out, rout)


@pytest.mark.parametrize("shape_name,seed,gate,stereo", mc.FUSE_CASES)
def test_fuse(oracle_mod, shape_name, seed, gate, stereo):
    """gate 1 = Fuse(pKF, vpMapPoints, th) with its chi2 test, gate 0 = Fuse(pKF, Scw, ...)."""
    frame, q, inv_sigma2 = mc.fuse_case(oracle_mod, shape_name, seed, stereo)
    bi, bd = oracle_mod.search_best(frame, q, gate, inv_sigma2, 50)
    rn, rbi = R.fuse(frame, q, q["th"], mbf=mc.MBF, scw=(gate == 0))
    assert (rbi >= 0).sum() > 300
    assert rn == (bi >= 0).sum() and np.array_equal(bi, rbi)


@pytest.mark.parametrize("shape_name,seed", mc.SIM3_CASES)
def test_search_by_sim3(oracle_mod, shape_name, seed):
    kf1, q2_in_1, kf2, q1_in_2 = mc.sim3_case(oracle_mod, shape_name, seed)
    n, out = oracle_mod.search_by_sim3(kf1, q2_in_1, kf2, q1_in_2)
    rn, rout = R.search_by_sim3(kf1, q2_in_1, kf2, q1_in_2, 7.5)
    assert rn > 200
    assert n == rn and np.array_equal(out, rout)


def test_distinctive_descriptors(oracle_mod):
    start, desc = mc.distinctive_case()
    idx = oracle_mod.distinctive_descriptors(start, desc)
    ref = R.distinctive_descriptors(start, desc)
    for s in range(len(start) - 1):
        if start[s + 1] > start[s]:
            assert np.array_equal(ref[s], desc[start[s] + idx[s]]), s


@pytest.mark.parametrize("k,L,levelsup", mc.VOCAB_CASES)
def test_vocabulary_transform(oracle_mod, k, L, levelsup):
    """the reference's own loadFromTextFile + TemplatedVocabulary::transform against the oracle's flat-tree descent, and
    Frame::ComputeBoW's BowVector / FeatureVector against the Python folding the GPU path uses (vocabulary.fold)."""
    vocab, desc = mc.vocab_case(k, L)
    voc = R.Vocabulary(vocab)
    w, nd, wt = oracle_mod.bow_transform(vocab, desc, levelsup)
    rw, rnd, rwt = voc.transform_features(desc, levelsup)
    assert np.array_equal(w, rw) and np.array_equal(nd, rnd) and np.array_equal(wt.view(np.uint64), rwt.view(np.uint64))
    w4, nd4, wt4 = oracle_mod.bow_transform(vocab, desc[100:700], 4)
    (ids, vals), fv = vocabulary.fold(w4, nd4, wt4)
    (rids, rvals), rfv = voc.compute_bow(desc[100:700])
    assert np.array_equal(ids, rids) and np.array_equal(vals.view(np.uint64), rvals.view(np.uint64))
    for key in ("node_id", "node_start", "node_feat"):
        assert np.array_equal(fv[key], rfv[key]), key
    voc.close()


@pytest.mark.parametrize("seed", mc.STEREO_SEEDS)
def test_stereo_frame_constructor(oracle_mod, seed):
    """The reference's whole stereo Frame constructor (two extractor threads, ComputeStereoMatches with the fork's in-loop
    outlier block, AssignFeaturesToGrid) against the oracle chain extract + stereo_match + assign_grid."""
    left, right = mc.stereo_case(seed)
    cam = mc.STEREO_CAM
    cfg = synth.PRESETS["euroc"]
    ref = R.stereo_frame(left, right, cfg, cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["bf"])
    oL, oR = oracle_mod.Extractor(*cfg), oracle_mod.Extractor(*cfg)
    kl, dl = oL.extract(left)
    kr, dr = oR.extract(right)
    assert ref.n == len(kl) and ref.n_right == len(kr)
    assert np.array_equal(ref.keys[:, 0], kl["x"]) and np.array_equal(ref.keys[:, 1], kl["y"])
    Lv = dict(x=kl["x"].copy(), y=kl["y"].copy(), octave=kl["octave"].astype(np.int32), desc=dl)
    Rv = dict(x=kr["x"].copy(), y=kr["y"].copy(), octave=kr["octave"].astype(np.int32), desc=dr)
    mb = float(np.float32(cam["bf"]) / np.float32(cam["fx"]))
    n, ur, dp = oracle_mod.stereo_match(oL, oR, Lv, Rv, synth.SHAPES["euroc"][0], cam["bf"], mb)
    assert (ref.u_right >= 0).sum() > 50
    assert np.array_equal(ur.view(np.uint32), ref.u_right.view(np.uint32))
    assert np.array_equal(dp.view(np.uint32), ref.depth.view(np.uint32))
    rows, cols = synth.SHAPES["euroc"]
    w_inv = float(np.float32(64.0) / np.float32(cols))
    h_inv = float(np.float32(48.0) / np.float32(rows))
    cs, ci = oracle_mod.assign_grid(kl["x"], kl["y"], 0.0, 0.0, w_inv, h_inv)
    assert np.array_equal(cs, ref.cell_start) and np.array_equal(ci, ref.cell_items)


@pytest.mark.parametrize("shape_name,seed,cam,dist", mc.UNDISTORT_CAMS)
def test_mono_frame_constructor_undistort_bounds_grid(oracle_mod, shape_name, seed, cam, dist):
    """Frame::UndistortKeyPoints + ComputeImageBounds + AssignFeaturesToGrid through the reference's monocular constructor."""
    img = synth.gen_frame(synth.SHAPES[shape_name], seed)
    cfg = synth.PRESETS[shape_name]
    ref = R.mono_frame(img, cfg, cam["fx"], cam["fy"], cam["cx"], cam["cy"], dist)
    kps, desc = oracle_mod.Extractor(*cfg).extract(img)
    assert ref.n == len(kps)
    ux, uy = oracle_mod.undistort_points(kps["x"], kps["y"], cam["fx"], cam["fy"], cam["cx"], cam["cy"], dist)
    assert np.array_equal(ux.view(np.uint32), ref.keys_un[:, 0].view(np.uint32))
    assert np.array_equal(uy.view(np.uint32), ref.keys_un[:, 1].view(np.uint32))
    rows, cols = synth.SHAPES[shape_name]
    if dist[0] != 0.0:
        cx_, cy_ = oracle_mod.undistort_points(np.array([0, cols, 0, cols], np.float32), np.array([0, 0, rows, rows], np.float32),
                                               cam["fx"], cam["fy"], cam["cx"], cam["cy"], dist)
        bounds = np.array([min(cx_[0], cx_[2]), min(cy_[0], cy_[1]), max(cx_[1], cx_[3]), max(cy_[2], cy_[3])], np.float32)
    else:
        bounds = np.array([0, 0, cols, rows], np.float32)
    assert np.array_equal(bounds, ref.bounds)
    w_inv = np.float32(64.0) / (bounds[2] - bounds[0])
    h_inv = np.float32(48.0) / (bounds[3] - bounds[1])
    cs, ci = oracle_mod.assign_grid(ux, uy, float(bounds[0]), float(bounds[1]), float(w_inv), float(h_inv))
    assert np.array_equal(cs, ref.cell_start) and np.array_equal(ci, ref.cell_items)


def test_golden_is_current(oracle_mod, golden_dir):
    """tests/golden/matcher_ref.npz (consumed by the -m gpu tests) equals what the compiled reference returns now."""
    import os

    gold = np.load(os.path.join(golden_dir, "matcher_ref.npz"))
    live = mc.reference_outputs(oracle_mod, R)
    assert sorted(gold.files) == sorted(live)
    for k, v in live.items():
        g = gold[k]
        assert g.dtype == v.dtype and g.shape == v.shape, k
        assert g.tobytes() == v.tobytes(), k
