"""GPU parity: the CUDA matchers (through the C-ABI) against the CPU oracle restatements of the reference loops.
Bit-exact: match indices, distances, second-best distances, match counts."""
import numpy as np
import pytest

from orbslam2commentedbyxcm_b200 import synth

import helpers

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    import orbslam2commentedbyxcm_b200 as pkg

    return pkg


def test_descriptor_distance(gpu, oracle_mod):
    rng = np.random.default_rng(0)
    L = oracle_mod.lib()
    for _ in range(200):
        a = rng.integers(0, 256, 32, dtype=np.uint8)
        b = rng.integers(0, 256, 32, dtype=np.uint8)
        d = gpu.ORBmatcher.DescriptorDistance(a, b)
        assert d == L.orbo_descriptor_distance(a.ctypes.data, b.ctypes.data) == helpers.popcount_dist(a, b)
    z = np.zeros(32, np.uint8)
    assert gpu.ORBmatcher.DescriptorDistance(z, ~z) == 256 and gpu.ORBmatcher.DescriptorDistance(z, z) == 0


@pytest.mark.parametrize("nq,ndb", [(1, 1), (37, 1000), (2000, 20000), (1025, 7777), (300, 0), (5, 3)])
def test_bruteforce_vs_oracle(gpu, oracle_mod, nq, ndb):
    db = synth.gen_descriptors(max(ndb, 1), seed=ndb + 1)[:ndb]
    if ndb:
        q, _ = synth.gen_queries(db, nq, seed=nq)
        # force exact duplicates and ties
        if ndb > 10:
            db[ndb // 2] = db[3]
            db[ndb - 1] = db[3]
    else:
        q = synth.gen_descriptors(nq, seed=5)
    m = gpu.ORBmatcher()
    best, idx, second = m.BruteForce(q, db)
    ob, oi, os_ = oracle_mod.bruteforce(q, db, threads=8)
    assert np.array_equal(best, ob)
    assert np.array_equal(idx, oi)
    assert np.array_equal(second, os_)
    m.close()


def test_bruteforce_low_entropy_ties(gpu, oracle_mod):
    """descriptors drawn from a tiny alphabet: many equal distances -> first-index and second-best semantics."""
    rng = np.random.default_rng(7)
    base = rng.integers(0, 256, (4, 32), dtype=np.uint8)
    db = base[rng.integers(0, 4, 5000)]
    q = base[rng.integers(0, 4, 300)]
    m = gpu.ORBmatcher()
    best, idx, second = m.BruteForce(q, db)
    ob, oi, os_ = oracle_mod.bruteforce(q, db, threads=8)
    assert np.array_equal(best, ob) and np.array_equal(idx, oi) and np.array_equal(second, os_)
    assert (best == 0).all() and (second == 0).all()
    m.close()


def test_bruteforce_full_size_properties(gpu):
    """2000 x 1M (config 5 on one GPU) through properties: the planted match is found with the planted distance,
    second >= best, and a sharded search + merge equals the unsharded one bit for bit."""
    import torch

    ndb, nq = 1_000_000, 2000
    db = synth.gen_descriptors(ndb)
    q, pick = synth.gen_queries(db, nq)
    m = gpu.ORBmatcher()
    best, idx, second = m.BruteForce(q, db)
    planted = np.unpackbits(np.bitwise_xor(q, db[pick]), axis=1).sum(1)
    assert np.array_equal(best, np.minimum(best, planted))  # never worse than the planted row
    hit = idx == pick
    assert hit.mean() > 0.99
    assert np.array_equal(best[hit], planted[hit])
    assert (second >= best).all()
    got = np.unpackbits(np.bitwise_xor(q, db[idx]), axis=1).sum(1)
    assert np.array_equal(got, best)
    # sharded: 8 contiguous shards searched separately on the device, then merged
    dev = torch.device("cuda:0")
    dq = torch.from_numpy(q).to(dev)
    ddb = torch.from_numpy(db).to(dev)
    parts = 8
    pb = torch.empty((parts, nq), dtype=torch.int32, device=dev)
    pi = torch.empty_like(pb)
    ps = torch.empty_like(pb)
    per = ndb // parts
    for s in range(parts):
        m.bruteforce_device(dq.data_ptr(), nq, ddb[s * per:].data_ptr(), per, s * per, pb[s].data_ptr(), pi[s].data_ptr(), ps[s].data_ptr())
    ob = torch.empty(nq, dtype=torch.int32, device=dev)
    oi = torch.empty_like(ob)
    os_ = torch.empty_like(ob)
    m.merge_partials_device(pb.data_ptr(), pi.data_ptr(), ps.data_ptr(), parts, nq, ob.data_ptr(), oi.data_ptr(), os_.data_ptr())
    m.sync()
    assert np.array_equal(ob.cpu().numpy(), best) and np.array_equal(oi.cpu().numpy(), idx) and np.array_equal(os_.cpu().numpy(), second)
    m.close()


@pytest.mark.parametrize("shape_name,seed,nmp,th,stereo", [
    ("kitti", 0, 4000, 1.0, False), ("kitti", 1, 4000, 3.0, True), ("tum", 2, 1500, 5.0, False), ("euroc", 3, 50, 1.0, True),
])
def test_search_by_projection_vs_oracle(gpu, oracle_mod, shape_name, seed, nmp, th, stereo):
    img, cfg, oex, kps, desc = helpers.extract_with_oracle(oracle_mod, shape_name, seed)
    rows, cols = synth.SHAPES[shape_name]
    frame = helpers.make_frame_view(oracle_mod, kps, desc, rows, cols, oex.scale_factors(), seed=seed, stereo=stereo)
    mps = helpers.make_mappoints(frame, nmp, seed=seed, stereo=stereo)
    m = gpu.ORBmatcher(0.8, True)
    n_gpu, out_gpu = m.SearchByProjection(frame, mps, th)
    n_ref, out_ref = oracle_mod.search_by_projection(frame, mps, th, 0.8)
    assert n_ref > 0.2 * nmp * 0.5
    assert n_gpu == n_ref
    assert np.array_equal(out_gpu, out_ref)
    m.close()


def test_search_by_projection_claim_chain(gpu, oracle_mod):
    """a long chain of map points that all want the same few keypoints: worst case for the fixed-point sweeps."""
    img, cfg, oex, kps, desc = helpers.extract_with_oracle(oracle_mod, "tum", 4)
    rows, cols = synth.SHAPES["tum"]
    frame = helpers.make_frame_view(oracle_mod, kps, desc, rows, cols, oex.scale_factors(), occupied_p=0.0, seed=4)
    tgt = int(np.argmax((kps["octave"] == 0)))
    nmp = 64
    rng = np.random.default_rng(0)
    mps = dict(
        x=np.full(nmp, frame["x"][tgt], np.float32), y=np.full(nmp, frame["y"][tgt], np.float32),
        xr=np.full(nmp, -1, np.float32), level=np.zeros(nmp, np.int32), view_cos=np.full(nmp, 0.9, np.float32),
        valid=np.ones(nmp, np.uint8), obs=(rng.random(nmp) < 0.7).astype(np.uint8),
        desc=np.ascontiguousarray(np.repeat(frame["desc"][tgt][None], nmp, 0)),
    )
    m = gpu.ORBmatcher(0.8, True)
    n_gpu, out_gpu = m.SearchByProjection(frame, mps, 15.0)
    n_ref, out_ref = oracle_mod.search_by_projection(frame, mps, 15.0, 0.8)
    assert n_gpu == n_ref and np.array_equal(out_gpu, out_ref)
    m.close()


@pytest.mark.parametrize("seed", [0, 1])
def test_stereo_rows_vs_oracle(gpu, oracle_mod, seed):
    left, right = synth.gen_stereo_pair(synth.SHAPES["euroc"], seed)
    cfg = synth.PRESETS["euroc"]
    oex = oracle_mod.Extractor(*cfg)
    kl, dl = oex.extract(left)
    sf = oex.scale_factors()
    kr, dr = oracle_mod.Extractor(*cfg).extract(right)
    L = dict(x=kl["x"].copy(), y=kl["y"].copy(), octave=kl["octave"].astype(np.int32), desc=dl)
    R = dict(x=kr["x"].copy(), y=kr["y"].copy(), octave=kr["octave"].astype(np.int32), desc=dr)
    fx, bf = 435.2, 47.9
    mb = bf / fx
    min_d, max_d = 0.0, float(np.float32(bf) / np.float32(mb))
    m = gpu.ORBmatcher()
    br, bd = m.StereoRows(L, R, sf, synth.SHAPES["euroc"][0], min_d, max_d)
    obr, obd = oracle_mod.stereo_rows(L, R, sf, synth.SHAPES["euroc"][0], min_d, max_d)
    assert (obr >= 0).sum() > 100
    assert np.array_equal(br, obr) and np.array_equal(bd, obd)
    m.close()


@pytest.mark.parametrize("seed,check_ori", [(0, True), (1, False), (2, True)])
def test_search_by_bow_vs_oracle(gpu, oracle_mod, seed, check_ori):
    img, cfg, oex, kps, desc = helpers.extract_with_oracle(oracle_mod, "tum", seed)
    rng = np.random.default_rng(seed)
    n = len(kps)
    # key frame = same features in shuffled order with perturbed descriptors and mostly consistent rotation
    perm = rng.permutation(n)
    flips = rng.random((n, 256)) < 0.06
    kf_desc = np.ascontiguousarray(desc[perm] ^ np.packbits(flips, axis=1, bitorder="little"))
    rot = np.where(rng.random(n) < 0.8, 24.0, rng.random(n) * 360).astype(np.float32)
    kf_angle = np.mod(kps["angle"][perm] + rot, 360).astype(np.float32)
    fvF, node_of = helpers.make_featvec(n, 400, seed)
    # the KF feature falls in the same node as its source feature most of the time
    kf_node = np.where(rng.random(n) < 0.9, node_of[perm], rng.integers(0, 400, n))
    ids = np.unique(kf_node)
    start, feat = [0], []
    for nid in ids:
        f = np.nonzero(kf_node == nid)[0]
        feat.extend(f.tolist())
        start.append(len(feat))
    kf = dict(desc=kf_desc, angle=kf_angle, has_mp=(rng.random(n) < 0.8).astype(np.uint8), node_id=ids.astype(np.int32),
              node_start=np.array(start, np.int32), node_feat=np.array(feat, np.int32))
    fr = dict(desc=np.ascontiguousarray(desc), angle=kps["angle"].astype(np.float32).copy(), **fvF)
    m = gpu.ORBmatcher(0.7, check_ori)
    n_gpu, out_gpu = m.SearchByBoW(kf, fr)
    n_ref, out_ref = oracle_mod.search_by_bow(kf, fr, 0.7, check_ori)
    assert n_ref > 50
    assert n_gpu == n_ref and np.array_equal(out_gpu, out_ref)
    m.close()


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_full_stereo_matching_vs_oracle(gpu, oracle_mod, seed):
    """Frame::ComputeStereoMatches complete: coarse search + SAD refinement on the device-resident pyramids + outlier
    removal; mvuRight / mvDepth bit-exact (float bits)."""
    left, right = synth.gen_stereo_pair(synth.SHAPES["euroc"], seed)
    cfg = synth.PRESETS["euroc"]
    oL, oR = oracle_mod.Extractor(*cfg), oracle_mod.Extractor(*cfg)
    kl, dl = oL.extract(left)
    kr, dr = oR.extract(right)
    gL, gR = gpu.ORBextractor(*cfg), gpu.ORBextractor(*cfg)
    gkl, gdl = gL(left)
    gkr, gdr = gR(right)
    assert gkl.tobytes() == kl.tobytes() and gkr.tobytes() == kr.tobytes()
    L = dict(x=kl["x"].copy(), y=kl["y"].copy(), octave=kl["octave"].astype(np.int32), desc=dl)
    R = dict(x=kr["x"].copy(), y=kr["y"].copy(), octave=kr["octave"].astype(np.int32), desc=dr)
    fx, bf = 435.2, 47.9
    mb = float(np.float32(bf) / np.float32(fx))
    n_ref, ur_ref, dp_ref = oracle_mod.stereo_match(oL, oR, L, R, synth.SHAPES["euroc"][0], bf, mb)
    m = gpu.ORBmatcher()
    ur, dp = m.ComputeStereoMatches(gL, gR, L, R, bf, mb)
    assert (ur_ref >= 0).sum() > 50
    assert np.array_equal(ur.view(np.uint32), ur_ref.view(np.uint32))
    assert np.array_equal(dp.view(np.uint32), dp_ref.view(np.uint32))
    m.close(); gL.close(); gR.close()


@pytest.mark.parametrize("k,L,levelsup", [(10, 3, 2), (10, 4, 4), (7, 5, 4), (3, 2, 4)])
def test_bow_transform_vs_oracle(gpu, oracle_mod, k, L, levelsup):
    """DBoW2 vocabulary descent on a synthetic k-ary tree: word id, node id `levelsup` above the leaves, weight."""
    from orbslam2commentedbyxcm_b200 import vocabulary

    vocab = vocabulary.synthetic_vocabulary(k, L, seed=k * 10 + L)
    # duplicate child descriptors so that the "first minimum in child order" rule is exercised
    vocab["node_desc"][2] = vocab["node_desc"][1]
    vocab["node_desc"][k + 3] = vocab["node_desc"][k + 2]
    desc = synth.gen_descriptors(1500, seed=5)
    desc[:50] = vocab["node_desc"][1]
    m = gpu.ORBmatcher()
    voc = vocabulary.ORBVocabulary(vocab, m)
    w, nd, wt = voc.transform_features(desc, levelsup)
    ow, ond, owt = oracle_mod.bow_transform(vocab, desc, levelsup)
    assert np.array_equal(w, ow) and np.array_equal(nd, ond) and np.array_equal(wt.view(np.uint64), owt.view(np.uint64))
    if L - levelsup <= 0:
        assert (nd == 0).all()
    # the folded FeatureVector feeds SearchByBoW: frame vs itself must match every non-stopped feature to itself
    (ids, vals), fv = voc.transform(desc[100:600], levelsup)
    assert abs(vals.sum() - 1.0) < 1e-9 and np.all(np.diff(ids) > 0)
    d = np.ascontiguousarray(desc[100:600])
    ang = np.zeros(len(d), np.float32)
    kf = dict(desc=d, angle=ang, has_mp=np.ones(len(d), np.uint8), **fv)
    fr = dict(desc=d, angle=ang, **fv)
    n, out = gpu.ORBmatcher(0.9, True).SearchByBoW(kf, fr)
    on, oout = oracle_mod.search_by_bow(kf, fr, 0.9, True)
    assert n == on and np.array_equal(out, oout)
    voc.close(); m.close()


def _window_queries(frame, kps, nq, seed, mode, nlevels=8):
    """source features 'after projection': mode 'forward' / 'backward' / 'band' = the three GetFeaturesInArea level
    patterns of SearchByProjection(Cur, Last), 'kf' = the (Cur, KF, found) overload (band, every claim blocks, no gate)."""
    rng = np.random.Generator(np.random.PCG64(seed + 7))
    n = len(frame["x"])
    pick = rng.integers(0, n, nq)
    pick[nq * 3 // 4:] = pick[rng.integers(0, nq * 3 // 4, nq - nq * 3 // 4)]  # several sources per keypoint
    octv = np.clip(frame["octave"][pick] + rng.integers(-1, 2, nq), 0, nlevels - 1).astype(np.int32)
    th = 7.0 if mode != "kf" else 10.0
    radius = (np.float32(th) * frame["scale_factors"][octv]).astype(np.float32)
    if mode == "forward":
        lmin, lmax = octv.copy(), np.full(nq, -1, np.int32)
    elif mode == "backward":
        lmin, lmax = np.zeros(nq, np.int32), octv.copy()
    else:
        lmin, lmax = octv - 1, octv + 1
    flips = rng.random((nq, 256)) < rng.choice([0.03, 0.1, 0.3], nq)[:, None]
    x = (frame["x"][pick] + rng.normal(0, 3.0, nq)).astype(np.float32)
    rot = np.where(rng.random(nq) < 0.75, 12.0, rng.random(nq) * 360)
    return dict(x=x, y=(frame["y"][pick] + rng.normal(0, 3.0, nq)).astype(np.float32), radius=radius,
                min_level=lmin.astype(np.int32), max_level=lmax.astype(np.int32),
                xr=None if mode == "kf" else (x - rng.random(nq).astype(np.float32) * 40).astype(np.float32),
                angle=np.mod(kps["angle"][pick] + rot, 360).astype(np.float32), valid=(rng.random(nq) < 0.9).astype(np.uint8),
                blocks=np.ones(nq, np.uint8) if mode == "kf" else (rng.random(nq) < 0.8).astype(np.uint8),
                desc=np.ascontiguousarray(frame["desc"][pick] ^ np.packbits(flips, axis=1, bitorder="little")))


@pytest.mark.parametrize("shape_name,seed,mode,check_ori,max_dist", [
    ("kitti", 0, "forward", True, 100), ("kitti", 1, "backward", True, 100), ("tum", 2, "band", True, 100),
    ("euroc", 3, "band", False, 100), ("tum", 4, "kf", True, 64), ("kitti", 5, "kf", False, 100),
])
def test_search_window_vs_oracle(gpu, oracle_mod, shape_name, seed, mode, check_ori, max_dist):
    """the windowed search behind SearchByProjection(Cur, Last, th, bMono) and (Cur, KF, found, th, ORBdist)."""
    img, cfg, oex, kps, desc = helpers.extract_with_oracle(oracle_mod, shape_name, seed)
    rows, cols = synth.SHAPES[shape_name]
    frame = helpers.make_frame_view(oracle_mod, kps, desc, rows, cols, oex.scale_factors(), seed=seed, stereo=(mode != "kf"),
                                    occupied_p=0.15)
    frame["angle"] = kps["angle"].astype(np.float32).copy()
    q = _window_queries(frame, kps, 1500, seed, mode)
    m = gpu.ORBmatcher(0.9, check_ori)
    n_gpu, out_gpu = m.SearchWindow(frame, q, max_dist, 0)
    n_ref, out_ref = oracle_mod.search_window(frame, q, max_dist, 0, 0.9, check_ori)
    assert n_ref > 100
    if check_ori:
        assert (out_ref == -2).sum() > 0
    assert n_gpu == n_ref and np.array_equal(out_gpu, out_ref)
    m.close()


def test_distinctive_descriptors_vs_oracle(gpu, oracle_mod):
    """MapPoint::ComputeDistinctiveDescriptors batched: set sizes 0..70 incl. duplicates (median ties -> first index)."""
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
    m = gpu.ORBmatcher()
    got = m.ComputeDistinctiveDescriptors(start, desc)
    ref = oracle_mod.distinctive_descriptors(start, desc)
    assert np.array_equal(got, ref)
    m.close()


@pytest.mark.parametrize("seed,check_ori", [(0, True), (1, False), (2, True)])
def test_search_by_bow_kf_kf_vs_oracle(gpu, oracle_mod, seed, check_ori):
    """SearchByBoW(KeyFrame, KeyFrame): both sides need map points, strict < TH_LOW, output indexed by KF1 feature."""
    img, cfg, oex, kps, desc = helpers.extract_with_oracle(oracle_mod, "tum", seed + 10)
    rng = np.random.default_rng(seed)
    n = len(kps)
    perm = rng.permutation(n)
    flips = rng.random((n, 256)) < 0.06
    d1 = np.ascontiguousarray(desc[perm] ^ np.packbits(flips, axis=1, bitorder="little"))
    rot = np.where(rng.random(n) < 0.8, 40.0, rng.random(n) * 360).astype(np.float32)
    a1 = np.mod(kps["angle"][perm] + rot, 360).astype(np.float32)
    fv2, node_of = helpers.make_featvec(n, 300, seed)
    node1 = np.where(rng.random(n) < 0.9, node_of[perm], rng.integers(0, 300, n))
    ids = np.unique(node1)
    start, feat = [0], []
    for nid in ids:
        f = np.nonzero(node1 == nid)[0]
        feat.extend(f.tolist())
        start.append(len(feat))
    kf1 = dict(desc=d1, angle=a1, has_mp=(rng.random(n) < 0.8).astype(np.uint8), node_id=ids.astype(np.int32),
               node_start=np.array(start, np.int32), node_feat=np.array(feat, np.int32))
    kf2 = dict(desc=np.ascontiguousarray(desc), angle=kps["angle"].astype(np.float32).copy(), has_mp=(rng.random(n) < 0.85).astype(np.uint8), **fv2)
    m = gpu.ORBmatcher(0.8, check_ori)
    n_gpu, out_gpu = m.SearchByBoWKF(kf1, kf2)
    n_ref, out_ref = oracle_mod.search_by_bow_kf(kf1, kf2, 0.8, check_ori)
    assert n_ref > 50
    assert n_gpu == n_ref and np.array_equal(out_gpu, out_ref)
    m.close()
