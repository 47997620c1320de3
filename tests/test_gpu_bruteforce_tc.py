"""Tensor-core brute force (tcgen05 int8 GEMM on +-1 expanded descriptors + best / second epilogue out of TMEM,
csrc/orbx_match_tc.cuh) against the sequential scan with ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1983-2003):
best distance, FIRST index attaining it, second-smallest distance of the multiset.  Bit-exact (integers)."""
import numpy as np
import pytest

from orbslam2commentedbyxcm_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    import orbslam2commentedbyxcm_b200 as pkg
    return pkg


# ragged everything: rows not a multiple of the 256-row tile, queries not a multiple of the 128-row tile, chunks of one tile,
# a single row, fewer rows than one column half
@pytest.mark.parametrize("nq,ndb", [(1, 1), (37, 1000), (2000, 20000), (1025, 7777), (5, 3), (128, 256), (129, 257), (300, 130),
                                    (64, 100_003), (2048, 65536)])
def test_bruteforce_tc_vs_oracle(gpu, oracle_mod, nq, ndb):
    db = synth.gen_descriptors(ndb, seed=ndb + 1)
    q, _ = synth.gen_queries(db, nq, seed=nq)
    if ndb > 10:  # exact duplicates: first index must win, second == best
        db[ndb // 2] = db[3]
        db[ndb - 1] = db[3]
    m = gpu.ORBmatcher()
    prep = m.PrepareDatabase(db)
    best, idx, second = m.BruteForceTC(q, prep)
    ob, oi, os_ = oracle_mod.bruteforce(q, db, threads=8)
    assert np.array_equal(best, ob)
    assert np.array_equal(idx, oi)
    assert np.array_equal(second, os_)
    # and the integer-pipe kernel agrees
    b2, i2, s2 = m.BruteForce(q, db)
    assert np.array_equal(best, b2) and np.array_equal(idx, i2) and np.array_equal(second, s2)
    # idx_offset shifts the reported indices only
    b3, i3, s3 = m.BruteForceTC(q, prep, idx_offset=1000)
    assert np.array_equal(b3, best) and np.array_equal(i3, idx + 1000) and np.array_equal(s3, second)
    prep.close(); m.close()


def test_bruteforce_tc_extreme_distances(gpu, oracle_mod):
    """distance 0 and 256 (dot = +-256), all-equal rows, complementary rows."""
    z = np.zeros((1, 32), np.uint8)
    o = np.full((1, 32), 255, np.uint8)
    db = np.concatenate([np.repeat(o, 300, 0), z, np.repeat(o, 211, 0)])
    q = np.concatenate([z, o, z])
    m = gpu.ORBmatcher()
    prep = m.PrepareDatabase(db)
    best, idx, second = m.BruteForceTC(q, prep)
    ob, oi, os_ = oracle_mod.bruteforce(q, db, threads=2)
    assert np.array_equal(best, ob) and np.array_equal(idx, oi) and np.array_equal(second, os_)
    assert best.tolist() == [0, 0, 0] and idx.tolist() == [300, 0, 300] and second.tolist() == [256, 0, 256]
    prep.close(); m.close()


def test_bruteforce_tc_low_entropy_ties(gpu, oracle_mod):
    rng = np.random.default_rng(7)
    base = rng.integers(0, 256, (4, 32), dtype=np.uint8)
    db = base[rng.integers(0, 4, 5000)]
    q = base[rng.integers(0, 4, 300)]
    m = gpu.ORBmatcher()
    prep = m.PrepareDatabase(db)
    best, idx, second = m.BruteForceTC(q, prep)
    ob, oi, os_ = oracle_mod.bruteforce(q, db, threads=8)
    assert np.array_equal(best, ob) and np.array_equal(idx, oi) and np.array_equal(second, os_)
    prep.close(); m.close()


def test_bruteforce_tc_full_size_equals_integer_pipe_kernel(gpu):
    """config 5 on one GPU: 2000 x 1 M, tensor-core path == integer-pipe path == planted matches."""
    ndb, nq = 1_000_000, 2000
    db = synth.gen_descriptors(ndb)
    q, pick = synth.gen_queries(db, nq)
    m = gpu.ORBmatcher()
    prep = m.PrepareDatabase(db)
    best, idx, second = m.BruteForceTC(q, prep)
    b2, i2, s2 = m.BruteForce(q, db)
    assert np.array_equal(best, b2) and np.array_equal(idx, i2) and np.array_equal(second, s2)
    got = np.unpackbits(np.bitwise_xor(q, db[idx]), axis=1).sum(1)
    assert np.array_equal(got, best) and (idx == pick).mean() > 0.99
    prep.close(); m.close()


@pytest.mark.parametrize("nq,ndb", [(2000, 20000), (129, 257), (300, 130), (1025, 7777)])
def test_bruteforce_tc_cta_pair_variant(gpu, oracle_mod, nq, ndb, monkeypatch):
    """k_bruteforce_tc2 (cta_group::2: a CTA pair multiplies 256 queries x 256 rows as one UMMA of M = 256, each CTA loading half
    of every database stage) returns the same triples; selected by ORBX_TC_PAIR=1 (measured slower than the default kernel)."""
    monkeypatch.setenv("ORBX_TC_PAIR", "1")
    db = synth.gen_descriptors(ndb, seed=ndb)
    q, _ = synth.gen_queries(db, nq, seed=nq)
    m = gpu.ORBmatcher()
    prep = m.PrepareDatabase(db)
    b, i, s = m.BruteForceTC(q, prep, idx_offset=11)
    ob, oi, os_ = oracle_mod.bruteforce(q, db, threads=8)
    assert np.array_equal(b, ob) and np.array_equal(i, oi + 11) and np.array_equal(s, os_)
    prep.close(); m.close()
