"""CPU, world_size 2 over gloo: the multi-GPU protocol of the sharded brute-force search (contiguous database shards,
global index offsets, all-gather in rank order, sequential-scan-equivalent merge) and of frame sharding.  Per-shard
searches use the CPU oracle here; on the GPU box the same protocol runs with orbm_bruteforce / orbm_merge_partials
(tests/test_gpu_matcher.py::test_bruteforce_full_size_properties checks the CUDA merge against the unsharded search)."""
import os
import socket

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402

from orbslam2commentedbyxcm_b200 import sharding, synth  # noqa: E402


def merge_partials_model(best, idx, second):
    """numpy model of k_merge_partials (csrc/orbx_match.cu): parts ordered by ascending index range."""
    nparts, nq = best.shape
    ob = np.full(nq, 256, np.int32)
    oi = np.full(nq, -1, np.int32)
    os_ = np.full(nq, 256, np.int32)
    for p in range(nparts):
        valid = idx[p] >= 0
        better = valid & (best[p] < ob)
        worse = valid & ~better
        os_ = np.where(better, np.minimum(ob, second[p]), np.where(worse, np.minimum(os_, best[p]), os_))
        oi = np.where(better, idx[p], oi)
        ob = np.where(better, best[p], ob)
    return ob, oi, os_


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ndb, nq, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle

    db = synth.gen_descriptors(ndb, seed=3)
    db[ndb // 2 + 5] = db[7]  # duplicate across the shard boundary: the lower index must win
    q, _ = synth.gen_queries(db, nq, seed=4)
    q[0] = db[7]
    lo, hi = sharding.db_shard(ndb, rank, world)
    b, i, s = oracle.bruteforce(q, db[lo:hi], threads=2)
    i = np.where(i >= 0, i + lo, i).astype(np.int32)
    tb, ti, ts = (torch.from_numpy(x.copy()) for x in (b, i, s))
    gb = [torch.empty_like(tb) for _ in range(world)]
    gi = [torch.empty_like(tb) for _ in range(world)]
    gs = [torch.empty_like(tb) for _ in range(world)]
    dist.all_gather(gb, tb)
    dist.all_gather(gi, ti)
    dist.all_gather(gs, ts)
    ob, oi, os_ = merge_partials_model(torch.stack(gb).numpy(), torch.stack(gi).numpy(), torch.stack(gs).numpy())
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), b=ob, i=oi, s=os_)
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_bruteforce_protocol_world2(tmp_path, oracle_mod):
    ndb, nq, world = 3001, 64, 2
    mp.spawn(_worker, args=(world, _free_port(), ndb, nq, str(tmp_path)), nprocs=world, join=True)
    db = synth.gen_descriptors(ndb, seed=3)
    db[ndb // 2 + 5] = db[7]
    q, _ = synth.gen_queries(db, nq, seed=4)
    q[0] = db[7]
    rb, ri, rs = oracle_mod.bruteforce(q, db)
    for r in range(world):
        g = np.load(os.path.join(str(tmp_path), "rank%d.npz" % r))
        assert np.array_equal(g["b"], rb) and np.array_equal(g["i"], ri) and np.array_equal(g["s"], rs)
    assert ri[0] == 7 and rb[0] == 0 and rs[0] == 0


def _exchange_worker(rank, world, port, ndb, nq, calls, out_dir):
    """the fused-exchange protocol on CPU: every rank owns a slot buffer [2][world][nq_max]; per call each rank 'stores'
    its packed triples into slot set (epoch & 1) of every peer (emulated with an all-gather of the words), a reader
    accepts a slot only when it carries the current epoch, then merges in rank order."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle

    nq_max = nq + 3
    slots = np.zeros(2 * world * nq_max, np.uint64)  # epoch 0 = never written
    results = []
    for call in range(1, calls + 1):
        db = synth.gen_descriptors(ndb, seed=3 + call)
        q, _ = synth.gen_queries(db, nq, seed=4 + call)
        lo, hi = sharding.db_shard(ndb, rank, world)
        b, i, s = oracle.bruteforce(q, db[lo:hi], threads=2)
        i = np.where(i >= 0, i + lo, i).astype(np.int32)
        words = torch.from_numpy(sharding.pack_exchange_word(call, b, s, i).view(np.int64).copy())
        gathered = [torch.empty_like(words) for _ in range(world)]
        dist.all_gather(gathered, words)
        stale = slots.copy()
        for src in range(world):
            o = sharding.exchange_slot_offset(call, src, world, nq_max)
            slots[o:o + nq] = gathered[src].numpy().view(np.uint64)
        parts = []
        for src in range(world):
            o = sharding.exchange_slot_offset(call, src, world, nq_max)
            ep_old = sharding.unpack_exchange_word(stale[o:o + nq])[0]
            assert (ep_old != (call & 0x3FFF)).all()        # what was there before (call - 2 or never written) is not mistaken for ready
            ep, pb, ps, pi = sharding.unpack_exchange_word(slots[o:o + nq])
            assert (ep == (call & 0x3FFF)).all()
            parts.append((pb, pi, ps))
        ob, oi, os_ = merge_partials_model(*(np.stack([p[k] for p in parts]).astype(np.int32) for k in range(3)))
        results.append((ob, oi, os_))
    np.savez(os.path.join(out_dir, "xrank%d.npz" % rank), **{"c%d_%s" % (c, n): v for c, r in enumerate(results) for n, v in zip("bis", r)})
    dist.barrier()
    dist.destroy_process_group()


def test_fused_exchange_protocol_world2(tmp_path, oracle_mod):
    ndb, nq, world, calls = 2000, 40, 2, 3
    mp.spawn(_exchange_worker, args=(world, _free_port(), ndb, nq, calls, str(tmp_path)), nprocs=world, join=True)
    for call in range(1, calls + 1):
        db = synth.gen_descriptors(ndb, seed=3 + call)
        q, _ = synth.gen_queries(db, nq, seed=4 + call)
        rb, ri, rs = oracle_mod.bruteforce(q, db)
        for r in range(world):
            g = np.load(os.path.join(str(tmp_path), "xrank%d.npz" % r))
            assert np.array_equal(g["c%d_b" % (call - 1)], rb) and np.array_equal(g["c%d_i" % (call - 1)], ri)
            assert np.array_equal(g["c%d_s" % (call - 1)], rs)


def test_exchange_word_round_trip():
    rng = np.random.default_rng(1)
    b = rng.integers(0, 257, 1000).astype(np.int32)
    s = rng.integers(0, 257, 1000).astype(np.int32)
    i = rng.integers(-1, 2**31 - 1, 1000).astype(np.int32)
    i[:10] = -1
    for epoch in (1, 2, 16383, 16384, 16385, 40000):
        e, ub, us, ui = sharding.unpack_exchange_word(sharding.pack_exchange_word(epoch, b, s, i))
        assert (e == (epoch & 0x3FFF)).all() and np.array_equal(ub, b) and np.array_equal(us, s) and np.array_equal(ui, i)
    assert sharding.exchange_slot_offset(5, 1, 4, 100) != sharding.exchange_slot_offset(6, 1, 4, 100)
    assert sharding.exchange_slot_offset(5, 1, 4, 100) == sharding.exchange_slot_offset(7, 1, 4, 100)


def test_shard_plans():
    for n, w in [(10, 3), (4096, 8), (7, 8), (1_000_000, 8)]:
        ranges = [sharding.db_shard(n, r, w) for r in range(w)]
        assert ranges[0][0] == 0 and ranges[-1][1] == n
        assert all(ranges[i][1] == ranges[i + 1][0] for i in range(w - 1))
        frames = sorted(f for r in range(w) for f in sharding.frame_shard(n if n < 5000 else 100, r, w))
        assert frames == list(range(n if n < 5000 else 100))


def test_merge_model_is_associative_with_empty_parts():
    rng = np.random.default_rng(0)
    db = rng.integers(0, 4, (40, 32)).astype(np.uint8)  # heavy ties
    q = rng.integers(0, 4, (9, 32)).astype(np.uint8)
    import oracle

    rb, ri, rs = oracle.bruteforce(q, db)
    cuts = [0, 0, 13, 13, 29, 40]  # includes empty shards
    parts = []
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        b, i, s = oracle.bruteforce(q, db[lo:hi])
        parts.append((b, np.where(i >= 0, i + lo, i), s))
    ob, oi, os_ = merge_partials_model(*(np.stack([p[k] for p in parts]).astype(np.int32) for k in range(3)))
    assert np.array_equal(ob, rb) and np.array_equal(oi, ri) and np.array_equal(os_, rs)
