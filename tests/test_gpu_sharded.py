"""Sharded brute-force search with the fused merge + exchange over peer memory (orbm_bruteforce_sharded, include/orbx.h).

* one GPU: `world` contexts of ONE process share cuda:0 (orbm_exchange_connect_local); their kernels run concurrently
  on separate streams and exchange through ordinary device memory -- exercises the protocol (slots, epochs, parity,
  repeated calls, ragged shards, empty shard) against the unsharded oracle;
* two or more GPUs (skipped otherwise): one process per GPU, CUDA IPC handles exchanged through torch.distributed,
  triples travel as NVLink peer stores."""
import os
import socket

import numpy as np
import pytest

from orbslam2commentedbyxcm_b200 import sharding, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    import orbslam2commentedbyxcm_b200 as pkg

    pkg.load_library()
    return pkg


def _shard_bounds(ndb, world, ragged):
    if not ragged:
        return [sharding.db_shard(ndb, r, world) for r in range(world)]
    # uneven split with one EMPTY shard in the middle (world == 4)
    assert world == 4
    cuts = [0, int(ndb * 0.13), int(ndb * 0.13), int(ndb * 0.71), ndb]
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


@pytest.mark.parametrize("world,ragged", [(1, False), (2, False), (3, False), (4, True)])
def test_sharded_local_contexts_equal_oracle(gpu, oracle_mod, world, ragged):
    """`world` contexts on ONE GPU.  The fused call makes every rank's kernel wait for the others' stores, which needs
    the kernels to be co-scheduled -- guaranteed with one process per GPU, not for several streams of one process
    on one device (streams share hardware work queues).  So here the two phases of the same kernel are launched
    separately (all ranks publish, then all ranks collect); the fused call is covered at world 1 here and across
    real GPUs in test_sharded_across_gpus_ipc."""
    import torch

    dev = torch.device("cuda:0")
    ndb, nq = 60_000, 777
    db = synth.gen_descriptors(ndb, seed=3)
    q, _ = synth.gen_queries(db, nq, seed=4)
    db[ndb // 2] = db[5]          # exact duplicates across shards: the smaller global index must win
    db[ndb - 1] = db[5]
    ob, oi, os_ = oracle_mod.bruteforce(q, db, threads=8)
    ms = [gpu.ORBmatcher(0.6, True, device=0) for _ in range(world)]
    ptrs = [m.exchange_create(world, r, nq)[1] for r, m in enumerate(ms)]
    for m in ms:
        m.exchange_connect_local(ptrs)
    dq = torch.from_numpy(q).to(dev)
    ddb = torch.from_numpy(db).to(dev)
    bounds = _shard_bounds(ndb, world, ragged)
    outs = [[torch.empty(nq, dtype=torch.int32, device=dev) for _ in range(3)] for _ in range(world)]
    torch.cuda.synchronize()
    for rep in range(5):  # repeated calls: epochs advance, slot sets alternate
        for o in outs:
            for t in o:
                t.fill_(-7)
        torch.cuda.synchronize()
        order = list(range(world) if rep % 2 == 0 else reversed(range(world)))
        for phase in ((3,) if world == 1 else (1, 2)):
            for r in order:
                lo, hi = bounds[r]
                ms[r].bruteforce_sharded_device(dq.data_ptr(), nq, ddb.data_ptr() + 32 * lo, hi - lo, lo,
                                                *[t.data_ptr() for t in outs[r]], phases=phase)
            for m in ms:
                m.sync()
        for r, m in enumerate(ms):
            assert not m.exchange_timed_out(), ("timeout", rep, r)
        for r in range(world):  # every rank holds the full result
            b, i, s = (t.cpu().numpy() for t in outs[r])
            assert np.array_equal(b, ob) and np.array_equal(i, oi) and np.array_equal(s, os_), (rep, r)
    for m in ms:
        m.close()


def test_sharded_rejects_unconnected_and_oversized(gpu):
    import torch

    dev = torch.device("cuda:0")
    m = gpu.ORBmatcher(0.6, True, device=0)
    t = torch.zeros(64 * 32, dtype=torch.uint8, device=dev)
    o = [torch.empty(64, dtype=torch.int32, device=dev) for _ in range(3)]
    with pytest.raises(Exception):
        m.bruteforce_sharded_device(t.data_ptr(), 64, t.data_ptr(), 64, 0, *[x.data_ptr() for x in o])  # no exchange
    m.exchange_create(2, 0, 32)
    with pytest.raises(Exception):
        m.bruteforce_sharded_device(t.data_ptr(), 16, t.data_ptr(), 64, 0, *[x.data_ptr() for x in o])  # peers not connected
    m.exchange_create(1, 0, 32)
    with pytest.raises(Exception):
        m.bruteforce_sharded_device(t.data_ptr(), 64, t.data_ptr(), 64, 0, *[x.data_ptr() for x in o])  # nq > nq_max
    m.close()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q, db, expect, ret):
    import torch
    import torch.distributed as dist

    import orbslam2commentedbyxcm_b200 as pkg

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        dev = torch.device("cuda", rank)
        nq = len(q)
        m = pkg.ORBmatcher(0.6, True, device=rank)
        m.exchange_connect_group(world, rank, nq)
        lo, hi = sharding.db_shard(len(db), rank, world)
        dq = torch.from_numpy(q).to(dev)
        dshard = torch.from_numpy(db[lo:hi]).to(dev)
        out = [torch.empty(nq, dtype=torch.int32, device=dev) for _ in range(3)]
        ok = True
        for rep in range(4):
            dist.barrier()
            m.bruteforce_sharded_device(dq.data_ptr(), nq, dshard.data_ptr(), hi - lo, lo, *[t.data_ptr() for t in out])
            m.sync()
            ok = ok and not m.exchange_timed_out()
            for t, e in zip(out, expect):
                ok = ok and np.array_equal(t.cpu().numpy(), e)
        ret[rank] = ok
        dist.barrier()
        m.close()
    finally:
        dist.destroy_process_group()


def test_sharded_across_gpus_ipc(gpu, oracle_mod):
    import torch
    import torch.multiprocessing as mp

    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs >= 2 GPUs (gpurun --gpus 2)")
    world = min(ngpu, 8)
    ndb, nq = 200_000, 500
    db = synth.gen_descriptors(ndb, seed=11)
    q, _ = synth.gen_queries(db, nq, seed=12)
    db[ndb - 2] = db[1]
    expect = oracle_mod.bruteforce(q, db, threads=8)
    mgr = mp.get_context("spawn").Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), q, db, expect, ret), nprocs=world, join=True)
    assert all(ret.get(r) for r in range(world)), dict(ret)


def test_tag_wrap_with_varying_nq(gpu, oracle_mod):
    """the 14-bit slot tag wraps every 16384 searches and nq may change from call to call.  Drive more
    than 16384 searches with nq varying between 1 and nq_max over two local contexts and check, across the wrap,
    (a) the results, and (b) the invariant that makes stale slots harmless: before search e, no slot of set e & 1 already
    carries tag e & 0x3fff (without the scrub window the slot nq_max - 1, last written 16384 searches earlier, does)."""
    import torch

    dev = torch.device("cuda:0")
    world, nq_max, ndb = 2, 48, 512
    db = synth.gen_descriptors(ndb, seed=31)
    q, _ = synth.gen_queries(db, nq_max, seed=32)
    ob, oi, os_ = oracle_mod.bruteforce(q, db, threads=1)
    ms = [gpu.ORBmatcher(0.6, True, device=0) for _ in range(world)]
    ptrs = [m.exchange_create(world, r, nq_max)[1] for r, m in enumerate(ms)]
    for m in ms:
        m.exchange_connect_local(ptrs)
    dq = torch.from_numpy(q).to(dev)
    ddb = torch.from_numpy(db).to(dev)
    bounds = [sharding.db_shard(ndb, r, world) for r in range(world)]
    outs = [[torch.empty(nq_max, dtype=torch.int32, device=dev) for _ in range(3)] for _ in range(world)]

    def search(nq):
        for phase in (1, 2):
            for r in range(world):
                lo, hi = bounds[r]
                ms[r].bruteforce_sharded_device(dq.data_ptr(), nq, ddb.data_ptr() + 32 * lo, hi - lo, lo,
                                                *[t.data_ptr() for t in outs[r]], phases=phase)

    def check_invariant():
        for m in ms:
            words, done = m.debug_exchange_snapshot(world, nq_max)
            nxt = done + 1
            tags = (words[nxt & 1] >> np.uint64(50)).astype(np.int64)
            assert not np.any(tags == (nxt & 0x3fff)), "a slot already carries the tag of search %d" % nxt

    def check_result(nq):
        torch.cuda.synchronize()
        for r in range(world):
            assert not ms[r].exchange_timed_out()
            b, i, s = (t.cpu().numpy()[:nq] for t in outs[r])
            assert np.array_equal(b, ob[:nq]) and np.array_equal(i, oi[:nq]) and np.array_equal(s, os_[:nq])

    total = 2 * 16384 + 40
    probe = {1, 2, 3, 4, 16383, 16384, 16385, 16386, 16387, 2 * 16384, 2 * 16384 + 1, 2 * 16384 + 2, 2 * 16384 + 3}
    for e in range(1, total + 1):
        # full-width searches at 1, 2 and exactly one and two wraps later; tiny ones in between (with a few odd sizes)
        if e in (1, 2, 16385, 16386, 2 * 16384 + 1, 2 * 16384 + 2):
            nq = nq_max
        else:
            nq = 1 + (e * 7) % 5
        if e in probe or e % 4099 == 0:
            check_invariant()
        search(nq)
        if e in probe or e % 4099 == 0:
            check_result(nq)
    for m in ms:
        m.close()
