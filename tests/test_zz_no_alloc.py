"""GPU: orbm_bruteforce_noalloc (include/orbx.h) against orbm_bruteforce, through the C-ABI.

orbm_bruteforce_noalloc is the brute-force Hamming search for integrations that may not allocate device memory while they
run -- a tracker living inside a fixed memory budget beside another owner of the GPU, or a real-time loop that cannot
afford the device-wide stall of a cudaMalloc.  Its chunk partials sit in a 12 MiB array that is part of liborbx.so, so
it takes no context and allocates nothing; the price is a size limit and one shared scratch per device.

What is checked here: the result equals orbm_bruteforce's bit for bit (these are integer distances and indices; every
parity test of this repo is exact, so the tolerance is zero) on calls made back to back, where the second search is
smaller than the first and so leaves stale partials of the first in the array; the free device memory does not move
across the calls; and sizes beyond the fixed scratch are refused before anything is launched.

This file sorts last on purpose: pytest -x stops at the first failure, and a failure here must not hide the parity tests.
"""
import ctypes as C

import numpy as np
import pytest

from orbslam2commentedbyxcm_b200 import synth

pytestmark = pytest.mark.gpu

MAX_NQ = 4096        # ORBM_NOALLOC_MAX_NQ
MAX_NDB = 1 << 24    # ORBM_NOALLOC_MAX_NDB
ERR_CAPACITY = -4    # ORBX_ERR_CAPACITY
ERR_BAD_ARG = -1     # ORBX_ERR_BAD_ARG


def _case(nq, ndb, seed):
    db = synth.gen_descriptors(max(ndb, 1), seed=seed)[:ndb]
    if ndb:
        q, _ = synth.gen_queries(db, nq, seed=seed + 1)
        if ndb > 10:  # exact duplicates: the lowest index must win, second-best equals best
            db[ndb // 2] = db[3]
            db[ndb - 1] = db[3]
    else:
        q = synth.gen_descriptors(nq, seed=seed + 1)
    return q, db


def test_noalloc_equals_bruteforce_on_consecutive_calls():
    import torch

    import orbslam2commentedbyxcm_b200 as pkg
    from orbslam2commentedbyxcm_b200 import _lib

    L = pkg.load_library()
    dev = torch.device("cuda:0")
    m = pkg.ORBmatcher()  # only for the existing entry point; the new one takes no context
    # large, then small (stale partials of the large search stay in the scratch), then empty database, then large again
    shapes = [(2000, 50000, 11), (37, 1000, 23), (300, 0, 5), (MAX_NQ, 20000, 31), (2000, 50000, 11)]
    cases = []
    for nq, ndb, seed in shapes:
        q, db = _case(nq, ndb, seed)
        dq = torch.from_numpy(q).to(dev)
        ddb = torch.from_numpy(db).to(dev) if ndb else torch.zeros((1, 32), dtype=torch.uint8, device=dev)
        want = [torch.full((nq,), -7, dtype=torch.int32, device=dev) for _ in range(3)]
        got = [torch.full((nq,), -9, dtype=torch.int32, device=dev) for _ in range(3)]
        cases.append((nq, ndb, dq, ddb, want, got))
    torch.cuda.synchronize()
    # existing entry point first (device pointers, the context's own stream and cudaMalloc'ed scratch)
    for nq, ndb, dq, ddb, want, _ in cases:
        m.bruteforce_device(dq.data_ptr(), nq, ddb.data_ptr(), ndb, 100, *[w.data_ptr() for w in want])
    m.sync()
    # the first call on the device does the one-time symbol lookup; it is part of what must not allocate
    stream = torch.cuda.current_stream().cuda_stream
    free0, _ = torch.cuda.mem_get_info(dev)
    for nq, ndb, dq, ddb, _, got in cases:  # back to back on one stream, no synchronisation in between
        rc = L.orbm_bruteforce_noalloc(0, stream, dq.data_ptr(), nq, ddb.data_ptr(), ndb, 100, *[g.data_ptr() for g in got])
        _lib.check(rc, "orbm_bruteforce_noalloc")
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free1 >= free0, "device memory was allocated during orbm_bruteforce_noalloc: %d -> %d bytes free" % (free0, free1)
    for nq, ndb, _, _, want, got in cases:
        for name, w, g in zip(("best", "best_idx", "second"), want, got):
            assert torch.equal(w, g), "%s differs at nq = %d, ndb = %d" % (name, nq, ndb)
    # sanity of the comparison itself: the sentinels are gone and the empty database gives (256, -1, 256)
    assert int(cases[0][5][0].min()) >= 0 and int(cases[0][5][1].min()) >= 100
    assert (cases[2][5][0] == 256).all() and (cases[2][5][1] == -1).all() and (cases[2][5][2] == 256).all()
    m.close()


def test_noalloc_refuses_sizes_beyond_the_static_scratch():
    import torch

    import orbslam2commentedbyxcm_b200 as pkg

    L = pkg.load_library()
    dev = torch.device("cuda:0")
    buf = torch.zeros((8, 32), dtype=torch.uint8, device=dev)
    out = [torch.full((8,), -9, dtype=torch.int32, device=dev) for _ in range(3)]
    ptrs = [o.data_ptr() for o in out]
    # sizes are checked before any pointer is touched, so small buffers are enough to probe the limits
    assert L.orbm_bruteforce_noalloc(0, None, buf.data_ptr(), MAX_NQ + 1, buf.data_ptr(), 8, 0, *ptrs) == ERR_CAPACITY
    msg = L.orbx_last_error().decode()
    assert "orbm_bruteforce_noalloc" in msg and str(MAX_NQ) in msg
    assert L.orbm_bruteforce_noalloc(0, None, buf.data_ptr(), 8, buf.data_ptr(), MAX_NDB + 1, 0, *ptrs) == ERR_CAPACITY
    assert L.orbm_bruteforce_noalloc(0, None, None, 8, buf.data_ptr(), 8, 0, *ptrs) == ERR_BAD_ARG
    assert L.orbm_bruteforce_noalloc(64, None, buf.data_ptr(), 8, buf.data_ptr(), 8, 0, *ptrs) == ERR_BAD_ARG
    torch.cuda.synchronize()
    assert all(int((o != -9).sum()) == 0 for o in out), "a refused call wrote to its outputs"
    # and a call that fits still works afterwards (NULL stream = default stream)
    assert L.orbm_bruteforce_noalloc(0, None, buf.data_ptr(), 8, buf.data_ptr(), 8, 0, *ptrs) == 0
    torch.cuda.synchronize()
    assert (out[0] == 0).all() and (out[1] == 0).all() and (out[2] == 0).all()  # eight identical all-zero rows
