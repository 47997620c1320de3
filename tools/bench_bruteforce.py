"""Times the brute-force Hamming kernel variants (ORBX_BF_VARIANT = 0 plain 8 POPC, 1 = 6 POPC, 2 = 4 POPC carry-save tree)
on 2000 x 1M descriptors, device resident.  Run on the GPU box: python tools/bench_bruteforce.py"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

if len(sys.argv) > 1:
    import numpy as np
    import torch

    import orbslam2commentedbyxcm_b200 as pkg
    from orbslam2commentedbyxcm_b200 import synth

    nq, ndb = 2000, 1_000_000
    db = synth.gen_descriptors(ndb)
    q, _ = synth.gen_queries(db, nq)
    dev = torch.device("cuda:0")
    ddb, dq = torch.from_numpy(db).to(dev), torch.from_numpy(q).to(dev)
    b = torch.empty(nq, dtype=torch.int32, device=dev)
    i, s = torch.empty_like(b), torch.empty_like(b)
    m = pkg.ORBmatcher()
    st = torch.cuda.ExternalStream(m.stream, device=dev)
    for _ in range(3):
        m.bruteforce_device(dq.data_ptr(), nq, ddb.data_ptr(), ndb, 0, b.data_ptr(), i.data_ptr(), s.data_ptr())
    m.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(20):
        m.bruteforce_device(dq.data_ptr(), nq, ddb.data_ptr(), ndb, 0, b.data_ptr(), i.data_ptr(), s.data_ptr())
    e1.record(st)
    m.sync()
    ms = e0.elapsed_time(e1) / 20
    chk = int(b.sum().item()) ^ int(i.sum().item()) ^ int(s.sum().item())
    print("variant %s: %.3f ms  %.3e pairs/s  checksum %d" % (os.environ.get("ORBX_BF_VARIANT"), ms, nq * ndb / (ms * 1e-3), chk))
else:
    for v in ("0", "1", "3", "4", "2"):
        subprocess.call([sys.executable, __file__, "run"], env=dict(os.environ, ORBX_BF_VARIANT=v))
