"""2000 x 1M brute force: tensor-core kernel against the integer-pipe kernel, CUDA-event timed on the matcher stream."""
import json
import sys
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import orbslam2commentedbyxcm_b200 as pkg
from orbslam2commentedbyxcm_b200 import synth

ndb, nq = 1_000_000, 2000
db = synth.gen_descriptors(ndb)
q, pick = synth.gen_queries(db, nq)
dev = torch.device("cuda:0")
ddb = torch.from_numpy(db).to(dev)
dq = torch.from_numpy(q).to(dev)
m = pkg.ORBmatcher()
st = torch.cuda.ExternalStream(m.stream)
t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
t0.record(st)
prep = m.PrepareDatabase(d_db=ddb.data_ptr(), ndb=ndb)
t1.record(st); t1.synchronize()
out = {"db_expand_ms": t0.elapsed_time(t1)}
res = {}
names = ("tc",) if "--only-tc" in sys.argv else ("tc", "int")
for name in names:
    b = torch.empty(nq, dtype=torch.int32, device=dev); i = torch.empty_like(b); s = torch.empty_like(b)
    def run():
        if name == "tc":
            m.bruteforce_tc_device(prep, dq.data_ptr(), nq, 0, b.data_ptr(), i.data_ptr(), s.data_ptr())
        else:
            m.bruteforce_device(dq.data_ptr(), nq, ddb.data_ptr(), ndb, 0, b.data_ptr(), i.data_ptr(), s.data_ptr())
    for _ in range(3):
        run()
    m.sync()
    ts = []
    for _ in range(10):
        t0.record(st); run(); t1.record(st); t1.synchronize()
        ts.append(t0.elapsed_time(t1))
    res[name] = (b.cpu().numpy(), i.cpu().numpy(), s.cpu().numpy())
    out[name + "_ms"] = float(np.median(ts)); out[name + "_ms_min"] = float(np.min(ts))
    out[name + "_pairs_per_s"] = nq * ndb / (np.median(ts) * 1e-3)
out["equal"] = all(np.array_equal(a, c) for a, c in zip(res["tc"], res["int"])) if "int" in res else None
out["tc_int8_TOPs"] = 2.0 * 2048 * 1_000_000 * 256 / (out["tc_ms"] * 1e-3) / 1e12
print(json.dumps(out))
