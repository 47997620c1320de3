import sys, os, time, subprocess
sys.path.insert(0, '/root/repo')
import numpy as np, torch
import orbslam2commentedbyxcm_b200 as pkg
from orbslam2commentedbyxcm_b200 import synth
ndb, nq = 1_000_000, 2000
db = synth.gen_descriptors(ndb); q, _ = synth.gen_queries(db, nq)
dev = torch.device("cuda:0")
ddb = torch.from_numpy(db).to(dev); dq = torch.from_numpy(q).to(dev)
m = pkg.ORBmatcher(); st = torch.cuda.ExternalStream(m.stream)
prep = m.PrepareDatabase(d_db=ddb.data_ptr(), ndb=ndb)
b = torch.empty(nq, dtype=torch.int32, device=dev); i = torch.empty_like(b); s = torch.empty_like(b)
def run(): m.bruteforce_tc_device(prep, dq.data_ptr(), nq, 0, b.data_ptr(), i.data_ptr(), s.data_ptr())
for _ in range(5): run()
m.sync()
for reps in (1, 10, 100, 1000, 2000):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(reps): run()
    e1.record(st); e1.synchronize()
    clk = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_throttle_reasons.active", "--format=csv,noheader"], capture_output=True, text=True).stdout.strip()
    print("reps %5d: %.4f ms per search   [%s]" % (reps, e0.elapsed_time(e1) / reps, clk), flush=True)
