"""Prints the per-chunk device timeline of one pipelined orbx_extract_batch call (ORBX_PIPE_TRACE) and the wall time of
untraced calls.  usage: [ORBX_PIPE_CHUNK=32] python tools/pipe_trace.py [frames]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["ORBX_PIPE_TRACE"] = "1"
import numpy as np
import torch

import orbslam2commentedbyxcm_b200 as pkg
from orbslam2commentedbyxcm_b200 import synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
frames = np.stack([synth.gen_frame((480, 640), s) for s in range(8)])
frames = np.ascontiguousarray(np.tile(frames, (n // 8, 1, 1)))
ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7, device=0)
h = torch.from_numpy(frames).pin_memory()
out = ex.alloc_outputs(n, 480, 640, pinned=True)
for i in range(3):
    sys.stderr.write("---- call %d ----\n" % i)
    sys.stderr.flush()
    t = time.perf_counter()
    ex.extract_batch(h.numpy(), out=out)
    dt = time.perf_counter() - t
    print("call %d: %.3f ms (%.0f frames/s)" % (i, dt * 1e3, n / dt))
