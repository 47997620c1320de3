import sys, time, numpy as np, torch
sys.path.insert(0, '.')
import orbslam2commentedbyxcm_b200 as pkg
from orbslam2commentedbyxcm_b200 import synth
n = 256
frames = np.stack([synth.gen_frame((480, 640), s) for s in range(8)])
frames = np.ascontiguousarray(np.tile(frames, (n // 8, 1, 1)))
ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7, device=0)
h = torch.from_numpy(frames).pin_memory()
out = ex.alloc_outputs(n, 480, 640, pinned=True)
for i in range(5): ex.extract_batch(h.numpy(), out=out)
ts = []
for i in range(40):
    t = time.perf_counter(); ex.extract_batch(h.numpy(), out=out); ts.append(time.perf_counter() - t)
ts = np.array(ts) * 1e3
print("e2e ms median %.3f min %.3f -> %.0f frames/s (median)" % (np.median(ts), ts.min(), n / np.median(ts) * 1e3))
