import sys, time, threading, numpy as np, torch
sys.path.insert(0, '.')
import pynvml
import orbslam2commentedbyxcm_b200 as pkg
from orbslam2commentedbyxcm_b200 import synth
n = 256
nd = int(sys.argv[1]) if len(sys.argv) > 1 else 8
frames = np.stack([synth.gen_frame((480, 640), s) for s in range(nd)])
frames = np.ascontiguousarray(np.tile(frames, (n // nd, 1, 1)))
ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7, device=0)
h = torch.from_numpy(frames).pin_memory()
out = ex.alloc_outputs(n, 480, 640, pinned=True)
for i in range(5): ex.extract_batch(h.numpy(), out=out)
def run(tag):
    ts = []
    for i in range(40):
        t = time.perf_counter(); ex.extract_batch(h.numpy(), out=out); ts.append(time.perf_counter() - t)
    ts = np.array(ts) * 1e3
    print("%s: e2e ms median %.3f mean %.3f max %.3f -> %.0f frames/s (mean)" % (tag, np.median(ts), ts.mean(), ts.max(), n / ts.mean() * 1e3))
run("no sampler")
pynvml.nvmlInit(); hd = pynvml.nvmlDeviceGetHandleByIndex(0)
stop = False; q = []
def poll():
    while not stop:
        t = time.perf_counter(); pynvml.nvmlDeviceGetClockInfo(hd, pynvml.NVML_CLOCK_SM); pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(hd); q.append(time.perf_counter() - t); time.sleep(0.01)
th = threading.Thread(target=poll, daemon=True); th.start()
run("nvml 10ms")
stop = True; th.join()
print("nvml query ms: median %.3f max %.3f" % (np.median(q) * 1e3, max(q) * 1e3))
run("no sampler again")
