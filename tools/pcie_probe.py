import torch, time
x=torch.empty(78643200,dtype=torch.uint8).pin_memory(); d=torch.empty_like(x,device='cuda')
y=torch.empty(14795776,dtype=torch.uint8).pin_memory(); dy=torch.empty_like(y,device='cuda')
for n in (78643200, 9830400, 2457600):
    for _ in range(3): d[:n].copy_(x[:n],non_blocking=True)
    torch.cuda.synchronize(); t=time.perf_counter()
    for _ in range(20): d[:n].copy_(x[:n],non_blocking=True)
    torch.cuda.synchronize(); dt=(time.perf_counter()-t)/20
    print('H2D',n,'bytes %.1f GB/s'%(n/dt/1e9))
torch.cuda.synchronize(); t=time.perf_counter()
for _ in range(20): y.copy_(dy,non_blocking=True)
torch.cuda.synchronize(); dt=(time.perf_counter()-t)/20
print('D2H %.1f GB/s'%(14795776/dt/1e9))
# chunked copies as the pipelined host path issues them: 8 x 9.8 MB back to back on one stream, on two streams, and with
# a concurrent D2H stream
def chunked(nstreams, with_d2h):
    ss = [torch.cuda.Stream() for _ in range(nstreams)]
    sd = torch.cuda.Stream()
    c = 9830400
    def run():
        for k in range(8):
            with torch.cuda.stream(ss[k % nstreams]):
                d[k * c:(k + 1) * c].copy_(x[k * c:(k + 1) * c], non_blocking=True)
            if with_d2h:
                with torch.cuda.stream(sd):
                    y[: 1849472].copy_(dy[: 1849472], non_blocking=True)
    for _ in range(3): run()
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(20): run()
    torch.cuda.synchronize(); dt = (time.perf_counter() - t) / 20
    print('H2D 8 x 9.8 MB, %d stream(s), d2h %s: %.1f GB/s' % (nstreams, with_d2h, 8 * c / dt / 1e9))
for ns in (1, 2, 4):
    for w in (False, True):
        chunked(ns, w)
# the same copies while the extractor kernels run on another stream (device-resident input): does kernel activity slow the DMA?
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import orbslam2commentedbyxcm_b200 as pkg
from orbslam2commentedbyxcm_b200 import synth
frames = np.stack([synth.gen_frame((480, 640), s) for s in range(8)])
frames = np.ascontiguousarray(np.tile(frames, (8, 1, 1)))
ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7, device=0)
dfr = torch.from_numpy(frames).cuda()
cap = ex.max_keypoints(480, 640)
dk = torch.empty((64, cap, 24), dtype=torch.uint8, device='cuda'); dd = torch.empty((64, cap, 32), dtype=torch.uint8, device='cuda'); dc = torch.empty(64, dtype=torch.int32, device='cuda')
def kern():
    ex.extract_batch_device(dfr.data_ptr(), 64, 480, 640, dk.data_ptr(), dd.data_ptr(), dc.data_ptr(), cap)
try:
    for _ in range(3): kern()
    ex.sync()
    s1 = torch.cuda.Stream(); c = 9830400
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(20):
        for k in range(8):
            kern()
            with torch.cuda.stream(s1):
                d[k * c:(k + 1) * c].copy_(x[k * c:(k + 1) * c], non_blocking=True)
    s1.synchronize(); dt = (time.perf_counter() - t) / 20
    ex.sync()
    print('H2D 8 x 9.8 MB with extractor kernels running: %.1f GB/s' % (8 * c / dt / 1e9))
except Exception as e:
    print('kernel-overlap probe skipped:', e)
