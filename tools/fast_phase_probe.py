import sys, numpy as np, torch
sys.path.insert(0, '.')
import orbslam2commentedbyxcm_b200 as pkg
from orbslam2commentedbyxcm_b200 import synth
B = 256
frames = np.stack([synth.gen_frame((480, 640), s) for s in range(32)])
frames = np.ascontiguousarray(np.tile(frames, (B // 32, 1, 1)))
for cfg in [(1000, 1.2, 8, 20, 7), (1000, 1.2, 8, 20, 20), (1000, 1.2, 8, 7, 7)]:
    ex = pkg.ORBextractor(*cfg, device=0)
    cap = ex.max_keypoints(480, 640)
    d = torch.from_numpy(frames).cuda()
    k = torch.empty((B, cap, 6), dtype=torch.float32, device='cuda'); ds = torch.empty((B, cap, 32), dtype=torch.uint8, device='cuda'); c = torch.empty(B, dtype=torch.int32, device='cuda')
    ex.set_profiling(True)
    acc = {}
    for i in range(6):
        ex.extract_batch_device(d.data_ptr(), B, 480, 640, k.data_ptr(), ds.data_ptr(), c.data_ptr(), cap); ex.sync()
        if i:
            for kk, v in ex.stage_times().items(): acc[kk] = acc.get(kk, 0) + v / 5
    print(cfg, {kk: round(v, 4) for kk, v in acc.items()}, 'kps', float(c.float().mean()))
