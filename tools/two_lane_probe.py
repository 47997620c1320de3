"""Device-resident extraction throughput with ONE handle against TWO handles fed alternately (each handle has its own stream
and arenas; the kernels of consecutive batches then overlap: the barrier-bound quad-tree and kernel tails of one batch run
beside the pyramid / FAST of the other)."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import orbslam2commentedbyxcm_b200 as pkg
from orbslam2commentedbyxcm_b200 import synth

shape, cfg = synth.SHAPES["tum"], synth.PRESETS["tum"]
rows, cols = shape
dev = torch.device("cuda", 0)
for B in (256, 512):
    frames = bench.make_frames(shape, 32, B)
    d_imgs = torch.from_numpy(frames).to(dev)
    for nh in (1, 2, 3):
        exs = [pkg.ORBextractor(*cfg) for _ in range(nh)]
        cap = exs[0].max_keypoints(rows, cols)
        outs = [(torch.empty((B, cap, 6), dtype=torch.float32, device=dev), torch.empty((B, cap, 32), dtype=torch.uint8, device=dev),
                 torch.empty(B, dtype=torch.int32, device=dev)) for _ in range(nh)]
        def run(passes):
            for p in range(passes):
                ex, (k, d, c) = exs[p % nh], outs[p % nh]
                ex.extract_batch_device(d_imgs.data_ptr(), B, rows, cols, k.data_ptr(), d.data_ptr(), c.data_ptr(), cap)
            for ex in exs:
                ex.sync()
        run(6)
        passes = 8192 * 2 // B
        t0 = time.perf_counter(); run(passes); dt = time.perf_counter() - t0
        print("batch %d, %d handle(s): %.0f frames/s" % (B, nh, passes * B / dt), flush=True)
        for ex in exs:
            ex.close()
        del outs
