"""End-to-end frames/s of orbx_extract_batch (pinned host in, pinned host out) against the number of frames per call."""
import os, sys, time
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import orbslam2commentedbyxcm_b200 as pkg
from orbslam2commentedbyxcm_b200 import synth

shape, cfg = synth.SHAPES["tum"], synth.PRESETS["tum"]
rows, cols = shape
ex = pkg.ORBextractor(*cfg)
for n in (256, 512, 1024, 2048, 4096):
    frames = bench.make_frames(shape, 32, n)
    h = torch.from_numpy(frames).pin_memory()
    out = ex.alloc_outputs(n, rows, cols, pinned=True)
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < 0.4:
        ex.extract_batch(h.numpy(), out=out)
    reps = max(2, 16384 // n)
    t0 = time.perf_counter()
    for _ in range(reps):
        ex.extract_batch(h.numpy(), out=out)
    dt = time.perf_counter() - t0
    print("frames per call %5d: %.0f frames/s (%.3f ms per call)" % (n, n * reps / dt, 1e3 * dt / reps), flush=True)
    del h, out
