"""End-to-end frames/s of orbx_extract_batch (pinned host in / out, 512 frames per call) from ONE host thread against TWO host
threads with one extractor handle each (the reference's stereo front-end drives two ORBextractor objects from two threads):
the fill / drain of one call's pipeline overlaps the other call's transfers."""
import os, sys, time, threading
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import orbslam2commentedbyxcm_b200 as pkg
from orbslam2commentedbyxcm_b200 import synth

shape, cfg = synth.SHAPES["tum"], synth.PRESETS["tum"]
rows, cols = shape
for n in (256, 512):
    frames = bench.make_frames(shape, 32, n)
    for nth in (1, 2, 3):
        exs = [pkg.ORBextractor(*cfg) for _ in range(nth)]
        hs = [torch.from_numpy(frames.copy()).pin_memory() for _ in range(nth)]
        outs = [ex.alloc_outputs(n, rows, cols, pinned=True) for ex in exs]
        calls = max(8, 8192 // n)
        def work(i, reps):
            for _ in range(reps):
                exs[i].extract_batch(hs[i].numpy(), out=outs[i])
        def run(reps):
            ts = [threading.Thread(target=work, args=(i, reps)) for i in range(nth)]
            [t.start() for t in ts]; [t.join() for t in ts]
        t0 = time.perf_counter()
        while time.perf_counter() - t0 < 0.5:
            run(2)
        t0 = time.perf_counter(); run(calls); dt = time.perf_counter() - t0
        print("%d frames per call, %d thread(s): %.0f frames/s" % (n, nth, nth * calls * n / dt), flush=True)
        for ex in exs:
            ex.close()
        del hs, outs
