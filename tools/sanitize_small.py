"""Small end-to-end invocation of every kernel, for `compute-sanitizer --tool memcheck python tools/sanitize_small.py`."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import orbslam2commentedbyxcm_b200 as pkg  # noqa: E402
from orbslam2commentedbyxcm_b200 import synth, vocabulary  # noqa: E402
from orbslam2commentedbyxcm_b200.matcher import frame_grid  # noqa: E402

cfg = (500, 1.2, 8, 20, 7)
exL, exR = pkg.ORBextractor(*cfg), pkg.ORBextractor(*cfg)
left, right = synth.gen_stereo_pair((240, 320), 1)
for _ in range(3):  # plain launch, graph capture, graph replay
    kl, dl = exL(left)
    kr, dr = exR(right)
kb, db_, cb = exL.extract_batch(np.stack([left, right, left]))
m = pkg.ORBmatcher(0.8, True)
L = dict(x=kl["x"], y=kl["y"], octave=kl["octave"], desc=dl)
R = dict(x=kr["x"], y=kr["y"], octave=kr["octave"], desc=dr)
exL(left); exR(right)
print("stereo", int((m.ComputeStereoMatches(exL, exR, L, R, 40.0, 0.1)[0] >= 0).sum()))
print("bf", m.BruteForce(dl[:100], synth.gen_descriptors(5000))[0][:3])
n = len(kl)
w_inv, h_inv = 64.0 / 320, 48.0 / 240
cs, ci = frame_grid(kl["x"], kl["y"], 0.0, 0.0, w_inv, h_inv)
frame = dict(x=kl["x"].copy(), y=kl["y"].copy(), octave=kl["octave"].astype(np.int32), u_right=None, desc=dl, occupied=np.zeros(n, np.uint8),
             cell_start=cs, cell_items=ci, min_x=0.0, min_y=0.0, w_inv=w_inv, h_inv=h_inv, scale_factors=exL.GetScaleFactors(),
             angle=kl["angle"].copy())
mps = dict(x=kl["x"] + 1, y=kl["y"] - 1, xr=np.full(n, -1, np.float32), level=kl["octave"].astype(np.int32),
           view_cos=np.full(n, 0.9, np.float32), valid=np.ones(n, np.uint8), obs=np.ones(n, np.uint8), desc=dl)
print("proj", m.SearchByProjection(frame, mps, 3.0)[0])
q = dict(x=kl["x"] + 1, y=kl["y"], radius=np.full(n, 10, np.float32), min_level=kl["octave"].astype(np.int32) - 1,
         max_level=kl["octave"].astype(np.int32) + 1, xr=None, angle=kl["angle"].copy(), valid=np.ones(n, np.uint8),
         blocks=np.ones(n, np.uint8), desc=dl)
print("window", m.SearchWindow(frame, q, 100, 0)[0])
voc = vocabulary.ORBVocabulary(vocabulary.synthetic_vocabulary(10, 3), m)
(_, _), fv = voc.transform(dl, 2)
kf = dict(desc=dl, angle=kl["angle"].copy(), has_mp=np.ones(n, np.uint8), **fv)
fr = dict(desc=dl, angle=kl["angle"].copy(), **fv)
print("bow", m.SearchByBoW(kf, fr)[0])
print("distinct", m.ComputeDistinctiveDescriptors(np.array([0, 5, 40], np.int32), dl[:40])[:2])
print("sanitize_small done")
