"""Runs every matcher kernel once on realistic inputs (for `ncu --set full`): brute force 2000 x 1M, SearchByProjection on a
KITTI-shaped frame with 4000 map points, complete stereo matching on a EuRoC-shaped pair, SearchByBoW.  GPU box only."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import orbslam2commentedbyxcm_b200 as pkg  # noqa: E402
from orbslam2commentedbyxcm_b200 import synth  # noqa: E402
from orbslam2commentedbyxcm_b200.matcher import frame_grid  # noqa: E402

m = pkg.ORBmatcher(0.8, True)
db = synth.gen_descriptors(1_000_000)
q, _ = synth.gen_queries(db, 2000)
for _ in range(2):
    m.BruteForce(q, db)

shape, cfg = synth.SHAPES["kitti"], synth.PRESETS["kitti"]
ex = pkg.ORBextractor(*cfg)
kps, desc = ex(synth.gen_frame(shape, 0))
n = len(kps)
w_inv, h_inv = float(np.float32(64.0) / np.float32(shape[1])), float(np.float32(48.0) / np.float32(shape[0]))
cs, ci = frame_grid(kps["x"], kps["y"], 0.0, 0.0, w_inv, h_inv)
rng = np.random.Generator(np.random.PCG64(99))
nmp = 4000
pick = rng.integers(0, n, nmp)
flips = rng.random((nmp, 256)) < 0.08
frame = dict(x=kps["x"].copy(), y=kps["y"].copy(), octave=kps["octave"].astype(np.int32), u_right=None, desc=desc,
             occupied=np.zeros(n, np.uint8), cell_start=cs, cell_items=ci, min_x=0.0, min_y=0.0, w_inv=w_inv, h_inv=h_inv,
             scale_factors=ex.GetScaleFactors())
mps = dict(x=(kps["x"][pick] + rng.normal(0, 2, nmp)).astype(np.float32), y=(kps["y"][pick] + rng.normal(0, 2, nmp)).astype(np.float32),
           xr=np.full(nmp, -1, np.float32), level=kps["octave"][pick].astype(np.int32), view_cos=np.full(nmp, 0.9, np.float32),
           valid=np.ones(nmp, np.uint8), obs=np.ones(nmp, np.uint8),
           desc=np.ascontiguousarray(desc[pick] ^ np.packbits(flips, axis=1, bitorder="little")))
for _ in range(2):
    print("projection matches", m.SearchByProjection(frame, mps, 3.0)[0])

shape, cfg = synth.SHAPES["euroc"], synth.PRESETS["euroc"]
left, right = synth.gen_stereo_pair(shape, 0)
eL, eR = pkg.ORBextractor(*cfg), pkg.ORBextractor(*cfg)
kl, dl = eL(left)
kr, dr = eR(right)
L = dict(x=kl["x"], y=kl["y"], octave=kl["octave"], desc=dl)
R = dict(x=kr["x"], y=kr["y"], octave=kr["octave"], desc=dr)
for _ in range(2):
    ur, dp = m.ComputeStereoMatches(eL, eR, L, R, 47.9, 47.9 / 435.2)
print("stereo matches", int((ur >= 0).sum()))

# BoW: frame vs a perturbed copy of itself, 400 vocabulary nodes
nf = len(kl)
node_of = rng.integers(0, 400, nf)


def fv(node):
    ids = np.unique(node)
    start, feat = [0], []
    for nid in ids:
        f = np.nonzero(node == nid)[0]
        feat.extend(f.tolist())
        start.append(len(feat))
    return dict(node_id=ids.astype(np.int32), node_start=np.array(start, np.int32), node_feat=np.array(feat, np.int32))


perm = rng.permutation(nf)
kf = dict(desc=np.ascontiguousarray(dl[perm]), angle=kl["angle"][perm].astype(np.float32), has_mp=np.ones(nf, np.uint8), **fv(node_of[perm]))
fr = dict(desc=dl, angle=kl["angle"].astype(np.float32), **fv(node_of))
m2 = pkg.ORBmatcher(0.7, True)
for _ in range(2):
    print("bow matches", m2.SearchByBoW(kf, fr)[0])
