"""Shared builders of synthetic matcher inputs (flat views of Frame / MapPoint / FeatureVector)."""
import numpy as np

from orbslam2commentedbyxcm_b200 import synth


def popcount_dist(a, b):
    return int(np.unpackbits(np.bitwise_xor(a, b)).sum())


def make_frame_view(oracle_mod, kps, desc, rows, cols, scale_factors, occupied_p=0.1, seed=0, stereo=False):
    """Frame-side view from extractor output: undistorted keys = keys (no distortion), 64x48 grid as in Frame.cc."""
    rng = np.random.Generator(np.random.PCG64(seed))
    n = len(kps)
    x = kps["x"].astype(np.float32).copy()
    y = kps["y"].astype(np.float32).copy()
    min_x, min_y, max_x, max_y = 0.0, 0.0, float(cols), float(rows)
    w_inv = np.float32(64.0) / np.float32(max_x - min_x)
    h_inv = np.float32(48.0) / np.float32(max_y - min_y)
    cs, ci = oracle_mod.assign_grid(x, y, min_x, min_y, float(w_inv), float(h_inv))
    occ = (rng.random(n) < occupied_p).astype(np.uint8)
    ur = None
    if stereo:
        ur = np.where(rng.random(n) < 0.6, x - rng.random(n).astype(np.float32) * 40, -1).astype(np.float32)
    return dict(x=x, y=y, octave=kps["octave"].astype(np.int32).copy(), u_right=ur, desc=np.ascontiguousarray(desc),
                occupied=occ, cell_start=cs, cell_items=ci, min_x=min_x, min_y=min_y, w_inv=float(w_inv), h_inv=float(h_inv),
                scale_factors=np.asarray(scale_factors, np.float32))


def make_mappoints(frame, nmp, seed=0, dup_p=0.3, nlevels=8, stereo=False):
    """Map points projected near frame keypoints (descriptor = perturbed keypoint descriptor), a share of them
    deliberately projecting onto the SAME keypoint so the order-dependent claims matter."""
    rng = np.random.Generator(np.random.PCG64(seed + 99))
    n = len(frame["x"])
    pick = rng.integers(0, n, nmp)
    ndup = int(nmp * dup_p)
    pick[nmp - ndup:] = pick[rng.integers(0, max(1, nmp - ndup), ndup)]
    px = (frame["x"][pick] + rng.normal(0, 2.0, nmp)).astype(np.float32)
    py = (frame["y"][pick] + rng.normal(0, 2.0, nmp)).astype(np.float32)
    level = np.clip(frame["octave"][pick] + rng.integers(-1, 2, nmp), 0, nlevels - 1).astype(np.int32)
    flips = rng.random((nmp, 256)) < rng.choice([0.03, 0.08, 0.2], nmp)[:, None]
    desc = frame["desc"][pick] ^ np.packbits(flips, axis=1, bitorder="little")
    view_cos = rng.choice([0.9, 0.9985, 0.9999], nmp).astype(np.float32)
    valid = (rng.random(nmp) < 0.9).astype(np.uint8)
    obs = (rng.random(nmp) < 0.85).astype(np.uint8)
    if stereo and frame.get("u_right") is not None:
        xr = (px - rng.random(nmp).astype(np.float32) * 40).astype(np.float32)
    else:
        xr = np.full(nmp, -1, np.float32)
    return dict(x=px, y=py, xr=xr, level=level, view_cos=view_cos, valid=valid, obs=obs, desc=np.ascontiguousarray(desc))


def make_featvec(n_features, n_nodes_total, seed):
    """CSR FeatureVector: each feature goes to exactly one random node id; node ids ascending."""
    rng = np.random.Generator(np.random.PCG64(seed))
    node_of = rng.integers(0, n_nodes_total, n_features)
    ids = np.unique(node_of)
    start = [0]
    feat = []
    for nid in ids:
        f = np.nonzero(node_of == nid)[0]
        feat.extend(f.tolist())
        start.append(len(feat))
    return dict(node_id=ids.astype(np.int32), node_start=np.array(start, np.int32), node_feat=np.array(feat, np.int32)), node_of


def extract_with_oracle(oracle_mod, shape_name, seed, kind="strict"):
    img = synth.gen_frame(synth.SHAPES[shape_name], seed)
    cfg = synth.PRESETS[shape_name]
    ex = oracle_mod.Extractor(*cfg, kind=kind)
    kps, desc = ex.extract(img)
    return img, cfg, ex, kps, desc
