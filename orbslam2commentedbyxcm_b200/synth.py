"""Seeded synthetic inputs for the ORB front-end (no datasets are available offline).

Shapes follow BASELINE.json's configs: TUM 640x480, EuRoC 752x480 (stereo), KITTI 1241x376.
Generators are pure numpy (PCG64) so tests, bench and golden fixtures see identical bytes everywhere.
"""
import numpy as np

SHAPES = {"tum": (480, 640), "euroc": (480, 752), "kitti": (376, 1241)}
# (nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST) per BASELINE.json configs
PRESETS = {"tum": (1000, 1.2, 8, 20, 7), "euroc": (1200, 1.2, 8, 20, 7), "kitti": (2000, 1.2, 8, 20, 7)}


def _blur(img, sigma):
    r = max(1, int(3 * sigma + 0.5))
    k = np.exp(-0.5 * (np.arange(-r, r + 1) / sigma) ** 2)
    k = (k / k.sum()).astype(np.float32)
    p = np.pad(img, ((0, 0), (r, r)), mode="reflect")
    out = np.zeros_like(img)
    for i, w in enumerate(k):
        out += w * p[:, i : i + img.shape[1]]
    p = np.pad(out, ((r, r), (0, 0)), mode="reflect")
    out2 = np.zeros_like(img)
    for i, w in enumerate(k):
        out2 += w * p[i : i + img.shape[0], :]
    return out2


def gen_frame(shape, seed, n_rects=400, flat_regions=3):
    """u8 image: random 8x8 blocks + random filled rectangles + a few flat low-contrast regions, blurred
    (sigma 1) with N(0, 2^2) noise.  Textured enough for >= 5x the feature quota of FAST-20 candidates, with
    flat areas so the minThFAST retry fires."""
    h, w = shape
    rng = np.random.Generator(np.random.PCG64(seed))
    blk = rng.integers(40, 216, (h // 8 + 1, w // 8 + 1)).astype(np.float32)
    img = np.kron(blk, np.ones((8, 8), np.float32))[:h, :w].copy()
    for _ in range(n_rects):
        x0, y0 = int(rng.integers(0, w)), int(rng.integers(0, h))
        rw, rh = int(rng.integers(4, max(5, w // 5))), int(rng.integers(4, max(5, h // 4)))
        img[y0 : y0 + rh, x0 : x0 + rw] = float(rng.integers(0, 256))
    for _ in range(flat_regions):
        x0, y0 = int(rng.integers(0, max(1, w - w // 4))), int(rng.integers(0, max(1, h - h // 3)))
        rw, rh = w // 4, h // 3
        base = float(rng.integers(60, 200))
        tex = rng.integers(-9, 10, (rh // 6 + 1, rw // 6 + 1)).astype(np.float32)
        region = img[y0 : y0 + rh, x0 : x0 + rw]
        region[:] = base + np.kron(tex, np.ones((6, 6), np.float32))[: region.shape[0], : region.shape[1]]
    img = _blur(img, 1.0)
    img += rng.normal(0.0, 2.0, img.shape).astype(np.float32)
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def gen_stereo_pair(shape, seed, max_disp=60):
    """Left frame + a right frame resampled with a per-row-band constant disparity and independent noise."""
    left = gen_frame(shape, seed)
    h, w = shape
    rng = np.random.Generator(np.random.PCG64(seed + 1_000_003))
    right = np.empty_like(left)
    band = 16
    for y0 in range(0, h, band):
        d = int(rng.integers(0, max_disp + 1))
        rows = left[y0 : y0 + band]
        shifted = np.empty_like(rows)
        if d:
            shifted[:, : w - d] = rows[:, d:]
            shifted[:, w - d :] = rows[:, w - 1 : w]
        else:
            shifted[:] = rows
        right[y0 : y0 + band] = shifted
    noise = rng.normal(0.0, 1.5, right.shape)
    right = np.clip(np.rint(right.astype(np.float32) + noise), 0, 255).astype(np.uint8)
    return left, right


def gen_descriptors(n, seed=12345):
    """n x 32 uniformly random descriptor bytes."""
    rng = np.random.Generator(np.random.PCG64(seed))
    return rng.integers(0, 256, (n, 32), dtype=np.uint8)


def gen_queries(db, nq, seed=54321, flip_p=0.08):
    """nq rows of db picked at random, each bit flipped with probability flip_p (true match ~ 20 bits)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    pick = rng.integers(0, len(db), nq)
    q = db[pick].copy()
    flips = rng.random((nq, 256)) < flip_p
    q ^= np.packbits(flips, axis=1, bitorder="little")
    return q, pick.astype(np.int32)
