"""GPU parity: the CUDA extractor (through the C-ABI) against the CPU oracle, stage by stage and end to end.
Bit-exact everywhere: pyramid bytes, FAST candidate sets, blurred planes, keypoint records (coordinates, octave,
size, response, ANGLE BITS) in quad-tree order, descriptor bytes."""
import os

import numpy as np
import pytest

from orbslam2commentedbyxcm_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    import orbslam2commentedbyxcm_b200 as pkg

    return pkg


def _sorted_rows(a):
    a = np.asarray(a)
    if len(a) == 0:
        return a
    return a[np.lexsort((a[:, 0], a[:, 1], a[:, 2]))]


CASES = [
    ("tum", 0), ("euroc", 1), ("kitti", 2),
]


def test_tables_match_oracle(gpu, oracle_mod):
    for cfg in [(1000, 1.2, 8, 20, 7), (2000, 1.2, 8, 20, 7), (500, 1.5, 5, 30, 10), (1200, 1.2, 8, 20, 7)]:
        ex = gpu.ORBextractor(*cfg)
        o = oracle_mod.Extractor(*cfg)
        img = synth.gen_frame((200, 300), 0)
        o.extract(img)
        assert np.array_equal(ex.GetScaleFactors(), o.scale_factors())
        assert np.array_equal(ex.GetInverseScaleSigmaSquares(), o.inv_level_sigma2())
        assert np.array_equal(ex.mnFeaturesPerLevel, o.quotas())
        assert ex.GetLevels() == cfg[2] and ex.GetScaleFactor() == np.float32(cfg[1])
        ex.close()


@pytest.mark.parametrize("shape_name,seed", CASES)
def test_stages_and_output_bit_exact(gpu, oracle_mod, shape_name, seed):
    img = synth.gen_frame(synth.SHAPES[shape_name], seed)
    cfg = synth.PRESETS[shape_name]
    o = oracle_mod.Extractor(*cfg)
    okps, odesc = o.extract(img)
    ex = gpu.ORBextractor(*cfg)
    kps, desc = ex(img)
    nl = cfg[2]
    # K1 pyramid
    for l in range(nl):
        w, h = o.level_size(l)
        assert ex.level_size(l) == (h, w)
        assert np.array_equal(ex.pyramid_level_padded(l), o.pyramid(l)), "pyramid level %d" % l
    # K2 FAST candidates (unordered on the GPU)
    for l in range(nl):
        assert np.array_equal(_sorted_rows(ex.debug_candidates(l)), _sorted_rows(o.candidates(l))), "candidates level %d" % l
    # K5 blur
    for l in range(nl):
        ob = o.blurred(l)
        if ob is not None:
            assert np.array_equal(ex.debug_blurred(l), ob), "blur level %d" % l
    # K3/K4/K6 final output
    assert len(kps) == len(okps)
    for f in ("x", "y", "octave", "response", "size"):
        assert np.array_equal(kps[f], okps[f]), f
    assert np.array_equal(kps["angle"].view(np.uint32), okps["angle"].view(np.uint32)), "angle bits"
    assert kps.tobytes() == okps.tobytes()
    assert np.array_equal(desc, odesc)
    ex.close()


@pytest.mark.parametrize("name", ["small_a", "small_b", "tum_s0"])
def test_against_reference_golden(gpu, golden_dir, name):
    """CUDA path == the reference's own ORBextractor.cc output (committed fixture)."""
    g = np.load(os.path.join(golden_dir, "extract_ref_%s.npz" % name))
    shape = tuple(int(v) for v in g["shape"])
    cfg = g["cfg"]
    img = synth.gen_frame(shape, int(g["seed"]))
    ex = gpu.ORBextractor(int(cfg[0]), float(cfg[1]), int(cfg[2]), int(cfg[3]), int(cfg[4]))
    kps, desc = ex(img)
    assert kps.tobytes() == g["kps"].tobytes()
    assert np.array_equal(desc, g["desc"])
    ex.close()


def test_batch_equals_per_frame_oracle(gpu, oracle_mod):
    shape = (240, 320)
    cfg = (500, 1.2, 8, 20, 7)
    imgs = np.stack([synth.gen_frame(shape, s) for s in range(6)])
    ex = gpu.ORBextractor(*cfg)
    kps, desc, counts = ex.extract_batch(imgs)
    o = oracle_mod.Extractor(*cfg)
    for i in range(len(imgs)):
        ok, od = o.extract(imgs[i])
        assert counts[i] == len(ok)
        assert kps[i, : counts[i]].tobytes() == ok.tobytes()
        assert np.array_equal(desc[i, : counts[i]], od)
    # a second batch on the same handle (buffers reused, counters reset by the kernels themselves)
    kps2, desc2, counts2 = ex.extract_batch(imgs[::-1].copy())
    assert np.array_equal(counts2, counts[::-1])
    assert kps2[0, : counts2[0]].tobytes() == kps[5, : counts[5]].tobytes()
    ex.close()


def test_graph_and_plain_launch_agree(gpu):
    img = synth.gen_frame((480, 640), 3)
    ex = gpu.ORBextractor(1000, 1.2, 8, 20, 7)
    a = ex(img)
    ex.set_use_graph(False)
    b = ex(img)
    ex.set_profiling(True)
    c = ex(img)
    t = ex.stage_times()
    assert a[0].tobytes() == b[0].tobytes() == c[0].tobytes()
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[1], c[1])
    assert all(v >= 0 for v in t.values()) and t["fast"] > 0
    ex.close()


@pytest.mark.parametrize("shape,cfg,seed", [
    ((130, 170), (2000, 1.2, 3, 20, 7), 3),     # quota never reached: tree runs until every node holds one key
    ((240, 320), (50, 1.3, 5, 30, 10), 4),      # tiny quota, other scale / thresholds
    ((97, 401), (300, 1.2, 2, 20, 7), 5),       # wide strip: several quad-tree roots
    ((333, 201), (300, 1.2, 2, 12, 5), 6),      # tall image, odd sizes
    ((480, 640), (4000, 1.2, 8, 20, 7), 7),     # mono-initialisation style doubled quota
])
def test_odd_configurations(gpu, oracle_mod, shape, cfg, seed):
    img = synth.gen_frame(shape, seed)
    ok, od = oracle_mod.Extractor(*cfg).extract(img)
    ex = gpu.ORBextractor(*cfg)
    kps, desc = ex(img)
    assert kps.tobytes() == ok.tobytes()
    assert np.array_equal(desc, od)
    ex.close()


def test_flat_and_empty_inputs(gpu, oracle_mod):
    ex = gpu.ORBextractor(1000, 1.2, 8, 20, 7)
    flat = np.full((480, 640), 127, np.uint8)
    kps, desc = ex(flat)
    assert len(kps) == 0 and desc.shape == (0, 32)
    kps, desc = ex(np.zeros((0, 0), np.uint8))  # empty image: silent return (src/ORBextractor.cc:1517)
    assert len(kps) == 0
    # binary noise: many equal responses -> exercises the candidate-order tie-break
    rng = np.random.default_rng(5)
    img = (rng.integers(0, 2, (240, 320)) * 255).astype(np.uint8)
    ok, od = oracle_mod.Extractor(1000, 1.2, 8, 20, 7).extract(img)
    kps, desc = ex(img)
    assert kps.tobytes() == ok.tobytes() and np.array_equal(desc, od)
    # uniform noise at full size: nearly every pixel survives the FAST pre-test, so k_fast_band's survivor queue
    # overflows and the band is scored in rounds (the path typical images never take)
    img = rng.integers(0, 256, (480, 640)).astype(np.uint8)
    ok, od = oracle_mod.Extractor(1000, 1.2, 8, 20, 7).extract(img)
    kps, desc = ex(img)
    assert len(ok) > 900
    assert kps.tobytes() == ok.tobytes() and np.array_equal(desc, od)
    ex.close()


def test_fast_two_phase_paths(gpu, oracle_mod):
    """k_fast_band scores at iniTh first and repeats at minTh only for the cells that kept nothing.  Cases built to hit
    every branch of that: cells whose iniTh corners are all NMS ties (2x2 blocks: four equal scores, none kept, so the
    cell is retried although it has pixels above iniTh), cells with only minTh-level corners (low contrast), cells with
    both kinds side by side in one band, low-amplitude noise (dense survivors at minTh, none at iniTh: the retry
    overflows the survivor queue), and thresholds where the retry cannot add anything (minTh >= iniTh) or where the
    byte-wise pre-test is off (>= 128)."""
    rng = np.random.default_rng(11)
    H, W = 240, 320
    imgs = []
    a = np.full((H, W), 60, np.uint8)          # 2x2 bright blocks on a grid that puts some cells' only corners in ties
    for y in range(24, H - 24, 17):
        for x in range(24, W - 24, 23):
            a[y:y + 2, x:x + 2] = 160
    imgs.append(a)
    b = np.full((H, W), 100, np.uint8)         # left half: low-contrast dots (minTh only); right half: strong dots
    for y in range(22, H - 22, 11):
        for x in range(22, W - 22, 13):
            b[y, x] = 100 + (12 if x < W // 2 else 90)
    imgs.append(b)
    c = (100 + rng.integers(-9, 10, (H, W))).astype(np.uint8)   # noise of amplitude 9: nothing reaches iniTh
    c[60:180, 100:220] = (100 + rng.integers(-40, 41, (120, 120))).astype(np.uint8)  # a textured island
    imgs.append(c)
    for cfg in [(500, 1.2, 8, 20, 7), (500, 1.2, 4, 7, 20), (500, 1.2, 4, 12, 12), (300, 1.2, 3, 130, 7), (300, 1.2, 3, 20, 128)]:
        ex = gpu.ORBextractor(*cfg)
        orc = oracle_mod.Extractor(*cfg)
        for i, img in enumerate(imgs):
            ok, od = orc.extract(img)
            kps, desc = ex(img)
            assert kps.tobytes() == ok.tobytes() and np.array_equal(desc, od), (cfg, i, len(kps), len(ok))
        ex.close()


def test_too_small_image_is_an_error_not_a_crash(gpu):
    ex = gpu.ORBextractor(1000, 1.2, 8, 20, 7)
    with pytest.raises(gpu.OrbxError):
        ex(np.zeros((60, 60), np.uint8))
    ex.close()


def test_full_size_batch_properties(gpu):
    """KITTI-shaped batch at BASELINE size through size-independent properties: determinism across runs, every
    keypoint inside the 19-px border of its level, per-level counts >= quota when candidates allow, octave order."""
    shape = synth.SHAPES["kitti"]
    cfg = synth.PRESETS["kitti"]
    base = [synth.gen_frame(shape, s) for s in range(4)]
    imgs = np.stack([base[i % 4] for i in range(32)])
    ex = gpu.ORBextractor(*cfg)
    k1, d1, c1 = ex.extract_batch(imgs)
    k2, d2, c2 = ex.extract_batch(imgs)
    assert np.array_equal(c1, c2) and k1.tobytes() == k2.tobytes() and np.array_equal(d1, d2)
    for i in range(4, 32):
        assert c1[i] == c1[i % 4] and k1[i, : c1[i]].tobytes() == k1[i % 4, : c1[i]].tobytes()
        assert np.array_equal(d1[i, : c1[i]], d1[i % 4, : c1[i]])
    sf = ex.GetScaleFactors()
    for i in range(4):
        k = k1[i, : c1[i]]
        assert c1[i] >= cfg[0] * 0.9
        assert np.all(np.diff(k["octave"]) >= 0)
        lx, ly = k["x"] / sf[k["octave"]], k["y"] / sf[k["octave"]]
        for l in range(cfg[2]):
            rows, cols = ex.level_size(l)
            m = k["octave"] == l
            assert np.all(lx[m] >= 18.99) and np.all(lx[m] <= cols - 19.99)
            assert np.all(ly[m] >= 18.99) and np.all(ly[m] <= rows - 19.99)
    ex.close()


def test_two_extractors_from_two_threads(gpu, oracle_mod):
    """the reference calls two extractor objects concurrently from two std::threads in stereo (src/Frame.cc:127-131):
    each handle owns its stream and buffers, results must not depend on the interleaving."""
    import threading

    left, right = synth.gen_stereo_pair(synth.SHAPES["euroc"], 5)
    cfg = synth.PRESETS["euroc"]
    exL, exR = gpu.ORBextractor(*cfg), gpu.ORBextractor(*cfg)
    refL = oracle_mod.Extractor(*cfg).extract(left)
    refR = oracle_mod.Extractor(*cfg).extract(right)
    out = {}

    def work(name, ex, img):
        try:
            res = None
            for _ in range(25):
                res = ex(img)
            out[name] = res
        except Exception as e:  # surfaced below with its message
            out[name] = e

    ts = [threading.Thread(target=work, args=("L", exL, left)), threading.Thread(target=work, args=("R", exR, right))]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not isinstance(out["L"], Exception), out["L"]
    assert not isinstance(out["R"], Exception), out["R"]
    assert out["L"][0].tobytes() == refL[0].tobytes() and np.array_equal(out["L"][1], refL[1])
    assert out["R"][0].tobytes() == refR[0].tobytes() and np.array_equal(out["R"][1], refR[1])
    exL.close(); exR.close()


def test_pitched_host_input_and_shape_changes(gpu, oracle_mod):
    """cv::Mat inputs may be ROIs (step > cols); one handle must also survive changing image shapes."""
    import ctypes as C

    from orbslam2commentedbyxcm_b200 import _lib

    cfg = (500, 1.2, 8, 20, 7)
    ex = gpu.ORBextractor(*cfg)
    o = oracle_mod.Extractor(*cfg)
    big = synth.gen_frame((300, 500), 11)
    roi = big[20:260, 40:360]  # 240 x 320 view with step 500
    L = _lib.load()
    cap = ex.max_keypoints(240, 320)
    kps = np.zeros(cap, gpu.KP_DTYPE)
    desc = np.zeros((cap, 32), np.uint8)
    n = C.c_int()
    _lib.check(L.orbx_extract(ex._h, big.ctypes.data + 20 * 500 + 40, 240, 320, 500, kps.ctypes.data, desc.ctypes.data, cap,
                              C.byref(n)), "orbx_extract")
    ok, od = o.extract(np.ascontiguousarray(roi))
    assert n.value == len(ok) and kps[: n.value].tobytes() == ok.tobytes() and np.array_equal(desc[: n.value], od)
    for shape, seed in [((480, 640), 1), ((240, 320), 2), ((376, 1241), 3), ((240, 320), 4)]:
        img = synth.gen_frame(shape, seed)
        k, d = ex(img)
        ok, od = o.extract(img)
        assert k.tobytes() == ok.tobytes() and np.array_equal(d, od), shape
    ex.close()


def test_degenerate_configurations(gpu, oracle_mod):
    img = synth.gen_frame((240, 320), 6)
    for cfg in [(0, 1.2, 4, 20, 7), (1, 1.2, 1, 20, 7), (300, 1.2, 8, 255, 200), (300, 2.5, 2, 20, 7), (300, 1.05, 8, 20, 20)]:
        ok, od = oracle_mod.Extractor(*cfg).extract(img)
        ex = gpu.ORBextractor(*cfg)
        k, d = ex(img)
        assert k.tobytes() == ok.tobytes() and np.array_equal(d, od), cfg
        ex.close()


def test_pinned_outputs_and_chunked_pipeline(gpu, oracle_mod):
    """large host batches are cut into pipelined chunks (graph per chunk slot); with page-locked output arrays the
    results are written straight into them.  Both variants must equal the per-frame oracle."""
    shape, cfg = (240, 320), (400, 1.2, 8, 20, 7)
    distinct = [synth.gen_frame(shape, s) for s in range(5)]
    imgs = np.stack([distinct[i % 5] for i in range(80)])
    ex = gpu.ORBextractor(*cfg)
    o = oracle_mod.Extractor(*cfg)
    refs = [o.extract(d) for d in distinct]
    for pinned in (False, True):
        out = ex.alloc_outputs(len(imgs), *shape, pinned=pinned)
        for rep in range(3):  # plain launches, graph capture, graph replay
            kps, desc, counts = ex.extract_batch(imgs, out=out)
            for i in range(len(imgs)):
                ok, od = refs[i % 5]
                assert counts[i] == len(ok), (pinned, rep, i)
                assert kps[i, : counts[i]].tobytes() == ok.tobytes() and np.array_equal(desc[i, : counts[i]], od), (pinned, rep, i)
    ex.close()


def test_getters_after_pipelined_batch_track_resident_frames(gpu, oracle_mod):
    """a host batch of more than 48 frames is pipelined through two arena slots; the pyramid / debug
    getters must translate batch frame -> arena frame for the chunks still resident and refuse the others."""
    shape, cfg = (120, 160), (300, 1.2, 4, 20, 7)
    n = 64
    frames = np.stack([synth.gen_frame(shape, 100 + s) for s in range(n)])
    ex = gpu.ORBextractor(*cfg)
    ex.extract_batch(frames)
    resident, gone = [], []
    for f in range(n):
        try:
            lvl = ex.pyramid_level_padded(1, frame=f)
        except gpu.OrbxError as e:
            assert "no longer resident" in str(e)
            gone.append(f)
            continue
        resident.append(f)
        o = oracle_mod.Extractor(*cfg)
        o.extract(frames[f])
        assert np.array_equal(lvl, o.pyramid(1)), "pyramid of batch frame %d" % f
        ob = o.blurred(1)
        if ob is not None:
            assert np.array_equal(ex.debug_blurred(1, frame=f), ob), "blurred plane of batch frame %d" % f
        assert np.array_equal(_sorted_rows(ex.debug_candidates(1, frame=f)), _sorted_rows(o.candidates(1)))
    # chunk schedule 8, 24, 32: the last two chunks are frames 8..63
    assert resident == list(range(8, 64)) and gone == list(range(8))
    with pytest.raises(gpu.OrbxError):
        ex.pyramid_level_padded(0, frame=n)
    # a small batch afterwards: everything resident again
    ex.extract_batch(frames[:5])
    for f in range(5):
        o = oracle_mod.Extractor(*cfg)
        o.extract(frames[f])
        assert np.array_equal(ex.pyramid_level_padded(0, frame=f), o.pyramid(0))
    ex.close()


def test_checkerboard_of_corners_fits_the_candidate_lists(gpu, oracle_mod):
    """the 3x3 NMS runs per cell, so the candidate capacity is the per-cell bound summed over the grid.
    Dense isolated corners (a bright dot every other pixel of every other row) must not overflow it."""
    shape = (240, 320)
    img = np.full(shape, 40, np.uint8)
    img[::2, ::2] = 220
    cfg = (1000, 1.2, 4, 20, 7)
    o = oracle_mod.Extractor(*cfg)
    okps, odesc = o.extract(img)
    ex = gpu.ORBextractor(*cfg)
    kps, desc = ex(img)
    assert len(kps) == len(okps)
    assert np.array_equal(desc, odesc)
    for l in range(cfg[2]):
        assert np.array_equal(_sorted_rows(ex.debug_candidates(l)), _sorted_rows(o.candidates(l))), "candidates level %d" % l
    ex.close()


def test_device_api_reports_truncation(gpu):
    """the device entry point accepts a capacity below orbx_max_keypoints(); a frame with more keypoints
    is cut to it and the next orbx_sync() says so."""
    import torch

    shape, cfg = synth.SHAPES["tum"], synth.PRESETS["tum"]
    img = synth.gen_frame(shape, 3)
    ex = gpu.ORBextractor(*cfg)
    full, _ = ex(img)
    assert len(full) > 200
    dev = torch.device("cuda", 0)
    d_img = torch.from_numpy(img[None].copy()).to(dev)
    cap = 200
    d_kps = torch.zeros((1, cap, 6), dtype=torch.float32, device=dev)
    d_desc = torch.zeros((1, cap, 32), dtype=torch.uint8, device=dev)
    d_cnt = torch.zeros(1, dtype=torch.int32, device=dev)
    ex.extract_batch_device(d_img.data_ptr(), 1, shape[0], shape[1], d_kps.data_ptr(), d_desc.data_ptr(), d_cnt.data_ptr(), cap)
    with pytest.raises(gpu.OrbxError):
        ex.sync()
    assert int(d_cnt.cpu()[0]) == cap
    got = d_kps[0].cpu().numpy().view(gpu.KP_DTYPE).reshape(cap)
    assert np.array_equal(got, full[:cap])
    ex.sync()   # the flag is cleared by the report
    ex.close()


@pytest.mark.parametrize("shape_name", ["tum", "kitti", "euroc", "odd_scale_1.3", "ten_levels_1.1", "scale_1.9"])
def test_large_device_batch_stage_by_stage(gpu, oracle_mod, shape_name):
    """The kernel variants that only batches select (k_resize_rows_tma at 16 rows per thread, 128-thread quad-tree CTAs, 12
    keypoints per warp in k_describe, blur beside the quad-tree): a 160-frame device-resident batch, every stage of three
    frames (first, middle, last) against the oracle."""
    import torch

    custom = {
        # scale 1.3: a CTA's source window exceeds the TMA tile rows -> the level falls back to the global-load walk
        "odd_scale_1.3": ((333, 517), (700, 1.3, 5, 20, 7)),
        # more than 8 levels: every TMA-staged kernel falls back (tensor maps are passed for 8 levels)
        "ten_levels_1.1": ((300, 400), (800, 1.1, 10, 20, 7)),
        # close to the 2x limit of the word-based resize kernels
        "scale_1.9": ((480, 640), (600, 1.9, 4, 20, 7)),
    }
    shape, cfg = custom[shape_name] if shape_name in custom else (synth.SHAPES[shape_name], synth.PRESETS[shape_name])
    distinct = [synth.gen_frame(shape, 40 + s) for s in range(5)]
    n = 160
    imgs = np.stack([distinct[i % 5] for i in range(n)])
    ex = gpu.ORBextractor(*cfg)
    cap = ex.max_keypoints(*shape)
    dev = torch.device("cuda", 0)
    d_img = torch.from_numpy(imgs).to(dev)
    d_kps = torch.zeros((n, cap, 6), dtype=torch.float32, device=dev)
    d_desc = torch.zeros((n, cap, 32), dtype=torch.uint8, device=dev)
    d_cnt = torch.zeros(n, dtype=torch.int32, device=dev)
    ex.extract_batch_device(d_img.data_ptr(), n, shape[0], shape[1], d_kps.data_ptr(), d_desc.data_ptr(), d_cnt.data_ptr(), cap)
    ex.sync()
    counts = d_cnt.cpu().numpy()
    kps = d_kps.cpu().numpy().view(gpu.KP_DTYPE).reshape(n, cap)
    desc = d_desc.cpu().numpy()
    ref = {}
    for f in (0, 77, 159):
        s = f % 5
        o = oracle_mod.Extractor(*cfg)
        okps, odesc = o.extract(distinct[s])
        ref[s] = (okps, odesc)
        for l in range(cfg[2]):
            assert np.array_equal(ex.pyramid_level_padded(l, frame=f), o.pyramid(l)), "pyramid level %d frame %d" % (l, f)
            assert np.array_equal(_sorted_rows(ex.debug_candidates(l, frame=f)), _sorted_rows(o.candidates(l))), "candidates %d/%d" % (l, f)
            ob = o.blurred(l)
            if ob is not None:
                assert np.array_equal(ex.debug_blurred(l, frame=f), ob), "blur level %d frame %d" % (l, f)
    for f in range(n):
        if f % 5 not in ref:
            continue
        okps, odesc = ref[f % 5]
        assert counts[f] == len(okps)
        assert kps[f, : counts[f]].tobytes() == okps.tobytes()
        assert np.array_equal(desc[f, : counts[f]], odesc)
    ex.close()


_VARIANT_SCRIPT = r"""
import hashlib, sys
import numpy as np, torch
import orbslam2commentedbyxcm_b200 as pkg
from orbslam2commentedbyxcm_b200 import synth
shape, cfg = synth.SHAPES["tum"], synth.PRESETS["tum"]
distinct = [synth.gen_frame(shape, 70 + s) for s in range(4)]
n = 96
imgs = np.stack([distinct[i % 4] for i in range(n)])
ex = pkg.ORBextractor(*cfg)
cap = ex.max_keypoints(*shape)
dev = torch.device("cuda", 0)
d_img = torch.from_numpy(imgs).to(dev)
d_kps = torch.zeros((n, cap, 6), dtype=torch.float32, device=dev)
d_desc = torch.zeros((n, cap, 32), dtype=torch.uint8, device=dev)
d_cnt = torch.zeros(n, dtype=torch.int32, device=dev)
ex.extract_batch_device(d_img.data_ptr(), n, shape[0], shape[1], d_kps.data_ptr(), d_desc.data_ptr(), d_cnt.data_ptr(), cap)
ex.sync()
h = hashlib.sha256()
cnt = d_cnt.cpu().numpy(); kps = d_kps.cpu().numpy(); desc = d_desc.cpu().numpy()
for f in range(n):
    h.update(cnt[f].tobytes()); h.update(kps[f, : cnt[f]].tobytes()); h.update(desc[f, : cnt[f]].tobytes())
for l in range(cfg[2]):
    h.update(ex.pyramid_level_padded(l, frame=n - 1).tobytes()); h.update(ex.debug_blurred(l, frame=n - 1).tobytes())
k1, d1 = ex(distinct[0])   # single frame through the host path (latency variants of the kernels)
h.update(k1.tobytes()); h.update(d1.tobytes())
print("DIGEST", h.hexdigest())
"""


def test_kernel_variants_agree(gpu):
    """Every staging variant of the kernels (TMA bulk tensor copies by default; ORBX_*_TMA=0 / ORBX_RESIZE_ROWS=0 select the
    cp.async / global-load / destination-row-walk variants that also serve more than 8 pyramid levels; ORBX_PDL=1 launches the
    chain with programmatic dependent launch) produces the same bytes: a 96-frame device batch + one host frame, each variant
    in its own process (the switches are read once per process)."""
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    digests = {}
    for name, env in [("default", {}), ("fast_cp_async", {"ORBX_FAST_TMA": "0"}), ("blur_global", {"ORBX_BLUR_TMA": "0"}),
                      ("resize_global", {"ORBX_RESIZE_TMA": "0"}), ("resize_words", {"ORBX_RESIZE_ROWS": "0"}),
                      ("pdl_on", {"ORBX_PDL": "1"}), ("pdl_off", {"ORBX_PDL": "0"})]:
        e = dict(os.environ)
        e.update(env)
        e["PYTHONPATH"] = root + os.pathsep + e.get("PYTHONPATH", "")
        out = subprocess.run([sys.executable, "-c", _VARIANT_SCRIPT], env=e, capture_output=True, text=True, timeout=300)
        assert out.returncode == 0, (name, out.stderr[-2000:])
        line = [l for l in out.stdout.splitlines() if l.startswith("DIGEST")]
        assert line, (name, out.stdout[-500:], out.stderr[-500:])
        digests[name] = line[0].split()[1]
    assert len(set(digests.values())) == 1, digests
