"""GPU parity of the frame preparation between extraction and matching, through the C-ABI:
Frame::UndistortKeyPoints (src/Frame.cc:586-628) and Frame::AssignFeaturesToGrid / PosInGrid (:351-370, :558-567).
Bit-exact against the CPU oracle and against the committed cv2 4.13 fixtures (tests/golden/undistort_cv2.npz)."""
import os

import numpy as np
import pytest

from orbslam2commentedbyxcm_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    import orbslam2commentedbyxcm_b200 as pkg
    return pkg


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def test_undistort_equals_cv2_fixtures_and_oracle(gpu, oracle_mod, golden_dir):
    g = np.load(os.path.join(golden_dir, "undistort_cv2.npz"))
    m = gpu.ORBmatcher()
    ncases = sum(1 for k in g.files if k.startswith("cam"))
    for c in range(ncases):
        fx, fy, cx, cy = (float(v) for v in g["cam%d" % c])
        d = g["dist%d" % c]
        pts, ref = g["pts%d" % c], g["res%d" % c]
        xu, yu = m.UndistortKeyPoints(pts[:, 0], pts[:, 1], fx, fy, cx, cy, d)
        assert np.array_equal(_bits(xu), _bits(ref[:, 0])), "case %d x vs cv2" % c
        assert np.array_equal(_bits(yu), _bits(ref[:, 1])), "case %d y vs cv2" % c
        ox, oy = oracle_mod.undistort_points(pts[:, 0], pts[:, 1], fx, fy, cx, cy, d)
        assert np.array_equal(_bits(xu), _bits(ox)) and np.array_equal(_bits(yu), _bits(oy))
    m.close()


def test_undistort_random_cameras_and_edge_cases(gpu, oracle_mod):
    rng = np.random.default_rng(33)
    m = gpu.ORBmatcher()
    for trial in range(8):
        fx, fy = (float(v) for v in 300 + 500 * rng.random(2))
        cx, cy = float(320 + 60 * rng.standard_normal()), float(240 + 40 * rng.standard_normal())
        nd = (4, 5, 8, 12)[trial % 4]
        d = (rng.standard_normal(nd) * ([0.3, 1.0, 0.01, 0.01, 1.0, 0.1, 0.1, 0.1, 0.01, 0.01, 0.01, 0.01][:nd])).astype(np.float32)
        n = (1, 7, 2000, 5000)[trial % 4]
        x = (rng.random(n) * 1300 - 30).astype(np.float32)      # also outside the image: strong distortion, icdist < 0 path
        y = (rng.random(n) * 800 - 30).astype(np.float32)
        xu, yu = m.UndistortKeyPoints(x, y, fx, fy, cx, cy, d)
        ox, oy = oracle_mod.undistort_points(x, y, fx, fy, cx, cy, d)
        assert np.array_equal(_bits(xu), _bits(ox)) and np.array_equal(_bits(yu), _bits(oy)), trial
    # k1 == 0: the reference copies the keypoints (src/Frame.cc:589-592), whatever the other coefficients are
    x = rng.random(100).astype(np.float32) * 640
    y = rng.random(100).astype(np.float32) * 480
    xu, yu = m.UndistortKeyPoints(x, y, 500.0, 500.0, 320.0, 240.0, np.array([0, 0.5, 0.1, 0.1, 0.2], np.float32))
    assert np.array_equal(xu, x) and np.array_equal(yu, y)
    xu, yu = m.UndistortKeyPoints(x[:0], y[:0], 500.0, 500.0, 320.0, 240.0, np.array([0.1, 0, 0, 0], np.float32))
    assert len(xu) == 0 and len(yu) == 0
    m.close()


@pytest.mark.parametrize("shape_name", ["tum", "kitti"])
def test_assign_grid_equals_oracle_on_extracted_frames(gpu, oracle_mod, shape_name):
    rows, cols = synth.SHAPES[shape_name]
    cfg = synth.PRESETS[shape_name]
    ex = gpu.ORBextractor(*cfg)
    kps, _ = ex(synth.gen_frame((rows, cols), 2))
    ex.close()
    x, y = kps["x"].copy(), kps["y"].copy()
    m = gpu.ORBmatcher()
    # image bounds as in Frame::ComputeImageBounds without distortion, and a shifted / shrunk box so that keypoints fall
    # outside the grid (PosInGrid returns false for them)
    for (mnx, mny, mxx, mxy) in [(0.0, 0.0, float(cols), float(rows)), (25.5, 31.25, cols - 40.0, rows - 17.0)]:
        w_inv = float(np.float32(64.0) / np.float32(mxx - mnx))
        h_inv = float(np.float32(48.0) / np.float32(mxy - mny))
        cs, ci = m.AssignFeaturesToGrid(x, y, mnx, mny, w_inv, h_inv)
        os_, oi = oracle_mod.assign_grid(x, y, mnx, mny, w_inv, h_inv)
        assert np.array_equal(cs, os_) and np.array_equal(ci, oi)
    m.close()


def test_assign_grid_crowded_cells_ties_and_empty(gpu, oracle_mod):
    rng = np.random.default_rng(8)
    m = gpu.ORBmatcher()
    # thousands of keypoints in a handful of cells (long per-cell lists must come out in index order), exact .5 grid
    # positions (round half away from zero), negative and far-out coordinates
    n = 6000
    x = (rng.integers(0, 6, n) * 10.0 + rng.random(n)).astype(np.float32)
    y = (rng.integers(0, 4, n) * 10.0 + rng.random(n)).astype(np.float32)
    x[:50] = np.arange(50, dtype=np.float32) * 5.0 + 5.0     # (x - 0) * 0.1 = k + 0.5 exactly
    y[:50] = 2.5
    x[50:60] = -3.0
    y[60:70] = 1e6
    cs, ci = m.AssignFeaturesToGrid(x, y, 0.0, 0.0, 0.1, 0.1)
    os_, oi = oracle_mod.assign_grid(x, y, 0.0, 0.0, 0.1, 0.1)
    assert np.array_equal(cs, os_) and np.array_equal(ci, oi)
    cs, ci = m.AssignFeaturesToGrid(x[:0], y[:0], 0.0, 0.0, 0.1, 0.1)
    assert cs.shape == (64 * 48 + 1,) and not cs.any() and len(ci) == 0
    m.close()


def test_device_pointer_mode_feeds_search_by_projection(gpu, oracle_mod):
    """mem_kind = 1: undistort + grid on device buffers, asynchronous on the context's stream; the CSR grid read back is
    the one SearchByProjection consumes (same arrays as the host mode)."""
    import torch

    rng = np.random.default_rng(4)
    dev = torch.device("cuda:0")
    n = 3000
    x = (rng.random(n) * 640).astype(np.float32)
    y = (rng.random(n) * 480).astype(np.float32)
    d = np.array([0.2624, -0.9531, -0.0054, 0.0026, 1.1633], np.float32)
    fx, fy, cx, cy = 517.3, 516.5, 318.6, 255.3
    m = gpu.ORBmatcher()
    dx, dy = torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev)
    dxu, dyu = torch.empty_like(dx), torch.empty_like(dy)
    dstart = torch.empty(64 * 48 + 1, dtype=torch.int32, device=dev)
    ditems = torch.empty(n, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    m.undistort_device(dx.data_ptr(), dy.data_ptr(), n, fx, fy, cx, cy, d, dxu.data_ptr(), dyu.data_ptr())
    w_inv, h_inv = float(np.float32(64.0) / np.float32(640.0)), float(np.float32(48.0) / np.float32(480.0))
    m.assign_grid_device(dxu.data_ptr(), dyu.data_ptr(), n, 0.0, 0.0, w_inv, h_inv, dstart.data_ptr(), ditems.data_ptr())
    m.sync()
    ox, oy = oracle_mod.undistort_points(x, y, fx, fy, cx, cy, d)
    assert np.array_equal(_bits(dxu.cpu().numpy()), _bits(ox)) and np.array_equal(_bits(dyu.cpu().numpy()), _bits(oy))
    os_, oi = oracle_mod.assign_grid(ox, oy, 0.0, 0.0, w_inv, h_inv)
    cs = dstart.cpu().numpy()
    assert np.array_equal(cs, os_) and np.array_equal(ditems.cpu().numpy()[: cs[-1]], oi)
    m.close()
