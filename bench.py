#!/usr/bin/env python
"""bench.py -- ORB front-end hot path on B200: frames/s of ORB extraction (640x480, 1000 features) + Hamming pairs/s.

  python bench.py --gpus N --steps K --warmup W            our CUDA path (one rank per GPU; torchrun for N > 1)
  python bench.py --impl reference --gpus N --steps K ...  the reference's own CPU ORBextractor on the host cores

A "step" = INNER passes of the extractor over a BATCH-frame buffer of synthetic frames (INNER x BATCH frames per GPU, weak
scaling), so that the K timed steps cover >= 0.5 s.
  value : frames/s with the frames already resident in HBM (device in, device out, CUDA events on the handle's stream)
  e2e   : frames/s through the C-ABI host entry point orbx_extract_batch (pinned host input, H2D + kernels + D2H +
          the copy into caller arrays inside the timed region)
  roofline      : dominant kernel stage, algorithmic bytes / its measured duration vs MEASURED_PEAKS
  cpu_baseline  : the CPU oracle / reference build on the host cores, on a bounded sample of the same workload
Prints ONE JSON line (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from orbslam2commentedbyxcm_b200 import sharding, synth  # noqa: E402

WORKLOAD = "tum"  # BASELINE.json metric: "ORB extract frames/s @640x480 1k feats"
WORKLOAD_NAME = "C1 TUM 640x480, 1000 feats, scale 1.2, 8 levels, FAST 20/7"
# identical in both arms' JSON lines (the driver compares them); what one step holds is in "step_detail"
CONFIG = {"workload": WORKLOAD_NAME, "parallelism": "frames sharded by rank, no data-path collective",
          "l2": "inputs larger than L2: every pass streams a frame buffer + pyramid of more than 126 MB"}
BATCH = 512       # frames per GPU per pass (one device buffer, larger than L2 with its pyramid; measured 256 -> 283 k,
                  # 512 -> 290 k, 1024 -> 294 k frames/s: the small upper pyramid levels fill the GPU better)
INNER = 16        # passes per step: a step is INNER x BATCH frames per GPU, so 20 steps time >= 0.5 s
E2E_INNER = 8     # host calls (E2E_BATCH frames each) per end-to-end step
E2E_BATCH = 512   # frames per end-to-end host call (measured sweep, tools/e2e_batch_sweep.py: 256 -> 148 k, 512 -> 160 k,
                  # >= 1024 -> 122-128 k frames/s: the pipeline's fill / drain amortises, larger pinned buffers copy slower)
C4_JOBS = 16      # repetitions of the 4096-frame config-4 job inside its timed region
N_DISTINCT = 32   # distinct synthetic frames (seeds 0..31), tiled to BATCH


def level_geometry(rows, cols, nlevels, scale_factor):
    s = np.float32(1.0)
    out = []
    for l in range(nlevels):
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float32(cols) * inv)), int(np.rint(np.float32(rows) * inv))))
        s = np.float32(s * np.float32(scale_factor))
    return out


def algorithmic_bytes(rows, cols, cfg):
    """algorithmic bytes per frame and per stage."""
    nf, sf, nl = cfg[0], cfg[1], cfg[2]
    lv = level_geometry(rows, cols, nl, sf)
    interior = [w * h for w, h in lv]
    padded = [(w + 38) * (h + 38) for w, h in lv]
    pyr = rows * cols + padded[0] + sum(interior[l - 1] + padded[l] for l in range(1, nl))
    return {
        "pyramid": pyr,
        "fast": sum(interior) + 12 * 10 * nf,
        "blur": 2 * sum(interior),
        "octree": 8 * 10 * nf,       # the quad-tree reads 8 B per FAST candidate (~10 x nfeatures)
        "desc": nf * (749 + 16) + nf * (512 + 32),   # IC_Angle disc reads + rBRIEF samples/descriptor (fused in k_describe)
    }


class ClockSampler:
    """SM clock and throttle reasons sampled every ~10 ms while the timed regions run.  NVML in-process (initialised before
    the timed region: an `nvidia-smi -lms` child spends its first second initialising the driver, which showed up as
    noise in the host-side end-to-end timing and produced only a couple of samples); nvidia-smi is the fallback."""

    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []
        self.nvml = None
        self.samples = []
        self.stop_flag = False
        try:
            import pynvml
            pynvml.nvmlInit()
            # NVML enumerates physical devices: map the CUDA ordinal through CUDA_VISIBLE_DEVICES when it lists indices
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            ids = [v.strip() for v in vis.split(",") if v.strip()]
            phys = int(ids[index]) if index < len(ids) and ids[index].isdigit() else index
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def start(self):
        if self.nvml is not None:
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _poll(self):
        n = self.nvml
        while not self.stop_flag:
            try:
                mhz = float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM))
                why = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
                self.samples.append((mhz, why))
            except Exception:
                pass
            time.sleep(0.01)

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.nvml is not None:
            n = self.nvml
            self.stop_flag = True
            self.t.join(timeout=1)
            bits = {"hw_slowdown": n.nvmlClocksThrottleReasonHwSlowdown,
                    "hw_thermal_slowdown": n.nvmlClocksThrottleReasonHwThermalSlowdown,
                    "sw_thermal_slowdown": n.nvmlClocksThrottleReasonSwThermalSlowdown,
                    "sw_power_cap": n.nvmlClocksThrottleReasonSwPowerCap}
            sm = [m for m, _ in self.samples]
            reasons = sorted(k for k, b in bits.items() if any(w & b for _, w in self.samples))
            return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": self.max_mhz, "reasons": reasons,
                    "samples": len(sm), "source": "nvml"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 6:
                continue
            try:
                sm.append(float(p[0])); mx.append(float(p[1]))
            except ValueError:
                continue
            for n, v in zip(names, p[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi"}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def measured_peaks_all():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p))
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}


def measured_traffic(stage, frames):
    """(DRAM bytes per launch of a stage, source): NOT measured by this run -- dram__bytes_read + dram__bytes_write of a committed
    `ncu --set full` capture (per frame, 64-frame launch) scaled to this launch's frame count; (None, why) without one."""
    for name in ("r2_traffic.json", "r1_traffic.json"):
        p = os.path.join(ROOT, "profiles", name)
        try:
            d = json.load(open(p))
            return int(d["bytes_per_frame"][stage] * frames), "ncu capture profiles/%s, extrapolated per frame (not measured live)" % name
        except Exception:
            continue
    return None, "no ncu capture committed"


def measured_warp_inst(stage, frames):
    """(executed warp instructions per launch of a stage, source) from the committed ncu capture, scaled per frame; (None, why)."""
    p = os.path.join(ROOT, "profiles", "r2_traffic.json")
    try:
        d = json.load(open(p))["warp_inst_per_frame"]
        return float(d[stage]) * frames, d["_source"]
    except Exception:
        return None, "no ncu capture committed for this stage"


def matcher_reference():
    """(oracle.matchref, kind) with the -O3 build of the reference's own matcher side selected when it travelled with the
    snapshot (oracle/_ref/libmatchref_fast.so), else the -O2 test build, else None.  CPU baselines only."""
    try:
        from oracle import matchref as R
    except Exception:
        return None, None
    for v in ("ref_fast", "ref"):
        if R.use(v):
            return R, v
    return None, None


def hamming_cpu_baseline(pkg, local, q, db_sample):
    """the reference's SWAR DescriptorDistance scan and a __builtin_popcountll variant, 1 thread and all cores,
    on a bounded sample (all queries x the first rows of rank 0's shard); results checked equal to the GPU search of the sample."""
    R, kind = matcher_reference()
    if R is None:
        return {"unavailable": "oracle/_ref/libmatchref*.so not present"}
    threads = host_threads()
    m = pkg.ORBmatcher(0.6, True, device=local)
    gb, gi, gs = m.BruteForce(q, db_sample)
    m.close()
    nq1 = max(1, len(q) // 8)
    out = {"kind": "reference", "cores": threads, "unit": "descriptor pairs/s",
           "build": "src/ORBmatcher.cc unmodified, %s" % ("-O3 -march=x86-64-v3" if kind == "ref_fast" else "-O2"),
           "sample": "%d queries x %d rows on all cores, %d queries x %d rows on one thread" % (len(q), len(db_sample), nq1, len(db_sample))}
    for name, variant in (("reference_swar", 0), ("popcountll", 1)):
        b, i, s2, ms_all = R.bruteforce_scan(q, db_sample, threads=threads, variant=variant)
        if not (np.array_equal(b, gb) and np.array_equal(i, gi) and np.array_equal(s2, gs)):
            raise RuntimeError("bench: host %s scan and the GPU search disagree on the sample" % name)
        _b, _i, _s, ms_one = R.bruteforce_scan(q[:nq1], db_sample, threads=1, variant=variant)
        out[name] = {"pairs_per_s_all_cores": len(q) * len(db_sample) / (ms_all * 1e-3),
                     "pairs_per_s_one_thread": nq1 * len(db_sample) / (ms_one * 1e-3)}
    out["value"] = out["reference_swar"]["pairs_per_s_all_cores"]
    out["results_equal_gpu"] = True
    return out


def make_frames(shape, n_distinct, batch):
    base = np.stack([synth.gen_frame(shape, s) for s in range(n_distinct)])
    reps = (batch + n_distinct - 1) // n_distinct
    return np.concatenate([base] * reps)[:batch].copy()


def output_digest(kps, desc, counts, n):
    """sha256 over (count, keypoint records, descriptor rows) of the first n frames: the same bytes from the CUDA path and
    from the reference arm mean identical keypoints (coordinates, octave, angle bits, response, order) and descriptors."""
    import hashlib

    h = hashlib.sha256()
    for i in range(n):
        c = int(counts[i])
        h.update(np.int32(c).tobytes())
        h.update(np.ascontiguousarray(kps[i, :c]).tobytes())
        h.update(np.ascontiguousarray(desc[i, :c]).tobytes())
    return h.hexdigest()


def reference_digest(frames, cfg, cap):
    """digest of the reference's own ORBextractor.cc (oracle/_ref) on `frames`, or (None, why)."""
    import oracle

    ref = oracle.ref_extract_batch_full(frames, *cfg, cap=cap, threads=host_threads())
    if ref is None:
        return None, "oracle/_ref/liborbref_fast.so not present"
    rk, rd, rc = ref
    return output_digest(rk, rd, rc, len(frames)), "reference src/ORBextractor.cc (oracle/_ref/liborbref_fast.so)"


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# ------------------------------------------------------------------------------------------------------------------
def run_reference(args):
    """--impl reference: the reference's own ORBextractor.cc (oracle/_ref, compiled unmodified against the mini-cv shim) on
    the host cores, frame-parallel with one extractor object per thread; falls back to the restated oracle port."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle

    shape = synth.SHAPES[WORKLOAD]
    cfg = synth.PRESETS[WORKLOAD]
    threads = host_threads()
    kind = "ref_fast" if oracle.lib("ref_fast") is not None else "fast"
    sample = max(threads * 4, 16)  # frames per step: a bounded sample of the BATCH-frame workload
    frames = make_frames(shape, min(N_DISTINCT, sample), sample)
    for _ in range(max(args.warmup, 1)):
        oracle.extract_batch(frames[: threads * 2], *cfg, threads=threads, kind=kind)
    t0 = time.perf_counter()
    total_kp = 0
    for _ in range(args.steps):
        tot, _ = oracle.extract_batch(frames, *cfg, threads=threads, kind=kind)
        total_kp += tot
    dt = time.perf_counter() - t0
    fps = sample * args.steps / dt
    distinct = make_frames(shape, N_DISTINCT, N_DISTINCT)
    digest, digest_src = (reference_digest(distinct, cfg, 2 * cfg[0]) if kind == "ref_fast" else (None, "oracle port only"))
    line = {
        "impl": "reference", "metric": "orb_extract_frames_per_s_640x480_1000feat", "value": fps, "unit": "frames/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": dict(CONFIG),
        "step_detail": {"step": "%d-frame sample of the workload per step, frame-parallel on %d host threads" % (sample, threads)},
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": threads,
                         "kind": "reference" if kind == "ref_fast" else "port",
                         "sample": "%d frames x %d steps, %d host threads, %s" % (sample, args.steps, threads,
                                   "reference src/ORBextractor.cc + mini-cv shim (-O3 x86-64-v3)" if kind == "ref_fast" else "oracle port (-O3 x86-64-v3)")},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "keypoints_per_frame": total_kp / (sample * args.steps),
        "output_sha256": {"frames": N_DISTINCT, "digest": digest, "of": digest_src},
    }
    print(json.dumps(line))


def other_configs(pkg, torch, dev, local):
    """BASELINE.json configs 2 and 3 on one GPU (reported beside the headline; parity for them is in tests/):
    C2 EuRoC stereo 752x480 x2, 1200 feats: two extractor handles on two host threads' worth of streams + the complete
       Frame::ComputeStereoMatches (row search + SAD refinement on the device pyramids);
    C3 KITTI 1241x376, 2000 feats: batched extraction + SearchByProjection against a synthetic local map of 4000 points."""
    from orbslam2commentedbyxcm_b200.matcher import frame_grid

    out = {}
    # ---- C2 ----
    shape, cfg = synth.SHAPES["euroc"], synth.PRESETS["euroc"]
    left, right = synth.gen_stereo_pair(shape, 0)
    exL, exR = pkg.ORBextractor(*cfg, device=local), pkg.ORBextractor(*cfg, device=local)
    mt = pkg.ORBmatcher(0.6, True, device=local)
    oL, oR = exL.alloc_outputs(1, *shape), exR.alloc_outputs(1, *shape)
    fx, bf = 435.2, 47.9
    mb = float(np.float32(bf) / np.float32(fx))

    def stereo_frame():
        kl, dl, cl = exL.extract_batch(left[None], out=oL)
        kr, dr, cr = exR.extract_batch(right[None], out=oR)
        L = dict(x=kl[0, :cl[0]]["x"], y=kl[0, :cl[0]]["y"], octave=kl[0, :cl[0]]["octave"], desc=dl[0, :cl[0]])
        R = dict(x=kr[0, :cr[0]]["x"], y=kr[0, :cr[0]]["y"], octave=kr[0, :cr[0]]["octave"], desc=dr[0, :cr[0]])
        t = time.perf_counter()
        ur, dp = mt.ComputeStereoMatches(exL, exR, L, R, bf, mb)
        return time.perf_counter() - t, int(cl[0]), int(cr[0]), int((ur >= 0).sum())

    for _ in range(10):
        stereo_frame()
    tt, tm = [], []
    for _ in range(100):
        t0 = time.perf_counter()
        dm, nl, nr, nmatch = stereo_frame()
        tt.append(time.perf_counter() - t0)
        tm.append(dm)
    # the reference's own threading (src/Frame.cc:127-131): left and right extraction on two host threads, one extractor each
    from concurrent.futures import ThreadPoolExecutor
    pool = ThreadPoolExecutor(1)

    def stereo_frame_threads():
        fut = pool.submit(exR.extract_batch, right[None], oR)
        kl, dl, cl = exL.extract_batch(left[None], out=oL)
        kr, dr, cr = fut.result()
        L = dict(x=kl[0, :cl[0]]["x"], y=kl[0, :cl[0]]["y"], octave=kl[0, :cl[0]]["octave"], desc=dl[0, :cl[0]])
        R = dict(x=kr[0, :cr[0]]["x"], y=kr[0, :cr[0]]["y"], octave=kr[0, :cr[0]]["octave"], desc=dr[0, :cr[0]])
        ur, dp = mt.ComputeStereoMatches(exL, exR, L, R, bf, mb)
        return int((ur >= 0).sum())

    for _ in range(10):
        stereo_frame_threads()
    t2 = []
    for _ in range(100):
        t0 = time.perf_counter()
        nmatch2 = stereo_frame_threads()
        t2.append(time.perf_counter() - t0)
    pool.shutdown()
    if nmatch2 != nmatch:
        raise RuntimeError("bench: threaded stereo front-end disagrees with the sequential one")
    out["C2_euroc_stereo_752x480_1200"] = {
        "ms_per_stereo_frame_e2e": 1e3 * float(np.median(tt)), "ms_per_stereo_frame_e2e_two_threads": 1e3 * float(np.median(t2)),
        "ms_stereo_match_e2e": 1e3 * float(np.median(tm)),
        "keypoints_left": nl, "keypoints_right": nr, "stereo_matches": nmatch,
        "note": "extract L + extract R (host in/out; sequential calls, and on two host threads like src/Frame.cc:127-131) + orbm_stereo_match (row search + SAD + outliers)"}
    R, rkind = matcher_reference()
    if R is not None:
        # the reference's own stereo Frame constructor (src/Frame.cc:113-165: two extractor threads, ComputeStereoMatches,
        # grid) on the same pair; its match count must equal the GPU path's
        R.set_timing(True)
        rows_, cols_ = shape
        tc_, tm_ = [], []
        for _ in range(12):
            ref = R.stereo_frame(left, right, cfg, fx, fx, cols_ / 2.0, rows_ / 2.0, bf)
            tc_.append(R.last_ms(1)); tm_.append(R.last_ms(2))
        R.set_timing(False)
        nref = int((ref.u_right >= 0).sum())
        if nref != nmatch or ref.n != nl:
            raise RuntimeError("bench: reference stereo frame (%d keypoints, %d matches) != GPU (%d, %d)" % (ref.n, nref, nl, nmatch))
        out["C2_euroc_stereo_752x480_1200"]["cpu_baseline"] = {
            "kind": "reference", "cores": 2, "unit": "ms per stereo frame", "value": float(np.median(tc_)),
            "compute_stereo_matches_ms": float(np.median(tm_)), "stereo_matches": nref, "matches_equal_gpu": True,
            "sample": "12 x Frame(imLeft, imRight, ...) of the unmodified src/Frame.cc + src/ORBextractor.cc over the mini-cv shim (%s); "
                      "ComputeStereoMatches (src/Frame.cc:673-885) timed once more on the constructed frame"
                      % ("-O3 -march=x86-64-v3" if rkind == "ref_fast" else "-O2")}
    exL.close(); exR.close()
    # ---- C3 ----
    shape, cfg = synth.SHAPES["kitti"], synth.PRESETS["kitti"]
    rows, cols = shape
    Bk = 64
    frames = make_frames(shape, 8, Bk)
    ex = pkg.ORBextractor(*cfg, device=local)
    cap = ex.max_keypoints(rows, cols)
    d_imgs = torch.from_numpy(frames).to(dev)
    d_kps = torch.empty((Bk, cap, 6), dtype=torch.float32, device=dev)
    d_desc = torch.empty((Bk, cap, 32), dtype=torch.uint8, device=dev)
    d_counts = torch.empty(Bk, dtype=torch.int32, device=dev)
    st = torch.cuda.ExternalStream(ex.stream, device=dev)
    for _ in range(3):
        ex.extract_batch_device(d_imgs.data_ptr(), Bk, rows, cols, d_kps.data_ptr(), d_desc.data_ptr(), d_counts.data_ptr(), cap)
    ex.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(10):
        ex.extract_batch_device(d_imgs.data_ptr(), Bk, rows, cols, d_kps.data_ptr(), d_desc.data_ptr(), d_counts.data_ptr(), cap)
    e1.record(st)
    ex.sync()
    fps = Bk * 10 / (e0.elapsed_time(e1) * 1e-3)
    kps, desc = ex(frames[0])
    n = len(kps)
    w_inv, h_inv = float(np.float32(64.0) / np.float32(cols)), float(np.float32(48.0) / np.float32(rows))
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
    mt2 = pkg.ORBmatcher(0.8, True, device=local)
    for _ in range(5):
        nm, _o = mt2.SearchByProjection(frame, mps, 3.0)
    ts = []
    for _ in range(50):
        t0 = time.perf_counter()
        nm, _o = mt2.SearchByProjection(frame, mps, 3.0)
        ts.append(time.perf_counter() - t0)
    out["C3_kitti_1241x376_2000"] = {
        "extract_frames_per_s_device_resident": fps, "frames_per_step": Bk, "keypoints_per_frame": n,
        "search_by_projection_ms_e2e": 1e3 * float(np.median(ts)), "map_points": nmp, "matches": int(nm),
        "note": "SearchByProjection(Frame, 4000 local map points, th=3) through the host C-ABI (uploads included)"}
    if R is not None:
        # the reference's own ORBmatcher::SearchByProjection(Frame, MapPoints, th) (src/ORBmatcher.cc:61-173) on the same frame and
        # map points, timed inside the harness around the call alone
        tr = []
        for _ in range(12):
            nref, _m = R.search_by_projection(frame, mps, 3.0, 0.8)
            tr.append(R.last_ms(0))
        if int(nref) != int(nm):
            raise RuntimeError("bench: reference SearchByProjection found %d matches, the GPU path %d" % (nref, nm))
        out["C3_kitti_1241x376_2000"]["cpu_baseline"] = {
            "kind": "reference", "cores": 1, "unit": "ms per SearchByProjection call", "value": float(np.median(tr)),
            "matches": int(nref), "matches_equal_gpu": True,
            "sample": "12 calls of the unmodified src/ORBmatcher.cc:61-173 on Frame / MapPoint objects built from the same arrays (%s)"
                      % ("-O3 -march=x86-64-v3" if rkind == "ref_fast" else "-O2")}
    ex.close(); mt.close(); mt2.close()
    return out


# ------------------------------------------------------------------------------------------------------------------
def c4_leg(pkg, torch, dev, local, rank, world, barrier):
    """BASELINE.json config 4: re-extraction of 4096 synthetic 1241x376 frames, 2000 features, sharded by frame across the
    ranks (strong scaling: 4096 / N frames per rank, no data-path collective), through the reference-facing host call
    orbx_extract_batch with pinned host input and output.  Returns (seconds, frames of this rank, keypoints per frame)."""
    shape, cfg = synth.SHAPES["kitti"], synth.PRESETS["kitti"]
    rows, cols = shape
    total = 4096
    mine = len(sharding.frame_shard(total, rank, world))
    call = min(512, mine)
    ncalls = (mine + call - 1) // call
    frames = make_frames(shape, 8, call)
    # two host threads with one extractor handle (and one set of page-locked buffers) each share the rank's calls, like the
    # end-to-end section of run_ours (the reference's own threading, src/Frame.cc:127-131)
    nth = 2
    exs = [pkg.ORBextractor(*cfg, device=local) for _ in range(nth)]
    bigs = [torch.from_numpy(frames.copy()).pin_memory() for _ in range(nth)]
    bouts = [e.alloc_outputs(call, rows, cols, pinned=True) for e in exs]
    jobs = C4_JOBS * world         # the 4096-frame job repeated so that the timed region is >= 0.5 s at every N
    sizes = []
    for _ in range(jobs):
        done = 0
        for i in range(ncalls):
            n = min(call, mine - done)
            sizes.append(n)
            done += n

    def worker(t, todo):
        e, big, bout = exs[t], bigs[t], bouts[t]
        for n in todo:
            e.extract_batch(big.numpy()[:n], out=(bout[0][:n], bout[1][:n], bout[2][:n]))

    def run(parts):
        ts = [threading.Thread(target=worker, args=(t, parts[t])) for t in range(nth)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()

    t_w = time.perf_counter()
    while time.perf_counter() - t_w < 0.3:   # host path warm-up by time (see the e2e section of run_ours)
        run([[call], [call]])
    barrier()
    t0 = time.perf_counter()
    run([sizes[t::nth] for t in range(nth)])
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / jobs
    barrier()
    kp = float(bouts[0][2][: min(call, mine)].mean())
    # parity of the timed calls: the 8 distinct frames at the head of thread 0's last output batch against the reference's own
    # ORBextractor.cc (rank 0; every rank extracts the same 8 distinct frames tiled)
    parity = None
    if rank == 0:
        nd = min(8, call, mine)
        ours = output_digest(bouts[0][0], bouts[0][1], bouts[0][2], nd)
        ref_digest, ref_src = reference_digest(frames[:nd], cfg, exs[0].max_keypoints(rows, cols))
        parity = {"frames": nd, "digest": ours, "reference_digest": ref_digest, "reference": ref_src,
                  "equal": (ours == ref_digest) if ref_digest else None}
        if ref_digest and ours != ref_digest:
            raise RuntimeError("bench: config 4 outputs differ from the reference's ORBextractor.cc on the %d distinct frames" % nd)
    for e in exs:
        e.close()
    del bigs, bouts
    return dt, mine, kp, parity


def bind_to_gpu_numa_node(index):
    """One process per GPU: pin the process to the CPUs next to its GPU (NVML's ideal affinity) before the page-locked host
    buffers are allocated, so that first touch places them on the GPU's NUMA node.  Matters when eight ranks stream frames
    from host memory at once; harmless otherwise.  ORBX_NO_AFFINITY=1 disables it."""
    if os.environ.get("ORBX_NO_AFFINITY"):
        return
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = [v.strip() for v in os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",") if v.strip()]
        phys = int(vis[index]) if index < len(vis) and vis[index].isdigit() else index
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(phys))
    except Exception:
        pass


def run_ours(args):
    import torch

    import orbslam2commentedbyxcm_b200 as pkg

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    bind_to_gpu_numa_node(local)

    shape = synth.SHAPES[WORKLOAD]
    cfg = synth.PRESETS[WORKLOAD]
    rows, cols = shape
    B = args.batch
    frames = make_frames(shape, N_DISTINCT, B)
    # each rank gets its own rotation of the frame set (independent units, no data-path collective)
    frames = np.roll(frames, rank, axis=0).copy()

    ex = pkg.ORBextractor(*cfg, device=local)
    cap = ex.max_keypoints(rows, cols)
    stream = torch.cuda.ExternalStream(ex.stream, device=dev)

    # ---- device-resident buffers ----
    d_imgs = torch.from_numpy(frames).to(dev)
    d_kps = torch.empty((B, cap, 6), dtype=torch.float32, device=dev)
    d_desc = torch.empty((B, cap, 32), dtype=torch.uint8, device=dev)
    d_counts = torch.empty(B, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()

    # A second extractor object with its own stream, arenas and outputs, as the reference drives two ORBextractor objects from
    # two threads in stereo (src/Frame.cc:127-131): the passes of a step alternate between the two handles, so the kernels of
    # consecutive batches overlap (the barrier-bound quad-tree and the kernel tails of one batch run beside the pyramid / FAST
    # of the other; tools/two_lane_probe.py: 291 k -> 298 k frames/s at 512-frame batches).
    ex_b = pkg.ORBextractor(*cfg, device=local)
    stream_b = torch.cuda.ExternalStream(ex_b.stream, device=dev)
    d_kps_b, d_desc_b, d_counts_b = torch.empty_like(d_kps), torch.empty_like(d_desc), torch.empty_like(d_counts)

    def pass_device(which=0):
        if which == 0:
            ex.extract_batch_device(d_imgs.data_ptr(), B, rows, cols, d_kps.data_ptr(), d_desc.data_ptr(), d_counts.data_ptr(), cap)
        else:
            ex_b.extract_batch_device(d_imgs.data_ptr(), B, rows, cols, d_kps_b.data_ptr(), d_desc_b.data_ptr(), d_counts_b.data_ptr(), cap)

    def step_device():
        for i in range(args.inner):
            pass_device(i & 1)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step_device()
    ex.sync()
    ex_b.sync()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    e0, e1, e1b = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record(stream)          # both streams are idle here (barrier above)
    for _ in range(args.steps):
        step_device()
    e1.record(stream)
    e1b.record(stream_b)
    ex.sync()
    ex_b.sync()
    barrier()
    ms = max(e0.elapsed_time(e1), e0.elapsed_time(e1b))
    ex_b.close()
    launches = ex.last_launch_count() * args.steps * args.inner
    counts = d_counts.cpu().numpy()
    # parity of the timed batch: digest of the device-resident outputs of the N_DISTINCT distinct frames (rank 0 holds them
    # unrotated) against the reference's own extractor on the same frames
    parity = None
    if rank == 0:
        nd = min(N_DISTINCT, B)
        kps_h = d_kps[:nd].cpu().numpy().view(pkg.KP_DTYPE).reshape(nd, cap)
        desc_h = d_desc[:nd].cpu().numpy()
        ours_digest = output_digest(kps_h, desc_h, counts, nd)
        digest_b = output_digest(d_kps_b[:nd].cpu().numpy().view(pkg.KP_DTYPE).reshape(nd, cap), d_desc_b[:nd].cpu().numpy(),
                                 d_counts_b.cpu().numpy(), nd)
        if digest_b != ours_digest:
            raise RuntimeError("bench: the two extractor handles disagree")
        ref_digest, ref_src = reference_digest(frames[:nd], cfg, cap)
        parity = {"frames": nd, "digest": ours_digest, "reference_digest": ref_digest, "reference": ref_src,
                  "equal": (ours_digest == ref_digest) if ref_digest else None}
        if ref_digest and ours_digest != ref_digest:
            raise RuntimeError("bench: the timed batch's keypoints / descriptors differ from the reference extractor")

    d_kps_b = d_desc_b = d_counts_b = None
    # ---- per-stage timing (CUDA events between stages on the handle's stream; serialised, no graph) ----
    ex.set_profiling(True)
    stage_ms = {}
    reps = max(3, min(args.steps, 10))
    for i in range(reps + 1):
        pass_device()
        ex.sync()
        if i == 0:
            continue
        for k, v in ex.stage_times().items():
            stage_ms[k] = stage_ms.get(k, 0.0) + v / reps
    ex.set_profiling(False)

    # ---- end to end through the host C-ABI (pinned host input) ----
    EB = max(B, (args.e2e_batch // B) * B)   # frames per host call: the B-frame buffer tiled
    h_imgs = torch.from_numpy(np.concatenate([frames] * (EB // B), axis=0) if EB > B else frames).pin_memory()
    out = ex.alloc_outputs(EB, rows, cols, pinned=True)
    e2e_steps = max(2, args.steps)
    # warm-up: the first ~100 ms of back-to-back host calls run ~30 % slower than the steady state (PCIe link / host side
    # ramping up after the device-resident phase, tools/e2e_nvml_probe.py), so warm up by time, not only by count
    t_w = time.perf_counter()
    n_w = 0
    while n_w < max(args.warmup, 3) or time.perf_counter() - t_w < 0.4:
        ex.extract_batch(h_imgs.numpy(), out=out)
        n_w += 1
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps * args.e2e_inner):
        ex.extract_batch(h_imgs.numpy(), out=out)
    torch.cuda.synchronize()
    e2e_single_s = time.perf_counter() - t0
    barrier()
    # The headline end-to-end figure: the same host calls from TWO host threads, one extractor handle (and one set of page-locked
    # buffers) each -- how the reference itself drives the extractor (two ORBextractor objects on two std::threads,
    # src/Frame.cc:127-131).  The fill and drain of one call's H2D | kernels | D2H pipeline overlap the other call's transfers,
    # so the host link stays busy (tools/e2e_two_threads_probe.py: 158 k -> 171 k frames/s at 512 frames per call).
    ex_t = pkg.ORBextractor(*cfg, device=local)
    h_imgs_t = torch.from_numpy(h_imgs.numpy().copy()).pin_memory()
    out_t = ex_t.alloc_outputs(EB, rows, cols, pinned=True)
    e2e_threads = 2
    calls_per_thread = max(1, (5 * e2e_steps * args.e2e_inner) // (4 * e2e_threads))   # 25 % more calls: the region stays >= 0.5 s

    def e2e_worker(which, reps):
        for _ in range(reps):
            if which == 0:
                ex.extract_batch(h_imgs.numpy(), out=out)
            else:
                ex_t.extract_batch(h_imgs_t.numpy(), out=out_t)

    def e2e_run(reps):
        ts = [threading.Thread(target=e2e_worker, args=(i, reps)) for i in range(e2e_threads)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()

    t_w = time.perf_counter()
    while time.perf_counter() - t_w < 0.3:
        e2e_run(2)
    barrier()
    t0 = time.perf_counter()
    e2e_run(calls_per_thread)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    e2e_calls = calls_per_thread * e2e_threads
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    if rank == 0 and parity is not None:
        nd = parity["frames"]
        parity["e2e_digest_equal"] = (output_digest(out[0], out[1], out[2], nd) == parity["digest"] and
                                      output_digest(out_t[0], out_t[1], out_t[2], nd) == parity["digest"])
        if not parity["e2e_digest_equal"]:
            raise RuntimeError("bench: the host-path outputs differ from the device-resident outputs")
    ex_t.close()

    # ---- single-frame latency through the host C-ABI ----
    one = frames[:1].copy()
    out1 = ex.alloc_outputs(1, rows, cols)
    for _ in range(20):
        ex.extract_batch(one, out=out1)
    lat = []
    for _ in range(200):
        t = time.perf_counter()
        ex.extract_batch(one, out=out1)
        lat.append(time.perf_counter() - t)
    lat_ms = 1e3 * float(np.median(lat))

    # ---- brute-force Hamming (config 5: 2000 queries x 1M map descriptors, the map sharded by rank) ----
    # default path = tensor cores (tcgen05 int8 GEMM on +-1 expanded descriptors + best / second epilogue, the shard expanded
    # once by orbm_db_create); the integer-pipe kernel (LOP3 + POPC) is timed beside it and both must agree bit for bit
    m = pkg.ORBmatcher(0.6, True, device=local)
    ndb_total, nq = 1_000_000, 2000
    lo, hi = sharding.db_shard(ndb_total, rank, world)
    per = hi - lo
    db = synth.gen_descriptors(per, seed=12345 + rank)
    q, _ = synth.gen_queries(synth.gen_descriptors(4096, seed=777), nq)
    ddb = torch.from_numpy(db).to(dev)
    dq = torch.from_numpy(q).to(dev)
    pb = torch.empty(nq, dtype=torch.int32, device=dev)
    pi = torch.empty_like(pb)
    ps = torch.empty_like(pb)
    mstream = torch.cuda.ExternalStream(m.stream, device=dev)
    b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    m.PrepareDatabase(d_db=ddb.data_ptr(), ndb=per).close()   # first call: module load, tensor-map entry point
    b0.record(mstream)
    prep = m.PrepareDatabase(d_db=ddb.data_ptr(), ndb=per)
    b1.record(mstream)
    m.sync()
    expand_ms = b0.elapsed_time(b1)

    ob, oi, os_ = torch.empty_like(pb), torch.empty_like(pb), torch.empty_like(pb)
    ib, ii, is_ = torch.empty_like(pb), torch.empty_like(pb), torch.empty_like(pb)
    if dist is not None:
        # N > 1: the search, the exchange of the per-shard (best, idx, second) triples and the merge are ONE stream of
        # device work per rank: the shard search, then k_merge_exchange (P2P stores into every peer's buffer over NVLink,
        # wait for the peers' triples, deterministic merge) -- no NCCL call on the data path.
        m.exchange_connect_group(world, rank, nq)

        def bf_step():
            m.bruteforce_sharded_tc_device(prep, dq.data_ptr(), nq, lo, ob.data_ptr(), oi.data_ptr(), os_.data_ptr())

        def bf_int_step():
            m.bruteforce_sharded_device(dq.data_ptr(), nq, ddb.data_ptr(), per, lo, ib.data_ptr(), ii.data_ptr(), is_.data_ptr())
    else:
        def bf_step():
            m.bruteforce_tc_device(prep, dq.data_ptr(), nq, lo, ob.data_ptr(), oi.data_ptr(), os_.data_ptr())

        def bf_int_step():
            m.bruteforce_device(dq.data_ptr(), nq, ddb.data_ptr(), per, lo, ib.data_ptr(), ii.data_ptr(), is_.data_ptr())

    def timed(fn, reps):
        for _ in range(3):
            fn()
        m.sync()
        barrier()
        b0.record(mstream)
        for _ in range(reps):
            fn()
        b1.record(mstream)
        m.sync()
        barrier()
        return b0.elapsed_time(b1) / reps

    # burst figure first (the GPU has been idle since the last host call): 10 searches timed one by one.  The long run below
    # is power-capped on these boxes (tools/tc_sustained_probe.py: 0.37 ms per search for <= 100 searches at 1965 MHz, 0.42 ms
    # at ~1550 MHz / sw_power_cap once the tensor pipe has been busy for ~0.3 s), which is what MEASURED_PEAKS.json's
    # burst / sustained bf16 figures distinguish too.
    for _ in range(3):
        bf_step()
    m.sync()
    burst = []
    for _ in range(10):
        b0.record(mstream)
        bf_step()
        b1.record(mstream)
        m.sync()
        burst.append(b0.elapsed_time(b1))
    bf_burst_ms = float(np.median(burst))
    bf_ms = timed(bf_step, 1500 if world == 1 else 3000)      # >= 0.5 s of searches at every N (sustained)
    bf_int_ms = timed(bf_int_step, 150)
    if not (torch.equal(ib, ob) and torch.equal(ii, oi) and torch.equal(is_, os_)):
        raise RuntimeError("brute force: tensor-core and integer-pipe kernels disagree")
    unsharded_equal = None
    nccl_ms = None
    if dist is not None:
        if m.exchange_timed_out():
            raise RuntimeError("sharded search: a peer never answered")
        # like-for-like baseline: the same shard search, then ONE all-gather of the packed (best, idx, second) triples and the
        # merge kernel, all enqueued on the matcher's stream, CUDA-event timed, no host synchronisation inside
        packed = torch.empty((3, nq), dtype=torch.int32, device=dev)
        gathered = torch.empty((world, 3, nq), dtype=torch.int32, device=dev)
        nb, ni, ns = torch.empty_like(pb), torch.empty_like(pb), torch.empty_like(pb)

        def nccl_step():
            m.bruteforce_tc_device(prep, dq.data_ptr(), nq, lo, packed[0].data_ptr(), packed[1].data_ptr(), packed[2].data_ptr())
            with torch.cuda.stream(mstream):
                dist.all_gather_into_tensor(gathered, packed)
                parts = gathered.permute(1, 0, 2).contiguous()     # [3][world][nq]: the layout orbm_merge_partials reads
            m.merge_partials_device(parts[0].data_ptr(), parts[1].data_ptr(), parts[2].data_ptr(), world, nq, nb.data_ptr(),
                                    ni.data_ptr(), ns.data_ptr())

        nccl_ms = timed(nccl_step, 1000)
        torch.cuda.synchronize()
        if not (torch.equal(nb, ob) and torch.equal(ni, oi) and torch.equal(ns, os_)):
            raise RuntimeError("sharded search: fused exchange and NCCL all-gather + merge disagree")
        # independent answer: rank 0 regenerates EVERY shard and scans the concatenated map unsharded on its own GPU with the
        # integer-pipe kernel (no shard search, no exchange, no merge across ranks); the fused-exchange result must equal it
        bf_step()
        m.sync()
        if rank == 0:
            full = np.concatenate([synth.gen_descriptors(sharding.db_shard(ndb_total, r, world)[1] - sharding.db_shard(ndb_total, r, world)[0],
                                                         seed=12345 + r) for r in range(world)])
            dfull = torch.from_numpy(full).to(dev)
            ub, ui, us = torch.empty_like(pb), torch.empty_like(pb), torch.empty_like(pb)
            torch.cuda.synchronize()
            m.bruteforce_device(dq.data_ptr(), nq, dfull.data_ptr(), ndb_total, 0, ub.data_ptr(), ui.data_ptr(), us.data_ptr())
            m.sync()
            if not (torch.equal(ub, ob) and torch.equal(ui, oi) and torch.equal(us, os_)):
                raise RuntimeError("sharded search: the fused-exchange result differs from the unsharded scan of the whole map")
            unsharded_equal = True
            del dfull, full

    # ---- config 4 (strong scaling over the ranks) ----
    c4_s, c4_frames, c4_kp, c4_parity = (None, 0, None, None)
    if not args.no_extra:
        c4_s, c4_frames, c4_kp, c4_parity = c4_leg(pkg, torch, dev, local, rank, world, barrier)

    # ---- max over ranks ----
    t = torch.tensor([ms, e2e_s, bf_ms, bf_int_ms, nccl_ms or 0.0, c4_s or 0.0], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_s, bf_ms, bf_int_ms, nccl_ms_max, c4_s = t.tolist()
    if nccl_ms is not None:
        nccl_ms = nccl_ms_max

    if rank == 0:
        peak, peak_src = measured_peaks()
        ab = algorithmic_bytes(rows, cols, cfg)
        stages = {}
        for k in ("pyramid", "fast", "octree", "blur", "desc"):
            t_ms = stage_ms.get(k, 0.0)
            gbs = ab[k] * B / (t_ms * 1e-3) / 1e9 if t_ms > 0 else 0.0
            stages[k] = {"ms": t_ms, "algo_bytes_per_frame": ab[k], "GB_per_s": gbs, "frac_of_hbm_peak": gbs / peak}
        dom = max(("pyramid", "fast", "octree", "blur", "desc"), key=lambda k: stage_ms.get(k, 0.0))
        frames_step = B * args.inner
        fps = world * frames_step * args.steps / (ms * 1e-3)
        e2e_fps = world * EB * e2e_calls / e2e_s
        e2e_single_fps = world * EB * args.e2e_inner * e2e_steps / e2e_single_s
        d2h = out[0].nbytes + out[1].nbytes + out[2].nbytes
        traffic, traffic_src = measured_traffic(dom, B)
        winst, winst_src = measured_warp_inst(dom, B)
        sm_hz = ((clocks or {}).get("sm_mhz") or 1965.0) * 1e6
        issue_peak = 148 * 4 * sm_hz     # one warp instruction per scheduler (4 per SM) per clock
        issue = None if winst is None else {
            "warp_instructions_per_launch": winst, "achieved": winst / (stage_ms[dom] * 1e-3), "peak": issue_peak,
            "unit": "warp instructions/s", "frac": winst / (stage_ms[dom] * 1e-3) / issue_peak,
            "source": winst_src + "; peak = 148 SMs x 4 schedulers x SM clock under load"}
        five = sum(ab[k] for k in ("pyramid", "fast", "blur", "desc"))   # pyramid + FAST + blur + orientation + descriptor
        peaks = measured_peaks_all()
        # tensor roof of the int8 search: int8 dense = 2 x bf16 dense on B200 (4.5 vs 2.25 PFLOP/s nominal); the measured
        # denominator is 2 x the driver's measured cuBLAS bf16 burst figure
        tc_ops = 2.0 * nq * per * 256   # useful operations: the 48 zero rows that pad 2000 queries to 16 tiles of 128 are not counted
        tc_tops = world * tc_ops / (bf_ms * 1e-3) / 1e12
        int8_peak_burst = 2.0 * peaks["bf16_tflops"]
        int8_peak = 2.0 * peaks.get("bf16_tflops_sustained", peaks["bf16_tflops"])   # the timed region is a long back-to-back run
        # integer-pipe roof of the POPC kernel (profiles/r1_bruteforce_variants.md): 6 POPC per pair, 16 POPC lanes / clk / SM
        popc_roof = 148 * 16 * (clocks["sm_mhz"] or 1965.0) * 1e6 / 6.0 if clocks else None
        line = {
            "metric": "orb_extract_frames_per_s_640x480_1000feat", "value": fps, "unit": "frames/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": dict(CONFIG),
            "step_detail": {
                "step": "%d frames per GPU per step = %d passes over a %d-frame device buffer (%d distinct seeded frames tiled), alternating between two extractor handles (two streams)"
                        % (frames_step, args.inner, B, N_DISTINCT),
                "frames_per_gpu_per_step": frames_step,
                "l2": "working set per pass (%d MB of frames + pyramid + blurred planes) exceeds the 126 MB L2" %
                      (B * (rows * cols + 2 * 1.2e6) / 1e6)},
            "e2e": {"value": e2e_fps, "unit": "frames/s", "h2d_bytes_per_step": int(h_imgs.numel()) * args.e2e_inner,
                    "d2h_bytes_per_step": int(d2h) * args.e2e_inner, "steps": e2e_calls / float(args.e2e_inner),
                    "calls_per_step": args.e2e_inner, "frames_per_call": EB, "timed_region_s": e2e_s,
                    "host_threads": e2e_threads,
                    "how": "orbx_extract_batch from %d host threads, one extractor handle and one set of page-locked buffers each "
                           "(the reference drives two ORBextractor objects from two threads, src/Frame.cc:127-131)" % e2e_threads,
                    "single_thread_value": e2e_single_fps, "single_thread_timed_region_s": e2e_single_s},
            "gpu_launches": launches,
            "clocks": clocks,
            "parity_checked": bool(parity and parity["equal"]),
            "parity": parity,
            "roofline": {
                "kernel": dom, "bound": "hbm", "achieved": stages[dom]["GB_per_s"], "peak": peak, "unit": "GB/s",
                "frac": stages[dom]["frac_of_hbm_peak"], "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": ab[dom] * B,
                "limited_by": "issue slots (the kernel is instruction-bound, not memory-bound: see issue.frac)" if issue and issue["frac"] > stages[dom]["frac_of_hbm_peak"] else "hbm",
                "issue": issue,
                "whole_step_GB_per_s": five * frames_step * args.steps / (ms * 1e-3) / 1e9,
            },
            "stages": stages,
            "keypoints_per_frame": float(counts.mean()),
            "latency_ms_single_frame_e2e": lat_ms,
            "hamming": {
                "pairs_per_s": world * nq * per / (bf_ms * 1e-3), "queries": nq, "db_rows_per_gpu": per,
                "ms_per_search": bf_ms, "unit": "descriptor pairs/s",
                "expand_ms_once_per_shard": expand_ms,
                "pairs_per_s_expansion_amortised_over_100_searches": world * nq * per / ((bf_ms + expand_ms / 100.0) * 1e-3),
                "kernel": "k_bruteforce_tc (tcgen05.mma kind::i8, TMA, TMEM; shard expanded once in %.2f ms)" % expand_ms,
                "roofline": {"bound": "tensor", "achieved": tc_tops / world, "peak": int8_peak, "unit": "TOP/s (int8, per GPU)",
                             "frac": tc_tops / world / int8_peak,
                             "peak_source": "2 x measured bf16 dense SUSTAINED (MEASURED_PEAKS.json bf16_tflops_sustained: the timed "
                                            "region is >= 0.5 s of back-to-back searches, power-capped); nominal int8 4500"},
                "ms_per_search_burst": bf_burst_ms,
                "roofline_burst": {"bound": "tensor", "achieved": tc_ops / (bf_burst_ms * 1e-3) / 1e12, "peak": int8_peak_burst,
                                   "unit": "TOP/s (int8, per GPU)", "frac": tc_ops / (bf_burst_ms * 1e-3) / 1e12 / int8_peak_burst,
                                   "peak_source": "2 x measured bf16 dense burst (MEASURED_PEAKS.json bf16_tflops); searches timed one by one"},
                "integer_pipe_kernel": {
                    "ms_per_search": bf_int_ms, "pairs_per_s": world * nq * per / (bf_int_ms * 1e-3),
                    "roofline": {"bound": "int-pipe", "achieved": nq * per / (bf_int_ms * 1e-3), "peak": popc_roof,
                                 "unit": "pairs/s per GPU", "frac": (nq * per / (bf_int_ms * 1e-3) / popc_roof) if popc_roof else None,
                                 "peak_source": "148 SMs x 16 POPC / clk x SM clock under load / 6 POPC per pair"}},
                "results_equal": True,
                "unsharded_scan_equal": unsharded_equal,
                "exchange": ("none (one shard)" if world == 1 else
                             "fused: k_merge_exchange stores the per-query triples into every peer's buffer over NVLink "
                             "P2P and merges; checked equal to NCCL all-gather + merge"),
                "nccl_allgather_path_ms_per_search": nccl_ms,
                "nccl_allgather_path": None if nccl_ms is None else
                    "same shard search + ONE all_gather_into_tensor of the packed triples + merge kernel on the matcher stream, "
                    "CUDA-event timed, no host synchronisation"},
        }
        if c4_s:
            line.setdefault("configs", {})["C4_batched_4096x1241x376_2000"] = {
                "frames_per_s_e2e": 4096 / c4_s, "seconds_for_4096_frames": c4_s, "jobs_timed": C4_JOBS * world,
                "timed_region_s": c4_s * C4_JOBS * world, "frames_per_rank": c4_frames,
                "keypoints_per_frame": c4_kp, "scaling": "strong (4096 frames sharded by frame over %d rank(s))" % world,
                "parity": c4_parity,
                "note": "pinned host in / pinned host out through orbx_extract_batch from two host threads per rank (one extractor handle each), calls of <= 512 frames (8 distinct seeded "
                        "frames tiled), H2D | kernels | D2H pipelined; max over ranks"}
        if world == 1 and not args.no_extra:
            line.setdefault("configs", {}).update(other_configs(pkg, torch, dev, local))
        # CPU baseline on rank 0 at N = 1 only (bounded sample)
        if world == 1 and not args.no_cpu_baseline:
            line["hamming"]["cpu_baseline"] = hamming_cpu_baseline(pkg, local, q, db[:100000])
            import oracle

            threads = host_threads()
            kind = "ref_fast" if oracle.lib("ref_fast") is not None else "fast"
            sample = max(4 * threads, 16)
            oracle.extract_batch(frames[:threads], *cfg, threads=threads, kind=kind)
            reps_cpu, t_cpu = 0, 0.0
            while t_cpu < 10.0 and reps_cpu < 1000:
                t0 = time.perf_counter()
                oracle.extract_batch(frames[:sample], *cfg, threads=threads, kind=kind)
                t_cpu += time.perf_counter() - t0
                reps_cpu += 1
            line["cpu_baseline"] = {
                "value": sample * reps_cpu / t_cpu, "unit": "frames/s", "cores": threads,
                "kind": "reference" if kind == "ref_fast" else "port",
                "sample": "%d frames x %d passes (%.1f s), %d host threads, %s" % (
                    sample, reps_cpu, t_cpu, threads,
                    "reference src/ORBextractor.cc + mini-cv shim" if kind == "ref_fast" else "oracle port"),
            }
        print(json.dumps(line))
    ex.close()
    m.close()
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--inner", type=int, default=INNER, help="passes over the device buffer per timed step")
    ap.add_argument("--e2e-inner", type=int, default=E2E_INNER, help="host calls per end-to-end step")
    ap.add_argument("--e2e-batch", type=int, default=E2E_BATCH, help="frames per end-to-end host call")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the C2 / C3 side measurements")
    args = ap.parse_args()
    # stdout carries exactly one JSON line: anything a library writes to file descriptor 1 (NCCL prints its version
    # there on some boxes) goes to stderr instead
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(json_fd, "w")
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
