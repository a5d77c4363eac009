#!/usr/bin/env python
"""bench.py — frame-pairs/sec of the two-view VO hot path (match + E-RANSAC + recoverPose +
triangulation + trajectory chaining) on N B200s, beside the reference's cv2 CPU path.

Contract: `python bench.py --gpus N --steps K --warmup W [--impl reference]` prints ONE JSON line.
Workload = BASELINE.json configs[1]: a 1000-frame synthetic drone sequence, ORB 2000 keypoints,
999 consecutive pairs per GPU (weak scaling: every rank processes its own 1000-frame block of the
sequence; the only collective is the all_gather of per-pair poses for chaining).
A step = one pass of the whole chain over the rank's 999 pairs.
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

N_FRAMES = 1000
N_KP = 2000
SEED = 20261104 + 1     # seed = 20261104 + config index
# dram__bytes_read.sum + dram__bytes_write.sum of the 999-pair K1f launch (ncu --set full,
# profiles/r2t_k1f_full_ncu_raw.csv: 595.19 MB read + 24.75 MB written)
K1F_DRAM_BYTES_PER_LAUNCH = 619940352   # 595.19 MB read + 24.75 MB written
K1F_TRAFFIC_SOURCE = "ncu --set full dram__bytes_read.sum + dram__bytes_write.sum of this launch " \
                     "(profiles/r2t_k1f_full_ncu_raw.csv); algorithmic: 1000 frames x (2048 rows x 128 B " \
                     "+ 10 column tiles x 33 280 B) = 595 MB of operands read once + 24 MB of row / column " \
                     "keys written"
METRIC = "frame-pairs/sec (match+E-RANSAC+pose)"
WORKLOAD = "configs[1]: 1000-frame synthetic drone sequence, ORB 2000 kp, consecutive-pair VO " \
           "with trajectory chaining"


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled while the timed region runs."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown," \
        "clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 6:
                continue
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if r[2 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx,
                "samples": len(sm), "reasons": sorted(reasons)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)), "MEASURED_PEAKS.json"
        except Exception:
            pass
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------- CPU reference arm
def _cpu_pair_worker(args):
    import cv2
    from oracle import reference_path as RP
    cv2.setNumThreads(1)
    d1, k1, d2, k2, K = args
    t0 = time.perf_counter()
    RP.pair_cv2(d1, k1, d2, k2, K, cv2.NORM_HAMMING)
    return time.perf_counter() - t0


def cpu_reference_rate(seq, n_sample, workers):
    """The reference's per-pair cv2 chain (oracle/reference_path.pair_cv2 = the calls of
    src/image_pair.py:234,280,304,332) on `n_sample` pairs of the same workload, one
    single-threaded worker per host core (pairs are independent, so this is the CPU's best case).
    Returns pairs/s over the wall time of the pool."""
    import multiprocessing as mp
    jobs = [(seq.desc[b], seq.kp[b], seq.desc[b + 1], seq.kp[b + 1], seq.K) for b in range(n_sample)]
    # spawn, not fork: the parent may already have started cv2's thread pool (the cv2 figures of
    # the configs block) and holds a CUDA context; a forked child of such a process deadlocks in
    # cv2.setNumThreads
    ctx = mp.get_context("spawn")
    with ctx.Pool(workers) as pool:
        pool.map(_cpu_pair_worker, jobs[:workers])           # warm-up (imports, page-in)
        t0 = time.perf_counter()
        per = pool.map(_cpu_pair_worker, jobs, chunksize=1)
        wall = time.perf_counter() - t0
    return n_sample / wall, float(np.mean(per))


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU path (cv2 chain) on the host cores."""
    if rank != 0:
        return
    from visual_odometry_b200.synth import make_sequence
    import cv2
    workers = os.cpu_count() or 1
    n_sample = max(workers * 2, min(999, workers * 6))
    seq = make_sequence(n_sample + 1, N_KP, SEED)
    rates = []
    for _ in range(args.warmup):
        cpu_reference_rate(seq, min(n_sample, workers), workers)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        r, _ = cpu_reference_rate(seq, n_sample, workers)
        rates.append(r)
    wall = time.perf_counter() - t0
    value = float(np.mean(rates))
    sample = f"{n_sample} consecutive pairs of the same ORB-2000 sequence per step, " \
             f"{workers} single-threaded cv2 workers (multiprocessing), cv2 {cv2.__version__}"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": "pairs/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * wall / max(args.steps, 1), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8 popcount Hamming + f64 geometry (cv2 on the host cores)",
        "data": "synthetic", "config": {"workload": WORKLOAD, "sample": sample},
        "cpu_baseline": {"value": value, "unit": "pairs/s", "cores": workers, "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": "pairs/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0}), flush=True)


def k2_roofline(pipe, n_pairs, ms_per_step, hbm_peak):
    """K2 (Sampson scoring) as achieved HBM GB/s, the form north_star asks for: every hypothesis
    sample scored reads its pair's correspondences once, 32 bytes each.  Samples scored
    = the iterations RANSAC ran per pair, rounded up to the round it stopped in (16, 80, 336, 1000)."""
    it = pipe.n_iters[:n_pairs].cpu().numpy().astype("int64")
    m = pipe.n_match[:n_pairs].cpu().numpy().astype("int64")
    bound = list(np.cumsum(pipe.ransac_schedule(n_pairs)))
    scored = np.array(bound)[np.searchsorted(np.array(bound), np.minimum(it, bound[-1]))]
    nbytes = float((scored * m).sum()) * 32.0
    gbs = nbytes / (ms_per_step * 1e-3) / 1e9 if ms_per_step > 0 else None
    return {"kernel": "ransac_score_kernel (K2)", "bound": "hbm", "achieved": gbs, "peak": hbm_peak,
            "unit": "GB/s", "frac": (gbs / hbm_peak) if gbs and hbm_peak else None,
            "algorithmic": "samples scored x matches x 32 B, summed over the pairs of one step",
            "samples_scored_per_pair_mean": float(scored.mean()), "matches_per_pair_mean": float(m.mean()),
            "note": "the correspondences of a pair (<= 64 KB) stay in L1/L2 across its samples: the real "
                    "limiter is the fp64 pipe (about 35 flop per sample and correspondence), not HBM"}


def cpu_baseline_forms(seq, workers):
    """The three CPU forms BASELINE.md section 2 promises, on bounded samples of the same
    workload, plus what materialising the reference's per-match `Match` tuples costs per pair
    (excluded from the pairs/s on both sides; the reference pays it in
    src/image_pair.py:243-252, our drop-in ImagePair in visual_odometry_b200/image_pair.py)."""
    import cv2
    from oracle import reference_path as RP
    out = {}

    def serial(n_pairs, threads):
        cv2.setNumThreads(threads)
        RP.pair_cv2(seq.desc[0], seq.kp[0], seq.desc[1], seq.kp[1], seq.K, cv2.NORM_HAMMING)
        t0 = time.perf_counter()
        for b in range(n_pairs):
            RP.pair_cv2(seq.desc[b], seq.kp[b], seq.desc[b + 1], seq.kp[b + 1], seq.K, cv2.NORM_HAMMING)
        return n_pairs / (time.perf_counter() - t0)

    out["cv2_1_thread_pairs_s"] = serial(12, 1)
    out["cv2_all_threads_pairs_s"] = serial(24, workers)
    out["cv2_threads_used"] = int(cv2.getNumThreads())
    cv2.setNumThreads(1)
    # Match tuples of one pair, built the way the reference builds them
    from collections import namedtuple
    Match = namedtuple("Match", "featureid1 featureid2 keypoint1 keypoint2 descriptor1 descriptor2 distance color")
    ms = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True).match(seq.desc[0], seq.desc[1])
    kp1 = [tuple(p) for p in seq.kp[0]]
    kp2 = [tuple(p) for p in seq.kp[1]]
    t0 = time.perf_counter()
    for _ in range(5):
        tuples = [Match((0, m.queryIdx), (1, m.trainIdx), kp1[m.queryIdx], kp2[m.trainIdx],
                        seq.desc[0][m.queryIdx], seq.desc[1][m.trainIdx], m.distance,
                        np.random.random(3)) for m in ms]
    out["match_tuples_ms_per_pair"] = (time.perf_counter() - t0) / 5 * 1e3
    out["matches_per_pair"] = len(tuples)
    out["note"] = "pairs/s above exclude the per-match Python tuples on both sides; the drop-in " \
                  "ImagePair builds the same tuples from device results at the same cost"
    return out


def bench_configs(dev, peaks):
    """BASELINE.json configs other than the headline one, measured in this run on this GPU
    (device-resident, CUDA events, 2 warm-up + 3 timed passes each): C1 single pair, C3 SIFT on
    tcgen05 with 4096 fixed hypotheses, C4 ORB 4000, C5 triangulation + batched PnP, N2
    reprojection errors.  Parity for these sizes lives in tests/."""
    import ctypes
    import torch
    from visual_odometry_b200 import _lib, geometry as G
    from visual_odometry_b200.pipeline import PairPipeline
    from visual_odometry_b200.synth import make_sequence
    lib = _lib.load()
    hbm = peaks.get("hbm_gbs")
    res = {}

    def timed(fn, warm=2, reps=3):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize(dev)
        return e0.elapsed_time(e1) / reps

    def prof(kid):
        ms, n = ctypes.c_double(0), ctypes.c_int(0)
        lib.vo_profile_read(kid, ctypes.byref(ms), ctypes.byref(n))
        return ms.value / max(n.value, 1)

    def pipeline_case(frames, n_kp, n_pairs, seed, **kw):
        seq = make_sequence(frames, n_kp, seed, descriptor=kw.get("descriptor", "orb"))
        pipe = PairPipeline(n_kp, n_pairs, frames, K=seq.K, device=dev, **kw)
        pairs = np.stack([np.arange(n_pairs) % (frames - 1), np.arange(n_pairs) % (frames - 1) + 1], 1)
        pipe.stage_host(seq.desc, seq.kp, seq.nkp, pairs.astype(np.int32))
        pipe.run_staged(frames, n_pairs)
        lib.vo_profile_enable(1)
        ms = timed(lambda: pipe.run_device(n_pairs))
        out = {"pairs": n_pairs, "keypoints": n_kp, "ms_per_pass": ms, "pairs_per_s": n_pairs / (ms * 1e-3),
               "mean_matches": float(pipe.n_match[:n_pairs].float().mean()),
               "mean_inliers": float(pipe.n_inl[:n_pairs].float().mean()),
               "match_kernel_ms": prof(2 if kw.get("descriptor") == "sift" else 0)}
        lib.vo_profile_enable(0)
        del pipe
        torch.cuda.empty_cache()
        return out

    try:
        r = pipeline_case(2, 1000, 1, SEED - 1)
        r["what"] = "configs[0]: single pair ORB 1000 kp, latency of one device-resident pass (popc kernel)"
        res["C1"] = r
        r = pipeline_case(9, 5000, 64, SEED + 1, descriptor="sift", max_iters=4096, adaptive=False,
                          match_mode=1, ratio=0.75)
        flop = 2.0 * 5000 * 5000 * 128 * 64
        tf = flop / (r["match_kernel_ms"] * 1e-3) / 1e12 if r["match_kernel_ms"] else None
        r["what"] = "configs[2]: SIFT 5000 kp, L2 knn + ratio 0.75 on tcgen05, E-RANSAC 4096 fixed hypotheses, 64 pairs"
        r["roofline_k6"] = {"kernel": "l8::match_kernel (K6i: integer SIFT data as a uint8 tcgen05 kind::i8 contraction; "
                                      "other float data runs the bf16 kernel l2_top2_tc_kernel)",
                            "bound": "tensor", "achieved": tf,
                            "peak": peaks.get("bf16_tflops"), "unit": "TFLOP/s",
                            "frac": (tf / peaks["bf16_tflops"]) if tf and peaks.get("bf16_tflops") else None,
                            "peak_kind": "bf16_tflops (burst: the kernel is timed alone)",
                            "algorithmic": "2*N1*N2*128 flop per pair x 64 pairs"}
        res["C3"] = r
        r = pipeline_case(65, 4000, 256, SEED + 2)
        ops = 2.0 * 4000 * 4000 * 256 * 256
        r["what"] = "configs[3] per-GPU work: ORB 4000 kp, 256 pairs (of the 4096), tcgen05 fp4 matching"
        r["match_tops_algorithmic"] = ops / (r["match_kernel_ms"] * 1e-3) / 1e12 if r["match_kernel_ms"] else None
        r["match_popc32_per_s_algorithmic"] = 8.0 * 4000 * 4000 * 256 / (r["match_kernel_ms"] * 1e-3) if r["match_kernel_ms"] else None
        res["C4"] = r
    except Exception as ex:   # a config that fails must not take the headline down
        res["error_pipeline"] = repr(ex)

    try:   # C5: 100 k points triangulated (K4, HBM bound at this size: 64 B/point)
        rng = np.random.default_rng(SEED + 3)
        n = 100_000
        Kc = np.array([[1400.0, 0, 960], [0, 1400.0, 540], [0, 0, 1]])
        X = np.c_[rng.uniform(-12, 12, n), rng.uniform(-7, 7, n), rng.uniform(8, 40, n)]
        th = np.array([0.02, -0.03, 0.01])
        ang = np.linalg.norm(th)
        kx = np.array([[0, -th[2], th[1]], [th[2], 0, -th[0]], [-th[1], th[0], 0]]) / ang
        R = np.eye(3) + np.sin(ang) * kx + (1 - np.cos(ang)) * kx @ kx
        t = np.array([0.6, 0.05, 0.1])
        proj = lambda R, t: (lambda x: x[:, :2] / x[:, 2:])((Kc @ (R @ X.T + t[:, None])).T)
        tt = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.float64)).to(dev)
        P1, P2 = tt(Kc @ np.c_[np.eye(3), np.zeros(3)]), tt(Kc @ np.c_[R, t])
        x1, x2 = tt(proj(np.eye(3), np.zeros(3))), tt(proj(R, t))
        ms = timed(lambda: G.triangulate(P1, P2, x1, x2, dehomog=True), warm=3, reps=10)
        gbs = 64.0 * n / (ms * 1e-3) / 1e9
        res["C5_triangulate"] = {"what": "configs[4]: 100 k points, one vo_triangulate launch", "ms": ms,
                                 "points_per_s": n / (ms * 1e-3),
                                 "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s",
                                              "frac": gbs / hbm if hbm else None,
                                              "algorithmic": "64 B per point (32 in, 32 out)",
                                              "note": "6.4 MB per launch: launch latency and the fp64 4x4 "
                                                      "null-vector solve bound it, not HBM"}}
        if hasattr(lib, "vo_pnp_ransac_batch"):
            from visual_odometry_b200 import pnp as PNP
            res["C5_pnp_batch"] = PNP.bench_batch(dev, X, Kc, rng, timed)
    except Exception as ex:
        res["error_c5"] = repr(ex)

    try:   # N2: reprojection errors of a large map (HBM bound)
        M, P, C = 20_000_000, 2_000_000, 18
        g = torch.Generator(device=dev).manual_seed(1)
        points = torch.rand(P, 3, dtype=torch.float64, device=dev, generator=g) * 20 + \
            torch.tensor([-10.0, -10, 10], device=dev, dtype=torch.float64)
        poses = torch.eye(4, dtype=torch.float64, device=dev).reshape(1, 16).repeat(C, 1).contiguous()
        obs_cam = ((torch.arange(M, device=dev) // 4096) % C).to(torch.int32)
        obs_pt = (torch.arange(M, device=dev) // 10).to(torch.int32)
        obs_xy = torch.rand(M, 2, dtype=torch.float64, device=dev, generator=g) * 1000
        Kd = torch.tensor([1400.0, 0, 960, 0, 1400, 540, 0, 0, 1], dtype=torch.float64, device=dev)
        sqerr = torch.empty(M, dtype=torch.float64, device=dev)
        keep = torch.empty(M, dtype=torch.uint8, device=dev)
        total = torch.zeros(1, dtype=torch.float64, device=dev)
        n_keep = torch.zeros(1, dtype=torch.int32, device=dev)
        wsb = lib.vo_reprojection_workspace_bytes(M)
        ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
        ms = timed(lambda: _lib.check(lib.vo_reprojection_errors(
            _lib.ptr(points), P, _lib.ptr(poses), C, _lib.ptr(obs_cam), _lib.ptr(obs_pt), _lib.ptr(obs_xy), M,
            _lib.ptr(Kd), 1.0, _lib.ptr(sqerr), _lib.ptr(keep), _lib.ptr(total), _lib.ptr(n_keep),
            _lib.ptr(ws), wsb, _lib.cur_stream())), warm=2, reps=5)
        gbs = 57.0 * M / (ms * 1e-3) / 1e9
        res["N2_reprojection"] = {"what": "20 M observations, 2 M points, 18 cameras (runs per camera, points in "
                                          "observation order)", "ms": ms,
                                  "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s",
                                               "frac": gbs / hbm if hbm else None,
                                               "algorithmic": "57 B per observation (24 indices + pixel, 24 point "
                                                              "gather, 9 out)"}}
        del points, obs_cam, obs_pt, obs_xy, sqerr, keep, ws
        torch.cuda.empty_cache()
    except Exception as ex:
        res["error_n2"] = repr(ex)

    # ---- N4: SIFT extraction (the reference's live detector, src/visual_slam.py:17), 32 frames 480 x 640
    try:
        from visual_odometry_b200 import synth
        from visual_odometry_b200.sift import SiftDetector
        F, H, W = 32, 480, 640
        imgs = torch.from_numpy(np.stack([synth.textured_image(H, W, 100 + i) for i in range(F)])).to(dev)
        det = SiftDetector(device=dev)
        out = det.detect_batch(imgs)
        lib.vo_profile_enable(2)
        ms = timed(lambda: det.detect_batch(imgs), warm=1, reps=3)
        parts = {}
        for kid, name in ((10, "pyramid"), (11, "extrema_orientation_order"), (12, "descriptors")):
            tot, cnt = ctypes.c_double(0), ctypes.c_int(0)
            lib.vo_profile_read(kid, ctypes.byref(tot), ctypes.byref(cnt))
            parts[name] = tot.value / max(cnt.value, 1)
        lib.vo_profile_enable(0)
        floats = sum(6 * w * h for w, h in det.octaves(W, H))
        # per frame: the pyramid written once, 5 of 6 layers read once as blur sources, all layers
        # read once by the extrema pass, the image read once
        alg = 4.0 * floats * (1 + 5.0 / 6 + 1) + H * W
        gbs = alg * F / (ms * 1e-3) / 1e9
        entry = {"what": "32 frames 480 x 640 -> key points + 128-byte descriptors in cv2's order "
                         "(cv2.SIFT_create() defaults)", "ms": ms, "frames_per_s": F / ms * 1e3,
                 "key_points_per_frame": float(out["n"].float().mean().item()), "stage_ms": parts,
                 "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s",
                              "frac": gbs / hbm if hbm else None,
                              "algorithmic": "%.1f MB per frame (pyramid of %d floats: 1 write + 5/6 blur "
                                             "reads + 1 extrema read)" % (alg / 1e6, floats)}}
        try:
            import cv2
            t0 = time.perf_counter()
            for i in range(4):
                cv2.SIFT_create().detectAndCompute(imgs[i].cpu().numpy(), None)
            entry["cv2_cpu_ms_per_frame"] = (time.perf_counter() - t0) / 4 * 1e3
            entry["cv2_threads"] = cv2.getNumThreads()
        except ImportError:
            pass
        res["N4_sift_extraction"] = entry
        del imgs, det, out
        torch.cuda.empty_cache()
    except Exception as ex:
        res["error_n4"] = repr(ex)
    return res


def strong_c4(dev, rank, world, n_pairs, n_kp, steps):
    """configs[3] as STRONG scaling: ONE shared 4096-pair / 4000-kp sequence, cut into contiguous
    pair blocks (dist.shard_range) with the one-frame halo (dist.frames_for_block); each rank
    uploads only its block's frames, the per-pair [R|t] are all-gathered and chained.  Rank 0 also
    runs the whole sequence unsharded once (untimed) and compares the two trajectories."""
    import torch
    import torch.distributed as dist
    from visual_odometry_b200.dist import shard_range, frames_for_block, PoseGather
    from visual_odometry_b200.pipeline import PairPipeline
    from visual_odometry_b200.synth import make_sequence
    seq = make_sequence(n_pairs + 1, n_kp, SEED + 2)            # the same sequence on every rank
    lo, hi = shard_range(n_pairs, rank, world)
    f_lo, f_hi = frames_for_block(lo, hi)
    nb, nf = hi - lo, f_hi - f_lo
    pipe = PairPipeline(n_kp, nb, nf, K=seq.K, device=dev, chain=False)
    local_pairs = np.stack([np.arange(nb), np.arange(nb) + 1], 1).astype(np.int32)
    pipe.stage_host(seq.desc[f_lo:f_hi], seq.kp[f_lo:f_hi], seq.nkp[f_lo:f_hi], local_pairs)
    gather = PoseGather(nb, dev)
    last = torch.empty(4, 4, dtype=torch.float64, pin_memory=True)

    def after(k):
        poses = gather(pipe.R, pipe.t, time_it=True)
        last.copy_(poses[-1], non_blocking=True)

    pipe.run_stream(nf, nb, 2, after_compute=after)             # warm-up
    torch.cuda.synchronize(dev)
    gather.collective_ms()
    dist.barrier()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    h2d, d2h = pipe.run_stream(nf, nb, steps, after_compute=after)
    e1.record()
    torch.cuda.synchronize(dev)
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    coll = gather.collective_ms()
    traj = gather.poses.clone()
    out = {"what": "configs[3] strong scaling: 4096 pairs x ORB 4000 kp, ONE sequence cut into contiguous "
                   "blocks + one-frame halo; end to end (H2D of the block's frames, chain, all-gather, D2H)",
           "pairs_total": n_pairs, "pairs_per_rank": nb, "frames_per_rank": nf, "steps": steps,
           "ms_per_step": float(ms[0]) / steps, "value": n_pairs * steps / (float(ms[0]) * 1e-3),
           "unit": "pairs/s", "scaling": "strong", "collective_ms": coll,
           "h2d_bytes_per_step_per_rank": int(h2d)}
    del pipe
    torch.cuda.empty_cache()
    if rank == 0:
        full = PairPipeline(n_kp, n_pairs, n_pairs + 1, K=seq.K, device=dev, chain=True)
        full.stage_host(seq.desc, seq.kp, seq.nkp, seq.consecutive_pairs())
        full.run_staged(n_pairs + 1, n_pairs)
        torch.cuda.synchronize(dev)
        out["trajectory_max_abs_diff_vs_unsharded"] = float((full.poses[:n_pairs + 1] - traj).abs().max())
        del full
        torch.cuda.empty_cache()
    return out


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from visual_odometry_b200 import _lib
    from visual_odometry_b200.dist import PoseGather
    from visual_odometry_b200.pipeline import PairPipeline
    from visual_odometry_b200.synth import make_sequence
    import ctypes

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    B = N_FRAMES - 1
    seq = make_sequence(N_FRAMES, N_KP, SEED + rank)   # this rank's block of the sequence
    pipe = PairPipeline(N_KP, B, N_FRAMES, K=seq.K, device=dev, chain=(world == 1))
    pairs = seq.consecutive_pairs()
    pipe.stage_host(seq.desc, seq.kp, seq.nkp, pairs)
    pin = pipe._pinned()
    pipe.desc.copy_(pin["desc"]); pipe.kp.copy_(pin["kp"]); pipe.nkp.copy_(pin["nkp"])
    pipe.pairs.copy_(pin["pairs"])
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)    # > 126 MB L2
    # device-resident steps at N > 1: the all-gather + chaining of step k runs on a side stream
    # beside the matching of step k + 1 (two send / receive buffer sets); every step's collective
    # is executed and lies inside a timed interval (the step after it, or the closing interval)
    gathers = [PoseGather(B, dev), PoseGather(B, dev)] if world > 1 else None
    side = torch.cuda.Stream(device=dev) if world > 1 else None
    pending = []
    last_pose = torch.empty(4, 4, dtype=torch.float64, pin_memory=True)

    def step_device(k=0):
        if world == 1:
            pipe.run_device(B)
            return pipe.poses
        main = torch.cuda.current_stream()

        def gather_previous():
            # the all-gather of step k - 1 starts when THIS step's matching kernel has finished:
            # it then runs beside the RANSAC / pose kernels, whose grids do not mind a busy SM
            if pending:
                idx, ev = pending.pop()
                evm = torch.cuda.Event()
                evm.record(main)
                with torch.cuda.stream(side):
                    side.wait_event(ev)
                    side.wait_event(evm)
                    gathers[idx].collect()

        pipe.run_device(B, after_match=gather_previous)
        g = gathers[k & 1]
        g.pack(pipe.R, pipe.t)
        ev = torch.cuda.Event()
        ev.record(main)
        pending.append((k & 1, ev))
        main.wait_stream(side)
        return g.poses

    def finish_device():
        """the collective of the last step (nothing left to overlap it with)"""
        if pending:
            idx, _ = pending.pop()
            gathers[idx].collect()

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    def prof(kid):
        ms, n = ctypes.c_double(0), ctypes.c_int(0)
        lib.vo_profile_read(kid, ctypes.byref(ms), ctypes.byref(n))
        return ms.value, n.value

    popc_peak = ctypes.c_double(0.0)
    _lib.check(lib.vo_measure_popc_peak(ctypes.byref(popc_peak), _lib.cur_stream()))
    for k in range(max(args.warmup, 3)):
        step_device(k)
    finish_device()
    sync_all()

    # ---- device-resident steps (the `value`): K1f is timed in-stream (two event records) ----
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    lib.vo_profile_enable(1)
    evs = []
    sync_all()
    wall0 = time.perf_counter()
    for k in range(args.steps):
        flush.fill_(0)                                   # L2 flush between timed iterations
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        step_device(k)
        if k == args.steps - 1:
            finish_device()
        e1.record()
        evs.append((e0, e1))
    sync_all()
    wall = time.perf_counter() - wall0
    dev_ms = sum(a.elapsed_time(b) for a, b in evs)
    k1_ms, k1_n = prof(0)
    lib.vo_profile_enable(0)

    # ---- end to end through the public host-buffer API (PairPipeline.run_stream): every step
    # uploads its frames from pinned memory, runs the chain, downloads poses / counts; the upload
    # of step k+1 overlaps the compute of step k (two device input slots) ----
    done_ev = [None, None]
    packed = []                      # (buffer set, event after its pack) of the batch not yet gathered

    def gather_packed(after_event=None):
        # all-gather + chaining + read-back of the last pose on the side stream
        if packed:
            idx, ev = packed.pop()
            with torch.cuda.stream(side):
                side.wait_event(ev)
                if after_event is not None:
                    side.wait_event(after_event)
                poses = gathers[idx].collect(time_it=True)
                last_pose.copy_(poses[-1], non_blocking=True)
                done_ev[idx] = torch.cuda.Event()
                done_ev[idx].record(side)

    def after(k):
        # multi-GPU: pack batch k's poses on the compute stream; its all-gather is enqueued when
        # batch k + 1's matching kernel is in the stream (after_match) and starts when it has ended
        if world > 1:
            main = torch.cuda.current_stream()
            g = gathers[k & 1]
            if done_ev[k & 1] is not None:
                main.wait_event(done_ev[k & 1])          # batch k - 2 has left these buffers
            g.pack(pipe.R, pipe.t)
            ev = torch.cuda.Event()
            ev.record(main)
            packed.append((k & 1, ev))

    def after_match(k):
        if world > 1 and packed:
            evm = torch.cuda.Event()
            evm.record(torch.cuda.current_stream())
            gather_packed(evm)

    def join_side():
        if world > 1:
            gather_packed()                              # the last batch: nothing left to wait for
            torch.cuda.current_stream().wait_stream(side)

    pipe.run_stream(N_FRAMES, B, 3, after_compute=after, after_match=after_match)
    join_side()
    sync_all()
    if gathers:
        for g in gathers:
            g.collective_ms()
    # one step's upload alone, every rank at the same time: what the host side can deliver.  When
    # this is longer than the device step, the end-to-end rate is bound by host->device bandwidth.
    h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    h0.record()
    for name in ("desc", "kp", "nkp"):
        getattr(pipe, name).copy_(pin[name], non_blocking=True)
    h1.record()
    sync_all()
    h2d_alone_ms = h0.elapsed_time(h1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    h2d, d2h = pipe.run_stream(N_FRAMES, B, args.steps, after_compute=after, after_match=after_match)
    join_side()
    e1.record()
    sync_all()
    e2e_ms = e0.elapsed_time(e1)
    collective_ms = max(g.collective_ms() for g in gathers) if gathers else 0.0
    clocks = sampler.stop() if rank == 0 else None

    # ---- one more pass with every kernel scope timed (outside the timed regions) ----
    lib.vo_profile_enable(2)
    for k in range(3):
        step_device(k)
    finish_device()
    torch.cuda.synchronize(dev)
    names = {0: "hamming_f4_match (K1f)", 4: "hamming_f4_expand", 8: "match_compact", 9: "ransac_control",
             3: "ransac_solve (K3)", 1: "ransac_score (K2)", 5: "recover_pose (K5)", 6: "triangulate_pairs (K4)",
             7: "chain"}
    kernels_ms = {nm: prof(k)[0] / 3 for k, nm in names.items()}
    k2_ms = kernels_ms["ransac_score (K2)"]
    lib.vo_profile_enable(0)

    t = torch.tensor([dev_ms, e2e_ms, h2d_alone_ms], dtype=torch.float64, device=dev)
    per_rank = None
    if world > 1:
        allv = torch.empty(world, 3, dtype=torch.float64, device=dev)
        dist.all_gather_into_tensor(allv, t)
        per_rank = allv.cpu().numpy()
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms, h2d_alone_ms = float(t[0]), float(t[1]), float(t[2])

    strong = None
    if world > 1 and not args.no_configs:
        try:
            strong = strong_c4(dev, rank, world, 4096, 4000, 3)
        except Exception as ex:
            strong = {"error": repr(ex)}
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    steps = args.steps
    steps_for_bound = max(steps, 1)
    total_pairs = B * world * steps
    value = total_pairs / (dev_ms * 1e-3)
    k1_avg_s = (k1_ms / max(k1_n, 1)) * 1e-3
    peaks, peak_src = measured_peaks()
    # K1f: algorithmic work = every pair distance once = 2 * N1 * N2 * 256 ops per pair
    # (SURVEY 8d's N1*N2*8 popc32, one multiply-add per bit); the kernel EXECUTES both directions
    # over K = 320, rows padded to 2048 and columns to 10 tiles of 208: 2 * 2 * 2048 * 2080 * 320.
    alg_ops = 2.0 * N_KP * N_KP * 256 * B
    exe_ops = 2.0 * 2 * 2048 * 2080 * 320 * B
    int8_peak = 4.0 * peaks.get("bf16_tflops", 1590.0)     # (the fp4 peak; name kept)
    alg_tops = alg_ops / k1_avg_s / 1e12 if k1_avg_s > 0 else None
    exe_tops = exe_ops / k1_avg_s / 1e12 if k1_avg_s > 0 else None
    popc_rate = float(B) * N_KP * N_KP * 8.0 / k1_avg_s / 1e12 if k1_avg_s > 0 else None
    e2e_step = e2e_ms / steps
    out = {
        "metric": METRIC, "value": value, "unit": "pairs/s", "n_gpus": world, "steps": steps,
        "warmup": max(args.warmup, 3), "ms_per_step": dev_ms / steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "fp4 (e2m1) tensor-core Hamming (exact) + f64 geometry",
        "data": "synthetic",
        "config": {"workload": WORKLOAD, "pairs_per_gpu": B, "keypoints": N_KP,
                   "ransac": "conf 0.99, thr 1 px, max 1000 iters, adaptive (OpenCV semantics)",
                   "l2": "256 MB buffer written between timed steps (L2 flush, outside the events); "
                         "the e2e steps stream 80 MB of fresh input each (> L2 with the 595 MB of operands)",
                   "parallelism": f"pair-sharded x{world}, one all_gather of [R|t] per step (run on a side stream "
                                  f"beside the next step's matching; the last step's inside its own interval)"
                   if world > 1 else "single GPU"},
        "e2e": {"value": total_pairs / (e2e_ms * 1e-3), "unit": "pairs/s",
                "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": e2e_step,
                "h2d_exposed_ms": max(0.0, e2e_step - dev_ms / steps),
                "collective_ms": collective_ms,
                "h2d_alone_ms": h2d_alone_ms,
                "h2d_gbs_per_rank_all_ranks_uploading": (h2d / 1e9) / (h2d_alone_ms * 1e-3) if h2d_alone_ms > 0 else None,
                "bound": "host->device bandwidth (one step's upload alone takes longer than the device step)"
                if h2d_alone_ms > dev_ms / steps_for_bound else "device",
                "api": "PairPipeline.run_stream: per step H2D from pinned host buffers, the chain, D2H of "
                       "poses and counts; the upload of step k+1 overlaps the compute of step k"},
        "gpu_launches": int((pipe.launches_per_run() + (2 if world > 1 else 0)) * steps),
        "clocks": clocks,
        "roofline": {"kernel": "hf4::match_kernel (K1f: Hamming as a tcgen05 kind::mxf4 (e2m1) contraction, "
                               "row operand in TMEM)",
                     "bound": "tensor", "achieved": alg_tops, "peak": int8_peak, "unit": "TOP/s",
                     "frac": (alg_tops / int8_peak) if alg_tops else None,
                     "traffic": K1F_DRAM_BYTES_PER_LAUNCH,
                     "algorithmic": "2*N1*N2*256 ops per pair x 999 pairs per launch (each pair "
                                    "distance once = SURVEY 8d's N1*N2*8 popc32)",
                     "executed": exe_tops, "executed_frac": (exe_tops / int8_peak) if exe_tops else None,
                     "executed_frac_of_nominal_9000": (exe_tops / 9000.0) if exe_tops else None,
                     "note": "the cross-check needs both directions, so the kernel runs TWO MMA passes "
                             "(rows of each side), K padded 256 -> 320 for the index step, rows "
                             "2000 -> 2048, columns 2000 -> 2080: executed = 2.66 x algorithmic. "
                             "Timing variants (profiles/r2t_k1f_check_and_timing_variants.log): every "
                             "128x208x64 MMA takes ~139 clk against a 104 clk floor with the epilogue "
                             "removed; the ranking adds ~20 %",
                     "peak_source": f"4 x bf16_tflops of {peak_src} (tcgen05 kind::mxf4 runs K = 64 per "
                                    "instruction in the cycles kind::f16 needs for K = 16; no fp4 "
                                    "entry is measured by the driver); nominal dense fp4 = 9000",
                     "traffic_source": K1F_TRAFFIC_SOURCE,
                     "avg_launch_ms": k1_avg_s * 1e3, "launches_timed": k1_n,
                     "share_of_step": (k1_ms / dev_ms) if dev_ms else None,
                     "hbm_peak_gbs": peaks.get("hbm_gbs"), "hbm_peak_source": peak_src},
        "roofline_popc": {"what": "the same launch in SURVEY 8d's popc accounting (N1*N2*8 popc32 per "
                                  "pair): the round-1 popc kernel reached 1.37x the POPC issue peak with "
                                  "a carry-save tree; the contraction now runs on the tensor pipe (int8 in K1t, fp4 in K1f)",
                          "achieved": popc_rate, "peak": popc_peak.value / 1e12, "unit": "Tpopc32/s",
                          "frac": (popc_rate / (popc_peak.value / 1e12)) if popc_rate and popc_peak.value else None},
        "kernels_ms_per_step": kernels_ms,
        "roofline_k2": k2_roofline(pipe, B, k2_ms, peaks.get("hbm_gbs")),
        "wall_s_timed_region": wall,
    }
    if per_rank is not None:
        # the headline uses the MAX over ranks; the spread shows what is straggler (one GPU or one
        # PCIe path slower than the others) and what is the collective
        out["per_rank"] = {"device_ms_per_step": [round(float(v) / steps, 4) for v in per_rank[:, 0]],
                           "e2e_ms_per_step": [round(float(v) / steps, 4) for v in per_rank[:, 1]],
                           "h2d_alone_ms": [round(float(v), 4) for v in per_rank[:, 2]]}
    if strong is not None:
        out["strong_c4"] = strong
    if world == 1 and not args.no_configs:
        del flush
        torch.cuda.empty_cache()
        out["configs"] = bench_configs(dev, peaks)
    if world == 1 and not args.no_cpu:
        workers = os.cpu_count() or 1
        n_sample = max(workers * 2, min(B, workers * 6))
        rate, per_pair = cpu_reference_rate(seq, n_sample, workers)
        import cv2
        out["cpu_baseline"] = {
            "value": rate, "unit": "pairs/s", "cores": workers, "kind": "port",
            "sample": f"{n_sample} consecutive pairs of this workload, {workers} single-threaded "
                      f"cv2 workers; {per_pair * 1e3:.1f} ms/pair/core; cv2 {cv2.__version__}"}
        out["cpu_baseline"]["forms"] = cpu_baseline_forms(seq, workers)
        out["cpu_baseline"]["forms"]["multiprocessing_pairs_s"] = rate
    print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-configs", action="store_true",
                    help="skip the `configs` block (C1/C3/C4/C5/N2) and the strong-scaling leg")
    args = ap.parse_args()
    rank, world, local_rank = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
