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
K1_DRAM_BYTES_PER_LAUNCH = 101683200   # ncu capture of the 999-pair launch (85.3 MB rd + 16.4 MB wr)
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
                 "--format=csv,noheader,nounits", "-lms", "100"],
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
    ctx = mp.get_context("fork")
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
        "scaling": "weak", "vs_baseline": None, "dtype": "u8 popc + f64 geometry",
        "data": "synthetic", "config": {"workload": WORKLOAD, "sample": sample},
        "cpu_baseline": {"value": value, "unit": "pairs/s", "cores": workers, "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": "pairs/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0}), flush=True)


def k2_roofline(pipe, n_pairs, ms_per_step, hbm_peak):
    """K2 (Sampson scoring) as achieved HBM GB/s, the form north_star asks for: every hypothesis
    sample scored reads its pair's correspondences once, 32 bytes each.  Samples scored
    = the iterations RANSAC ran per pair, rounded up to the round it stopped in (16, 32, 64, ...)."""
    it = pipe.n_iters[:n_pairs].cpu().numpy().astype("int64")
    m = pipe.n_match[:n_pairs].cpu().numpy().astype("int64")
    done, bound = 0, []
    cap = min(pipe.max_iters, max(256, 262144 // max(n_pairs, 1)))
    chunk, r = min(16, cap), 0
    while done < pipe.max_iters:
        chunk = min(chunk, pipe.max_iters - done)
        done += chunk
        bound.append(done)
        r += 1
        if r >= 2 and chunk < cap:
            chunk = min(chunk * 2, cap)
    import numpy as np
    scored = np.array(bound)[np.searchsorted(np.array(bound), np.minimum(it, bound[-1]))]
    nbytes = float((scored * m).sum()) * 32.0
    gbs = nbytes / (ms_per_step * 1e-3) / 1e9 if ms_per_step > 0 else None
    return {"kernel": "ransac_score_kernel (K2)", "bound": "hbm", "achieved": gbs, "peak": hbm_peak,
            "unit": "GB/s", "frac": (gbs / hbm_peak) if gbs and hbm_peak else None,
            "algorithmic": "samples scored x matches x 32 B, summed over the pairs of one step",
            "samples_scored_per_pair_mean": float(scored.mean()), "matches_per_pair_mean": float(m.mean()),
            "note": "the correspondences of a pair (<= 64 KB) stay in L1/L2 across its samples: the real "
                    "limiter is the fp64 pipe (about 35 flop per sample and correspondence), not HBM"}


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from visual_odometry_b200 import _lib
    from visual_odometry_b200 import geometry as G
    from visual_odometry_b200.dist import gather_relative_poses
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

    def step_device():
        pipe.run_device(B)
        if world > 1:
            Rg, tg = gather_relative_poses(pipe.R[:B], pipe.t[:B])
            return G.chain_poses(Rg, tg)
        return pipe.poses

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    popc_peak = ctypes.c_double(0.0)
    _lib.check(lib.vo_measure_popc_peak(ctypes.byref(popc_peak), _lib.cur_stream()))
    for _ in range(max(args.warmup, 3)):
        step_device()
    sync_all()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    lib.vo_profile_enable(1)
    evs = []
    sync_all()
    wall0 = time.perf_counter()
    for _ in range(args.steps):
        flush.fill_(0)                                   # L2 flush between timed iterations
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        step_device()
        e1.record()
        evs.append((e0, e1))
    sync_all()
    wall = time.perf_counter() - wall0
    dev_ms = sum(a.elapsed_time(b) for a, b in evs)
    k1_ms, k1_n = ctypes.c_double(0), ctypes.c_int(0)
    lib.vo_profile_read(0, ctypes.byref(k1_ms), ctypes.byref(k1_n))
    k2_ms, k2_n = ctypes.c_double(0), ctypes.c_int(0)
    lib.vo_profile_read(1, ctypes.byref(k2_ms), ctypes.byref(k2_n))
    lib.vo_profile_enable(0)

    # end to end through the public host-buffer API: H2D of the step's frames from pinned memory,
    # the chain, D2H of poses/counts — every step
    for _ in range(2):
        pipe.run_staged(N_FRAMES, B)
    sync_all()
    e2e_evs = []
    h2d = d2h = 0
    for _ in range(args.steps):
        flush.fill_(0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        h2d, d2h = pipe.run_staged(N_FRAMES, B)
        if world > 1:
            Rg, tg = gather_relative_poses(pipe.R[:B], pipe.t[:B])
            G.chain_poses(Rg, tg)[-1].cpu()
        e1.record()
        e2e_evs.append((e0, e1))
    sync_all()
    e2e_ms = sum(a.elapsed_time(b) for a, b in e2e_evs)
    clocks = sampler.stop() if rank == 0 else None

    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms = float(t[0]), float(t[1])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    steps = args.steps
    total_pairs = B * world * steps
    value = total_pairs / (dev_ms * 1e-3)
    k1_avg_s = (k1_ms.value / max(k1_n.value, 1)) * 1e-3
    popc_per_launch = float(B) * N_KP * N_KP * 8.0
    achieved = popc_per_launch / k1_avg_s / 1e12 if k1_avg_s > 0 else None
    peak = popc_peak.value / 1e12
    peaks, peak_src = measured_peaks()
    out = {
        "metric": METRIC, "value": value, "unit": "pairs/s", "n_gpus": world, "steps": steps,
        "warmup": max(args.warmup, 3), "ms_per_step": dev_ms / steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8 popc + f64 geometry",
        "data": "synthetic",
        "config": {"workload": WORKLOAD, "pairs_per_gpu": B, "keypoints": N_KP,
                   "ransac": "conf 0.99, thr 1 px, max 1000 iters, adaptive (OpenCV semantics)",
                   "l2": "256 MB buffer written between timed steps (L2 flush, outside the events)",
                   "parallelism": f"pair-sharded x{world}, one all_gather of [R|t] per step"
                   if world > 1 else "single GPU"},
        "e2e": {"value": total_pairs / (e2e_ms * 1e-3), "unit": "pairs/s",
                "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)},
        "gpu_launches": int(pipe.launches_per_run() * steps + (steps if world > 1 else 0)),
        "clocks": clocks,
        "roofline": {"kernel": "hamming_top2_kernel<4> (K1)", "bound": "int-popc",
                     "achieved": achieved, "peak": peak, "unit": "Tpopc32/s",
                     "frac": (achieved / peak) if achieved and peak else None,
                     "traffic": K1_DRAM_BYTES_PER_LAUNCH,
                     "algorithmic": "N1*N2*8 popc32 per pair x 999 pairs per launch (SURVEY 8d)",
                     "executed": (achieved * 4.5 / 8.0) if achieved else None,
                     "executed_frac": (achieved * 4.5 / 8.0 / peak) if achieved and peak else None,
                     "note": "frac > 1 is real: a carry-save adder tree (3-4 full adders = 6-8 LOP3 "
                             "on the ALU pipe) compresses the 8 XOR words so only 4.5 POPC per "
                             "256-bit distance are issued on average; `executed` counts those. ncu "
                             "(profiles/): XU and ALU pipes both > 80 % busy - together they are "
                             "the bound now",
                     "traffic_source": "ncu --set full dram__bytes_read.sum + dram__bytes_write.sum "
                                       "of this launch (profiles/r1h_k1_hamming_ncu_full_raw.csv); "
                                       "algorithmic 64 MB descriptors + 40 MB keys",
                     "peak_source": "vo_measure_popc_peak live on this GPU (xor+popc+add body); "
                                    "MEASURED_PEAKS.json has no integer entry",
                     "avg_launch_ms": k1_avg_s * 1e3, "launches_timed": k1_n.value,
                     "share_of_step": (k1_ms.value / dev_ms) if dev_ms else None,
                     "hbm_peak_gbs": peaks.get("hbm_gbs"), "hbm_peak_source": peak_src},
        "kernels_ms_per_step": {"hamming_top2": k1_ms.value / steps,
                                "ransac_score": k2_ms.value / steps},
        "roofline_k2": k2_roofline(pipe, B, k2_ms.value / steps, peaks.get("hbm_gbs")),
        "wall_s_timed_region": wall,
    }
    if world == 1 and not args.no_cpu:
        workers = os.cpu_count() or 1
        n_sample = max(workers * 2, min(B, workers * 6))
        rate, per_pair = cpu_reference_rate(seq, n_sample, workers)
        import cv2
        out["cpu_baseline"] = {
            "value": rate, "unit": "pairs/s", "cores": workers, "kind": "port",
            "sample": f"{n_sample} consecutive pairs of this workload, {workers} single-threaded "
                      f"cv2 workers; {per_pair * 1e3:.1f} ms/pair/core; cv2 {cv2.__version__}"}
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
    args = ap.parse_args()
    rank, world, local_rank = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
