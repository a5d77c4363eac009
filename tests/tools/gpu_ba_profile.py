"""B200: wall time of vo_bundle_adjust (40 LM iterations) at map sizes the reference reaches
(<= 18 cameras), beside the numpy oracle on the host.  Run under gpurun."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from ba_problem import make_ba                     # noqa: E402
from oracle import ba_restate as BA                # noqa: E402
from visual_odometry_b200.ba import bundle_adjust  # noqa: E402
from test_gpu_ba import _to_dev                    # noqa: E402

for C, P, nf in [(2, 800, 1), (6, 2000, 4), (18, 5000, 16), (18, 20000, 16), (24, 50000, 22)]:
    d = make_ba(C, P, 5, n_fixed=nf)
    M = len(d["obs_cam"])
    best = 1e9
    for rep in range(4):
        t = _to_dev(d)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = bundle_adjust(t["poses"], t["fixed"], t["points"], t["obs_cam"], t["obs_pt"],
                            t["obs_xy"], t["K"], 40)
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    line = f"C={C} P={P} M={M}: gpu {best*1e3:.2f} ms for {out['iterations']} iterations, " \
           f"{int(out['trials'].sum())} trials, chi2 {out['chi2_initial']:.1f} -> {out['chi2'][-1]:.1f}"
    if P <= 5000:
        pr = BA.BAProblem(**d)
        t0 = time.perf_counter()
        with np.errstate(all="ignore"):
            log = pr.optimize(40)
        line += f"; numpy oracle {1e3*(time.perf_counter()-t0):.0f} ms -> {log[-1][1]:.1f}"
    print(line, flush=True)
