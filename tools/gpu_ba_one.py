"""One vo_bundle_adjust call at a given size (for ncu launch lists). usage: C P n_fixed iters"""
import sys

import torch

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from ba_problem import make_ba                     # noqa: E402
from visual_odometry_b200.ba import bundle_adjust  # noqa: E402
from test_gpu_ba import _to_dev                    # noqa: E402

C, P, nf, it = (int(a) for a in sys.argv[1:5])
t = _to_dev(make_ba(C, P, 5, n_fixed=nf))
out = bundle_adjust(t["poses"], t["fixed"], t["points"], t["obs_cam"], t["obs_pt"], t["obs_xy"], t["K"], it)
torch.cuda.synchronize()
print(out["iterations"], out["chi2"][-1])
