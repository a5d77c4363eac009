import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    g = {k: z[k] for k in z.files}
    g["descriptor"] = str(g["descriptor"])
    if g["descriptor"] == "sift":
        g["desc"] = g["desc"].astype(np.float32)
    return g


GOLDEN_CASES = ["orb_300", "orb_1000_c1", "orb_700_refK", "sift_400"]


@pytest.fixture(params=GOLDEN_CASES)
def golden(request):
    return load_golden(request.param)


@pytest.fixture(scope="session")
def cuda_dev():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


def record_parity(test, **fields):
    """Per-test parity counts (mask mismatches, pose deviation on mismatching pairs, ...) merged
    into one JSON (default gpurun_out/parity.json; copied to profiles/parity_rNN.json): what the
    parity tests MEASURED, next to what they assert."""
    import json
    path = os.environ.get("VO_PARITY_JSON", os.path.join(ROOT, "gpurun_out", "parity.json"))
    try:
        os.makedirs(os.path.dirname(path), exist_ok=True)
        data = json.load(open(path)) if os.path.exists(path) else {}
        clean = {}
        for k, v in fields.items():
            if isinstance(v, (np.integer,)):
                v = int(v)
            elif isinstance(v, (np.floating,)):
                v = float(v)
            elif isinstance(v, np.ndarray):
                v = v.tolist()
            clean[k] = v
        data[test] = clean
        json.dump(data, open(path, "w"), indent=1, sort_keys=True)
    except OSError:
        pass
