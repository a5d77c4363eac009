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
