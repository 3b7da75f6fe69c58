import glob
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def golden_files():
    return sorted(glob.glob(os.path.join(GOLDEN_DIR, "c*_t*_s*.npz")))


def load_golden(path):
    """-> dict with C, T, x, arrays (exported layout, from the reference's own net.c), sd, outputs."""
    z = np.load(path)
    fx = {"C": int(z["C"]), "T": int(z["T"]), "seed": int(z["seed"]), "x": z["x"], "arrays": {}, "sd": {}}
    for k in z.files:
        if k.startswith("arr."):
            fx["arrays"][k[4:]] = z[k]
        elif k.startswith("sd."):
            fx["sd"][k[3:]] = z[k]
        elif k not in ("C", "T", "seed", "x"):
            fx[k] = z[k]
    return fx


@pytest.fixture(params=golden_files(), ids=lambda p: os.path.basename(p)[:-4])
def golden(request):
    return load_golden(request.param)
