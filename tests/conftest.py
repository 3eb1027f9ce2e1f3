import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
GOLDEN_CASES = ["c1_linear", "c1_linear_mid", "c2_swimmer", "c3_hopper", "c3_hopper_mid",
                "c5_adroit_mid", "c4_ant_mid", "c5_adroit_64"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    g = {k: z[k] for k in z.files}
    g["hidden"] = tuple(int(h) for h in g["hidden"])
    for k in ("n", "m", "N", "horizon", "seed"):
        g[k] = int(g[k])
    g["gamma"], g["gae_lambda"] = float(g["gamma"]), float(g["gae_lambda"])
    tr = {k[3:]: g[k] for k in g if k.startswith("tr_")}
    g["transforms"] = tr or None
    return g


def batch_of(g, tag):
    return {k[len(tag):]: v for k, v in g.items()
            if k.startswith(tag) and k[len(tag):] in
            ("observations", "actions", "rewards", "path_offsets", "terminated")}


def rel_l2(a, b):
    a = np.asarray(a, dtype=np.float64).ravel()
    b = np.asarray(b, dtype=np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


@pytest.fixture(params=GOLDEN_CASES)
def golden(request):
    return load_golden(request.param)
