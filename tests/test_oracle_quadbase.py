"""Pins the oracle's QuadraticBaseline (oracle/mjrl_oracle.py) bit-exactly against the LIVE
reference's mjrl/baselines/quadratic_baseline.py recorded in tests/golden/quadbase_*.npz
(tests/golden/make_golden.py --quadbase-only).  CPU only."""
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR
from oracle import mjrl_oracle as O

CASES = ["quadbase_pointmass", "quadbase_hopper"]


def load(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    return {k: z[k] for k in z.files}


def paths_of(g, pre):
    off = g[pre + "path_offsets"]
    return [dict(observations=g[pre + "observations"][off[i]:off[i + 1]],
                 rewards=g[pre + "rewards"][off[i]:off[i + 1]],
                 returns=g[pre + "returns"][off[i]:off[i + 1]],
                 terminated=bool(g[pre + "terminated"][i])) for i in range(len(off) - 1)]


@pytest.mark.parametrize("name", CASES)
def test_oracle_quadratic_baseline_bit_exact(name):
    g = load(name)
    base = O.QuadraticBaseline()
    for k in (0, 1):
        pre = "b%d_" % k
        paths = paths_of(g, pre)
        if k == 0:
            np.testing.assert_array_equal(O.quadratic_features(paths[:2]), g["features_head"])
            n = int(g["n"])
            assert g["features_head"].shape[1] == n + n * (n + 1) // 2 + 5
        np.testing.assert_array_equal(np.concatenate([base.predict(p) for p in paths]), g[pre + "pred_before"])
        errs = base.fit(paths, return_errors=True)
        np.testing.assert_array_equal(np.asarray(errs), g[pre + "errors"])
        np.testing.assert_array_equal(base._coeffs, g[pre + "coeffs"])
        np.testing.assert_array_equal(np.concatenate([base.predict(p) for p in paths]), g[pre + "pred_after"])
    assert g["b0_errors"][0] == 1.0          # never fit: zero predictions (quadratic_baseline.py:45-47)
