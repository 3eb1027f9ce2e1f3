"""Pins the oracle's MLPBaseline (oracle/mjrl_oracle.py) bit-exactly against the LIVE reference's
mjrl/baselines/mlp_baseline.py recorded in tests/golden/mlpbase_*.npz.  CPU only."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN_DIR
from oracle import mjrl_oracle as O

CASES = ["mlpbase_1step", "mlpbase_hopper"]


def load(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    return {k: z[k] for k in z.files}


def paths_of(g):
    off = g["path_offsets"]
    return [dict(observations=g["observations"][off[i]:off[i + 1]],
                 rewards=g["rewards"][off[i]:off[i + 1]],
                 returns=g["returns"][off[i]:off[i + 1]],
                 terminated=bool(g["terminated"][i])) for i in range(len(off) - 1)]


def flat(model):
    return np.concatenate([p.data.numpy().ravel() for p in model.parameters()]).astype(np.float32)


class Spec:
    def __init__(self, n):
        self.observation_dim = n


@pytest.mark.parametrize("name", CASES)
def test_oracle_mlpbaseline_bit_exact(name):
    g = load(name)
    paths = paths_of(g)
    torch.manual_seed(int(g["seed"]))
    base = O.MLPBaseline(Spec(int(g["n"])), reg_coef=float(g["reg_coef"]), batch_size=int(g["batch_size"]),
                         epochs=int(g["epochs"]))
    np.testing.assert_array_equal(flat(base.model), g["init_params"])
    np.testing.assert_array_equal(O.mlp_baseline_features(paths).astype(np.float32), g["features"])
    np.testing.assert_array_equal(np.concatenate([base.predict(p) for p in paths]), g["pred_before"])
    np.random.seed(int(g["seed"]) + 1)
    eb, ea = base.fit(paths, return_errors=True)
    np.testing.assert_array_equal(flat(base.model), g["fit_params"])
    np.testing.assert_array_equal(np.asarray([eb, ea], dtype=np.float64), g["errors"])
    np.testing.assert_array_equal(np.concatenate([base.predict(p) for p in paths]), g["pred_after"])
    base.fit(paths)   # Adam state persists across fit() calls
    np.testing.assert_array_equal(flat(base.model), g["fit2_params"])


def test_recorded_permutations_are_the_global_numpy_stream():
    g = load("mlpbase_hopper")
    np.random.seed(int(g["seed"]) + 1)
    for e in range(int(g["epochs"])):
        np.testing.assert_array_equal(np.random.permutation(int(g["N"])), g["perms"][e])
