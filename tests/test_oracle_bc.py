"""Pins the oracle's behaviour-cloning restatement bit-exactly against the LIVE reference's
mjrl/algos/behavior_cloning.py ("mle") and behavior_cloning_2.py ("mse") recorded in
tests/golden/bc_*.npz.  CPU only."""
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR
from mjrl_b200.utils import synthetic as syn
from oracle import mjrl_oracle as O

CASES = ["bc_mle_hopper", "bc_mse_adroit", "bc_mle_linear"]


def load(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    g = {k: z[k] for k in z.files}
    g["hidden"] = tuple(int(h) for h in g["hidden"])
    g["mode"] = str(g["mode"])
    return g


def paths_of(g):
    return syn.flat_to_paths({k: g[k] for k in ("observations", "actions", "rewards", "path_offsets",
                                                "terminated")})


@pytest.mark.parametrize("name", CASES)
def test_oracle_bc_bit_exact(name):
    g = load(name)
    pol = O.GaussianPolicy(int(g["n"]), int(g["m"]), g["hidden"], g["init_params"])
    paths = paths_of(g)
    state = O.BCState(pol, paths, mode=g["mode"], lr=1e-3)
    np.testing.assert_array_equal(pol.get_param_values(), g["params_after_init"])
    for k in ("in_shift", "in_scale", "out_shift", "out_scale"):
        np.testing.assert_array_equal(getattr(pol, k).numpy(), g["tr_" + k])
    np.random.seed(int(g["seed"]) + 9)
    out = O.bc_train(state, paths, epochs=int(g["epochs"]), mb_size=64)
    np.testing.assert_array_equal(out["new_params"], g["bc_new_params"])
    np.testing.assert_array_equal(out["losses"].astype(np.float64), g["bc_losses"])
    out2 = O.bc_train(state, paths, epochs=int(g["epochs"]), mb_size=64)
    np.testing.assert_array_equal(out2["new_params"], g["bc_new_params_2"])
    np.testing.assert_array_equal(out2["losses"].astype(np.float64), g["bc_losses_2"])


def test_recorded_minibatches_are_the_global_numpy_stream():
    g = load("bc_mle_hopper")
    np.random.seed(int(g["seed"]) + 9)
    for row in g["minibatches"]:
        np.testing.assert_array_equal(np.random.choice(int(g["N"]), size=64), row)
