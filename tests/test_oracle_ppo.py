"""Pins the oracle's PPO restatement bit-exactly against the LIVE reference's
mjrl/algos/ppo_clip.py recorded in tests/golden/ppo_*.npz.  CPU only."""
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR
from mjrl_b200.utils import synthetic as syn
from oracle import mjrl_oracle as O

CASES = ["ppo_hopper_mid", "ppo_linear"]


def load(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    g = {k: z[k] for k in z.files}
    g["hidden"] = tuple(int(h) for h in g["hidden"])
    g["transforms"] = {k[3:]: g[k] for k in g if k.startswith("tr_")} or None
    return g


def paths_of(g):
    b = {k: g[k] for k in ("observations", "actions", "rewards", "path_offsets", "terminated")}
    paths = syn.flat_to_paths(b)
    off = g["path_offsets"]
    for i, p in enumerate(paths):
        p["advantages"] = g["advantages"][off[i]:off[i + 1]]
    return paths


@pytest.mark.parametrize("name", CASES)
def test_oracle_ppo_bit_exact(name):
    g = load(name)
    pol = O.GaussianPolicy(int(g["n"]), int(g["m"]), g["hidden"], g["params"], **(g["transforms"] or {}))
    # the golden generator set the mid-training parameters through set_param_values(float32, True, True):
    # in the reference that makes old alias new (see GaussianPolicy.set_param_values); the default-init
    # case never called it, so its first iteration is proper clipped PPO and its second is aliased
    pol.aliased = g["transforms"] is not None
    paths = paths_of(g)
    np.random.seed(int(g["seed"]) + 9)
    out = O.ppo_update(pol, paths, epochs=int(g["epochs"]), mb_size=64, clip_coef=0.2, learn_rate=3e-4)
    np.testing.assert_array_equal(out["new_params"], g["ppo_new_params"])
    out2 = O.ppo_update(pol, paths, state=out["state"], epochs=int(g["epochs"]), mb_size=64)
    np.testing.assert_array_equal(out2["new_params"], g["ppo_new_params_2"])


def test_recorded_minibatches_are_the_global_numpy_stream():
    g = load("ppo_hopper_mid")
    np.random.seed(int(g["seed"]) + 9)
    for row in g["minibatches"]:
        np.testing.assert_array_equal(np.random.choice(int(g["N"]), size=64), row)
