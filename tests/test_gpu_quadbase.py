"""GPU parity of QuadraticBaseline (csrc/linbase.cu quadratic staging + predict) against the live
reference's outputs (tests/golden/quadbase_*.npz) and the oracle.  Tolerance (north_star): baseline
predictions within 1e-6 relative (fp64)."""
import copy
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR
from test_oracle_quadbase import load, paths_of

pytestmark = pytest.mark.gpu


def _mods():
    from mjrl_b200.baselines.quadratic_baseline import QuadraticBaseline
    from mjrl_b200.utils.gym_env import EnvSpec
    from mjrl_b200.batch import drop_cache
    return QuadraticBaseline, EnvSpec, drop_cache


@pytest.mark.parametrize("name", ["quadbase_pointmass", "quadbase_hopper"])
def test_quadratic_baseline_matches_reference(name):
    g = load(name)
    QuadraticBaseline, EnvSpec, drop_cache = _mods()
    n = int(g["n"])
    base = QuadraticBaseline(EnvSpec(n, 3, int(g["horizon"]), 1))
    for k in (0, 1):
        pre = "b%d_" % k
        drop_cache()
        paths = paths_of(g, pre)
        got = np.concatenate([base.predict(p) for p in paths])
        want = g[pre + "pred_before"]
        scale = max(1.0, float(np.abs(want).max()))
        assert np.max(np.abs(got - want)) / scale < 1e-6
        errs = base.fit(paths, return_errors=True)
        assert abs(errs[0] - g[pre + "errors"][0]) < 1e-6 and abs(errs[1] - g[pre + "errors"][1]) < 1e-6
        drop_cache()
        got = np.concatenate([base.predict(p) for p in paths])
        want = g[pre + "pred_after"]
        scale = max(1.0, float(np.abs(want).max()))
        assert np.max(np.abs(got - want)) / scale < 1e-6
        assert np.asarray(base._coeffs).shape == g[pre + "coeffs"].shape


@pytest.mark.parametrize("n,lens", [(3, [5, 40]), (6, [25] * 40), (12, [300, 211, 1]), (16, [700, 64])])
def test_quadratic_baseline_shapes_vs_oracle(n, lens):
    from oracle import mjrl_oracle as O
    from mjrl_b200.utils import process_samples as ps
    QuadraticBaseline, EnvSpec, drop_cache = _mods()
    drop_cache()
    rng = np.random.default_rng(n * 7 + len(lens))
    paths = [dict(observations=(4 * rng.standard_normal((T, n))).astype(np.float32).astype(np.float64),
                  actions=np.zeros((T, 2)), rewards=rng.standard_normal(T), terminated=False) for T in lens]
    ref = copy.deepcopy(paths)
    O.compute_returns(ref, 0.98)
    ps.compute_returns(paths, 0.98)
    gb, ob = QuadraticBaseline(EnvSpec(n, 2, max(lens), 1)), O.QuadraticBaseline()
    gb.fit(paths)
    ob.fit(ref)
    X = O.quadratic_features(ref)
    want = X @ ob._coeffs
    scale = max(1.0, float(np.abs(want).max()))
    if sum(lens) > 2 * X.shape[1]:
        assert np.max(np.abs(X @ np.asarray(gb._coeffs) - want)) / scale < 1e-6
    gb._coeffs = ob._coeffs.copy()       # predict kernel alone, with the oracle's coefficients
    drop_cache()
    got = np.concatenate([gb.predict(p) for p in paths])
    assert np.max(np.abs(got - want)) / scale < 1e-9


def test_npg_with_quadratic_baseline_like_point_mass_test():
    """The reference's own tests/point_mass_test.py configuration: NPG + MLP(32,32) +
    QuadraticBaseline, 40 paths x 25 steps, gamma 0.95, gae_lambda 0.97, step 0.1 -- two updates
    (the second one predicts with the fitted baseline) against the oracle."""
    from oracle import mjrl_oracle as O
    from mjrl_b200.utils import synthetic as syn
    from mjrl_b200.policies.gaussian_mlp import MLP
    from mjrl_b200.algos.npg_cg import NPG
    QuadraticBaseline, EnvSpec, drop_cache = _mods()
    n, m, hidden, horizon, N = 6, 2, (32, 32), 25, 1000
    spec = EnvSpec(n, m, horizon, 1)
    pol = MLP(spec, hidden_sizes=hidden, seed=123)
    flat = pol.get_param_values()
    o_pol, o_base = O.GaussianPolicy(n, m, hidden, flat), O.QuadraticBaseline()
    base = QuadraticBaseline(spec)
    agent = NPG(None, pol, base, normalized_step_size=0.1)
    for it in range(2):
        drop_cache()
        batch = syn.make_flat_batch(N, n, m, horizon, 77 + it, o_pol.get_param_values(), hidden)
        before = pol.get_param_values().astype(np.float64)
        o_out = O.full_npg_iteration(o_pol, o_base, syn.flat_to_paths(batch), 0.95, 0.97, normalized_step_size=0.1)
        agent.train_step(N, gamma=0.95, gae_lambda=0.97, paths=syn.flat_to_paths(batch))
        step_o = o_out["new_params"].astype(np.float64) - before
        step_g = pol.get_param_values().astype(np.float64) - before
        rel = np.linalg.norm(step_g - step_o) / np.linalg.norm(step_o)
        assert rel < 5e-3, (it, rel)          # 10 fp32 CG iterations: SURVEY H1 noise floor
        assert abs(agent.last_update["kl_dist"] - o_out["kl_dist"]) < 2e-3
        # keep the two trajectories of parameters identical for the next iteration
        pol.set_param_values(o_out["new_params"].copy(), True, True)
