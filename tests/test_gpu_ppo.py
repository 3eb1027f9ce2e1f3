"""GPU parity: PPO (algos/ppo_clip.py + csrc/ppo.cu) vs the live-reference goldens.

The chain of Adam steps is compared weight-for-weight (20-30 steps, lr 3e-4): tolerance 2e-5
absolute (a step is ~lr * g/|g|, so fp32 round-off in a near-zero gradient entry moves a weight by
a fraction of lr)."""
import os
import pickle

import numpy as np
import pytest
import torch

from conftest import GOLDEN_DIR
from mjrl_b200.utils import synthetic as syn

pytestmark = pytest.mark.gpu
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


def build(g):
    from mjrl_b200.utils.gym_env import EnvSpec
    from mjrl_b200.policies.gaussian_mlp import MLP
    from mjrl_b200.policies.gaussian_linear import LinearPolicy
    from mjrl_b200.baselines.linear_baseline import LinearBaseline
    from mjrl_b200.algos.ppo_clip import PPO
    spec = EnvSpec(int(g["n"]), int(g["m"]), int(g["horizon"]), 1)
    pol = MLP(spec, hidden_sizes=g["hidden"], seed=123) if g["hidden"] else LinearPolicy(spec, seed=123)
    np.testing.assert_array_equal(pol.get_param_values(), g["init_params"])
    if g["transforms"] is not None:
        pol.model.set_transformations(**g["transforms"])
        pol.old_model.set_transformations(**g["transforms"])
        pol.set_param_values(g["params"].copy(), set_new=True, set_old=True)   # float32: old aliases new
    return pol, PPO(None, pol, LinearBaseline(spec), clip_coef=0.2, epochs=int(g["epochs"]), mb_size=64,
                    learn_rate=3e-4)


@pytest.mark.parametrize("name", CASES)
def test_ppo_two_iterations_match_reference(name):
    from mjrl_b200.batch import drop_cache
    g = load(name)
    pol, agent = build(g)
    assert pol._old_aliases_new == (g["transforms"] is not None)
    paths = paths_of(g)
    np.random.seed(int(g["seed"]) + 9)
    stats = agent.train_from_paths(paths)
    np.testing.assert_allclose(stats, g["stats"], rtol=1e-12)
    p1 = pol.get_param_values()
    d1 = np.abs(p1 - g["ppo_new_params"]).max()
    moved = np.abs(g["ppo_new_params"] - g["params"]).max()
    print("%s: iteration 1 max|dp| = %.3e (moved %.3e)" % (name, d1, moved))
    assert d1 < 2e-5 and d1 < 0.02 * moved
    assert pol._old_aliases_new                      # set_param_values(params_after_opt fp32, True, True)
    agent = pickle.loads(pickle.dumps(agent))        # Adam moments survive pickling
    agent.policy = pol
    drop_cache()
    agent.train_from_paths(paths)
    d2 = np.abs(pol.get_param_values() - g["ppo_new_params_2"]).max()
    print("   iteration 2 (aliased old, continued Adam state) max|dp| = %.3e" % d2)
    assert d2 < 4e-5


def test_clip_weights_kernel():
    from mjrl_b200 import _lib
    from mjrl_b200._lib import c_float, c_int, ptr, stream_ptr
    ctx = _lib.context()
    dev = torch.device("cuda", ctx.device)
    rng = np.random.default_rng(0)
    B = 200
    lln = rng.standard_normal(B).astype(np.float32) * 0.3
    llo = rng.standard_normal(B).astype(np.float32) * 0.3
    llo[:20] = lln[:20]                                   # exact ties: LR == 1
    adv = rng.standard_normal(B).astype(np.float32)
    t = lambda a: torch.from_numpy(a).to(dev)
    w = torch.empty(B, dtype=torch.float32, device=dev)
    ssum = torch.empty(1, dtype=torch.float64, device=dev)
    d_lln, d_llo, d_adv = t(lln), t(llo), t(adv)     # keep the device buffers alive across the call
    ctx.call("mjrl_ppo_clip_weights", ptr(d_lln), ptr(d_llo), ptr(d_adv), c_int(B), c_float(0.2), ptr(w),
             ptr(ssum), stream_ptr())
    LR = torch.exp(torch.from_numpy(lln) - torch.from_numpy(llo)).requires_grad_(True)
    a = torch.from_numpy(adv)
    surr = torch.min(LR * a, torch.clamp(LR, 0.8, 1.2) * a)
    surr.sum().backward()
    np.testing.assert_allclose(w.cpu().numpy(), LR.grad.numpy(), rtol=0, atol=0)   # d surr / d LR = w
    assert abs(float(ssum.item()) - float(surr.sum())) < 1e-4
