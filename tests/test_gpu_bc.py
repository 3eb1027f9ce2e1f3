"""GPU parity: behaviour cloning (algos/behavior_cloning.py = the reference's behavior_cloning.py,
algos/behavior_cloning_2.py = behavior_cloning_2.py) vs the live-reference goldens.

The chain of Adam steps (lr 1e-3) is compared weight for weight after 2 x 20-30 steps: tolerance 5e-5
absolute (an Adam step is ~lr * g/|g|, so fp32 round-off in a near-zero gradient entry moves a weight
by a fraction of lr); the logged full-data losses to 1e-5 relative."""
import os
import pickle

import numpy as np
import pytest

from conftest import GOLDEN_DIR
from mjrl_b200.utils import synthetic as syn

pytestmark = pytest.mark.gpu
CASES = ["bc_mle_hopper", "bc_mse_adroit", "bc_mle_linear"]


def load(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    g = {k: z[k] for k in z.files}
    g["hidden"] = tuple(int(h) for h in g["hidden"])
    g["mode"] = str(g["mode"])
    return g


def build(g):
    from mjrl_b200.utils.gym_env import EnvSpec
    from mjrl_b200.policies.gaussian_mlp import MLP
    from mjrl_b200.policies.gaussian_linear import LinearPolicy
    from mjrl_b200.algos import behavior_cloning, behavior_cloning_2
    spec = EnvSpec(int(g["n"]), int(g["m"]), int(g["horizon"]), 1)
    pol = MLP(spec, hidden_sizes=g["hidden"], seed=123) if g["hidden"] else LinearPolicy(spec, seed=123)
    np.testing.assert_array_equal(pol.get_param_values(), g["init_params"])
    paths = syn.flat_to_paths({k: g[k] for k in ("observations", "actions", "rewards", "path_offsets",
                                                 "terminated")})
    mod = behavior_cloning if g["mode"] == "mle" else behavior_cloning_2
    return pol, mod.BC(paths, pol, epochs=int(g["epochs"]), batch_size=64, lr=1e-3)


@pytest.mark.parametrize("name", CASES)
def test_bc_two_trainings_match_reference(name):
    g = load(name)
    pol, agent = build(g)
    # __init__: transforms and (mse) the log_std reset are host arithmetic -> bit-exact
    np.testing.assert_array_equal(pol.get_param_values(), g["params_after_init"])
    for k in ("in_shift", "in_scale", "out_shift", "out_scale"):
        np.testing.assert_array_equal(getattr(pol.model, k).numpy(), g["tr_" + k])
        np.testing.assert_array_equal(getattr(pol.old_model, k).numpy(), g["tr_" + k])
    np.random.seed(int(g["seed"]) + 9)          # the global stream the reference drew its rows from
    agent.train()
    err = np.abs(pol.get_param_values() - g["bc_new_params"]).max()
    assert err < 5e-5, err
    np.testing.assert_array_equal(pol.get_old_param_values(), pol.get_param_values())
    np.testing.assert_allclose(agent.logger.log["loss"], g["bc_losses"], rtol=1e-5)
    agent = pickle.loads(pickle.dumps(agent))   # Adam moments survive a pickle round trip
    pol = agent.policy
    agent.train()
    err2 = np.abs(pol.get_param_values() - g["bc_new_params_2"]).max()
    assert err2 < 1e-4, err2
    k = len(g["bc_losses"])
    np.testing.assert_allclose(agent.logger.log["loss"][k:], g["bc_losses_2"], rtol=1e-5)
    print("%s: max |dp| %.2e / %.2e" % (name, err, err2))


def test_bc_injected_minibatches_equal_global_stream():
    g = load("bc_mle_hopper")
    pol, agent = build(g)
    steps = int(g["epochs"]) * (int(g["N"]) // 64)
    agent.train(minibatches=g["minibatches"][:steps])
    pol2, agent2 = build(g)
    np.random.seed(int(g["seed"]) + 9)
    agent2.train()
    np.testing.assert_array_equal(pol.get_param_values(), pol2.get_param_values())


def test_bc_loss_matches_oracle_on_host_arrays():
    from oracle import mjrl_oracle as O
    g = load("bc_mse_adroit")
    pol, agent = build(g)
    o_pol = O.GaussianPolicy(int(g["n"]), int(g["m"]), g["hidden"], g["params_after_init"],
                             **{k: g["tr_" + k] for k in ("in_shift", "in_scale", "out_shift", "out_scale")})
    obs, act = g["observations"][:100], g["actions"][:100]
    ref = float(O.bc_loss(o_pol, obs, act, "mse"))
    assert abs(float(agent.loss(obs, act)) - ref) < 1e-5 * abs(ref)


def test_bc_rejects_foreign_optimizer():
    import torch
    g = load("bc_mle_linear")
    pol, _ = build(g)
    from mjrl_b200.algos.behavior_cloning import BC
    paths = syn.flat_to_paths({k: g[k] for k in ("observations", "actions", "rewards", "path_offsets",
                                                 "terminated")})
    with pytest.raises(NotImplementedError):
        BC(paths, pol, optimizer=torch.optim.SGD(pol.trainable_params, lr=0.1))
    agent = BC(paths, pol, optimizer=torch.optim.Adam(pol.trainable_params, lr=3e-3))
    assert agent.hyper["lr"] == 3e-3
