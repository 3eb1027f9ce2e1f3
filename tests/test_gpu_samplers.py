"""Rollout path on the GPU (north_star (e)): the reference's own rollout loops
(fixtures generated from mjrl/samplers/base_sampler.py and evaluation_sampler.py by
tests/golden/make_golden.py --rollout-only) against (1) the single-environment mirror and (2) the
lock-step sampler fed the reference's noise stream; invariance of the Philox batch to slot and worker
counts; and `train_step` sampling for itself, with the batch filed on the device by the sampler."""
import copy
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import stub_env                                                          # noqa: E402
from mjrl_b200.algos.npg_cg import NPG                                   # noqa: E402
from mjrl_b200.baselines.linear_baseline import LinearBaseline           # noqa: E402
from mjrl_b200.policies.gaussian_linear import LinearPolicy              # noqa: E402
from mjrl_b200.policies.gaussian_mlp import MLP                          # noqa: E402
from mjrl_b200.samplers import base_sampler, trajectory_sampler         # noqa: E402
from mjrl_b200.samplers.lockstep import LockstepSampler                  # noqa: E402
from mjrl_b200.utils.get_environment import register_environment         # noqa: E402
from mjrl_b200.utils.gym_env import EnvSpec                              # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL = dict(rtol=1e-5, atol=1e-5)     # fp32 policy forward here vs torch CPU fp32 in the reference


def _fixture(name):
    fx = dict(np.load(os.path.join(GOLD, name + ".npz")))
    hidden = tuple(int(h) for h in fx["hidden"])
    spec = EnvSpec(5, 2, 30, 1)
    pol = MLP(spec, hidden_sizes=hidden, seed=1) if hidden else LinearPolicy(spec, seed=1)
    pol.set_param_values(fx["flat"].copy(), True, True)
    return fx, pol


def _check(fx, tag, paths):
    lens = [len(p["rewards"]) for p in paths]
    assert lens == list(fx[tag + "_lens"])                                   # segmentation: exact
    assert [bool(p["terminated"]) for p in paths] == list(fx[tag + "_terminated"])
    for key in ("observations", "actions", "rewards"):
        np.testing.assert_allclose(np.concatenate([p[key] for p in paths]), fx[tag + "_" + key], **TOL)
    np.testing.assert_allclose(np.concatenate([p["agent_infos"]["mean"] for p in paths]), fx[tag + "_mean"], **TOL)
    assert np.array_equal(np.concatenate([p["env_infos"]["t"] for p in paths]), fx[tag + "_info_t"])


@pytest.mark.parametrize("name", ["rollout_stub_mlp", "rollout_stub_linear"])
def test_single_env_rollout_matches_reference(name):
    fx, pol = _fixture(name)
    N, seed = int(fx["N"]), int(fx["pegasus_seed"])
    _check(fx, "s", base_sampler.do_rollout(N, pol, T=1e6, env=stub_env.make_stub(), pegasus_seed=seed))
    _check(fx, "e", base_sampler.do_evaluation_rollout(N, pol, T=1e6, env=stub_env.make_stub(), pegasus_seed=seed))


@pytest.mark.parametrize("name", ["rollout_stub_mlp", "rollout_stub_linear"])
@pytest.mark.parametrize("slots,workers", [(5, 0), (4, 2)])
def test_lockstep_with_reference_noise_matches_reference(name, slots, workers):
    fx, pol = _fixture(name)
    N, seed = int(fx["N"]), int(fx["pegasus_seed"])
    ls = LockstepSampler(env=stub_env.make_stub, num_slots=slots, num_workers=workers, worker_timeout=120)
    paths = ls.sample(N, pol, T=1e6, pegasus_seed=seed, noise="reference")
    _check(fx, "s", paths)
    # the device batch is the flat concatenation of exactly these paths
    b = paths.batch
    assert b.N == int(fx["s_lens"].sum()) and b.P == N
    assert np.array_equal(b.h_offsets, np.concatenate([[0], np.cumsum(fx["s_lens"])]))
    assert np.array_equal(b.obs.cpu().numpy(), np.concatenate([p["observations"] for p in paths]).astype(np.float32))
    assert np.array_equal(b.act.cpu().numpy(), np.concatenate([p["actions"] for p in paths]))
    assert np.array_equal(b.rew.cpu().numpy(), np.concatenate([p["rewards"] for p in paths]))
    assert np.array_equal(b.terminated.cpu().numpy().astype(bool), fx["s_terminated"])
    _check(fx, "e", ls.sample(N, pol, T=1e6, pegasus_seed=seed, mode="evaluation"))
    ls.close()


def test_philox_batch_is_independent_of_slots_and_workers():
    _, pol = _fixture("rollout_stub_mlp")
    runs = []
    for slots, workers in ((1, 0), (7, 0), (32, 0), (6, 3)):
        ls = LockstepSampler(env=stub_env.make_stub, num_slots=slots, num_workers=workers, worker_timeout=120)
        runs.append(ls.sample(20, pol, T=30, pegasus_seed=9))
        ls.close()
    for r in runs[1:]:
        for p, q in zip(runs[0], r):
            for key in ("observations", "actions", "rewards"):
                assert np.array_equal(p[key], q[key]), key
        assert torch.equal(runs[0].batch.obs, r.batch.obs) and torch.equal(runs[0].batch.act, r.batch.act)
    other = LockstepSampler(env=stub_env.make_stub, num_slots=7).sample(20, pol, T=30, pegasus_seed=10)
    assert not np.array_equal(other[0]["actions"], runs[0][0]["actions"])


def test_philox_noise_moments():
    _, pol = _fixture("rollout_stub_mlp")
    paths = LockstepSampler(env=stub_env.make_stub_long, num_slots=128, collect_env_infos=False).sample(
        2000, _long_policy(), T=100, pegasus_seed=1)
    act = np.concatenate([p["actions"] for p in paths]).astype(np.float64)
    mean = np.concatenate([p["agent_infos"]["mean"] for p in paths]).astype(np.float64)
    z = (act - mean) / np.exp(-0.5)
    n = z.size
    assert n > 20000
    assert abs(z.mean()) < 4 / np.sqrt(n) and abs(z.std() - 1) < 4 / np.sqrt(2 * n)
    assert abs(np.mean(z ** 3)) < 0.1 and abs(np.mean(z ** 4) - 3) < 0.2
    assert abs(np.corrcoef(z[:-1, 0], z[1:, 0])[0, 1]) < 0.03 and abs(np.corrcoef(z[:, 0], z[:, 1])[0, 1]) < 0.03
    assert paths[0]["env_infos"] == {}


def _long_policy():
    pol = MLP(EnvSpec(11, 3, 100, 1), hidden_sizes=(64, 64), seed=3, init_log_std=-0.5)
    return pol


def test_train_step_samples_for_itself_and_reuses_the_device_batch():
    register_environment("stub-v0", stub_env.make_stub)
    env = stub_env.make_stub()
    _, pol = _fixture("rollout_stub_mlp")
    pol2 = copy.deepcopy(pol)
    agents = []
    for p in (pol, pol2):
        bl = LinearBaseline(env.spec)
        agents.append(NPG(env, p, bl, normalized_step_size=0.05, seed=123, save_logs=False))
    # agent 0 samples inside train_step (lock-step sampler, 2 workers); the batch arrives resident
    stats = agents[0].train_step(N=16, env_name="stub-v0", T=30, num_cpu=2)
    assert len(stats) == 5 and stats[-1] == 16 and agents[0].seed == 123 + 16
    # agent 1 is handed the same trajectories as a plain list of dicts (re-flattened and uploaded)
    paths = trajectory_sampler.sample_paths_parallel(16, pol2, 30, "stub-v0", 123, 2, suppress_print=True)
    assert paths.batch is not None and paths.batch.h2d_bytes < 16 * 30 * 8 + 1024
    plain = [dict(p) for p in paths]
    stats2 = agents[1].train_step(N=16, env_name="stub-v0", T=30, num_cpu=2, paths=plain)
    assert np.array_equal(pol.get_param_values(), pol2.get_param_values())
    assert stats[:4] == stats2[:4]
    assert not np.array_equal(pol.get_param_values(), _fixture("rollout_stub_mlp")[0]["flat"].astype(np.float32))
    trajectory_sampler.shutdown()
