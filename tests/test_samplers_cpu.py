"""Host logic of the rollout path without a GPU: episode scheduling, the shared step
block, worker processes, path assembly.  The action computation is a stand-in here (the shipped
actor is `mjrl_rollout_act`, covered by tests/test_gpu_samplers.py)."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import stub_env                                                  # noqa: E402
from mjrl_b200.samplers import base_sampler                      # noqa: E402
from mjrl_b200.samplers.lockstep import LockstepSampler          # noqa: E402
from mjrl_b200.utils import tensor_utils                         # noqa: E402
from mjrl_b200.utils.get_environment import get_environment, register_environment, seed_environment  # noqa: E402


def _mean(W, obs):
    obs = np.asarray(obs, dtype=np.float32).reshape(-1, W.shape[1])
    return np.tanh((obs[:, None, :].astype(np.float64) * W[None]).sum(-1)).astype(np.float32)


class StubPolicy:
    """get_action with the reference's noise rule (gaussian_mlp.py:92-98)."""
    def __init__(self, n=5, m=2):
        self.n, self.m = n, m
        self.W = np.random.RandomState(7).randn(m, n) * 0.5
        self.log_std_val = np.full(m, -1.0)

    def get_action(self, o):
        mean = _mean(self.W, o)[0]
        noise = np.exp(self.log_std_val) * np.random.randn(self.m)
        return [mean + noise, dict(mean=mean, log_std=self.log_std_val, evaluation=mean)]


class StubActor:
    def __init__(self, policy, blk, pinned, rows):
        self.p, self.blk = policy, blk

    def __call__(self, sample, use_noise):
        blk = self.blk
        mean = _mean(self.p.W, blk.obs32)
        blk.mean[:] = mean
        if not sample:
            blk.act[:] = mean
            return
        if use_noise:
            eps = blk.noise.astype(np.float64)
        else:
            eps = np.stack([np.random.RandomState(int((int(s) * 1000003 + int(t)) % (2 ** 32))).randn(self.p.m)
                            for s, t in zip(blk.seeds, blk.tsteps)])
        blk.act[:] = (mean + np.exp(self.p.log_std_val) * eps).astype(np.float32)


def _same(a, b, exact=True):
    assert len(a) == len(b)
    for p, q in zip(a, b):
        assert p["terminated"] == q["terminated"]
        for k in ("observations", "actions", "rewards"):
            assert p[k].shape == q[k].shape, k
            if exact:
                assert np.array_equal(p[k], q[k]), k
            else:
                np.testing.assert_allclose(p[k], q[k], rtol=2e-6, atol=2e-6)


def test_stack_tensor_dict_list_nested():
    steps = [dict(a=1.0, b=dict(c=np.arange(2) + i, d=i)) for i in range(3)]
    out = tensor_utils.stack_tensor_dict_list(steps)
    assert out["a"].shape == (3,) and out["b"]["c"].shape == (3, 2)
    assert np.array_equal(out["b"]["d"], [0, 1, 2]) and np.array_equal(out["b"]["c"][2], [2, 3])


def test_registry_and_seeding_hooks():
    register_environment("stub-v0", stub_env.make_stub)
    env = get_environment("stub-v0")
    seed_environment(env, 5); a = env.reset()
    seed_environment(env, 5); b = env.reset()
    assert np.array_equal(a, b) and env.horizon == 30
    with pytest.raises((ImportError, Exception)):
        get_environment("not-registered-v0")


def test_lockstep_matches_serial_loop_with_reference_noise():
    """noise='reference' + the same policy arithmetic == base_sampler.do_rollout, value for value
    (fp32 actions in the lock-step store vs fp64 mean+noise in the serial loop: 2e-6)."""
    pol = StubPolicy()
    serial = base_sampler.do_rollout(9, pol, T=30, env=stub_env.make_stub(), pegasus_seed=100)
    lens = [len(p["rewards"]) for p in serial]
    assert min(lens) < 30 and max(lens) == 30, lens          # ragged: some episodes terminate early
    ls = LockstepSampler(env=stub_env.make_stub, num_slots=4, actor_factory=StubActor)
    batched = ls.sample(9, pol, T=30, pegasus_seed=100, noise="reference")
    _same(serial, batched, exact=False)
    assert [len(p["rewards"]) for p in batched] == lens
    p0, q0 = serial[0], batched[0]
    assert np.array_equal(p0["env_infos"]["t"], q0["env_infos"]["t"])
    assert q0["env_infos"]["state"]["x0"].shape == (lens[0],)
    assert q0["agent_infos"]["log_std"].shape == (lens[0], 2)
    np.testing.assert_allclose(p0["agent_infos"]["mean"], q0["agent_infos"]["mean"], rtol=2e-6, atol=2e-6)


def test_lockstep_evaluation_mode_matches_serial():
    pol = StubPolicy()
    serial = base_sampler.do_evaluation_rollout(5, pol, T=30, env=stub_env.make_stub(), pegasus_seed=3)
    batched = LockstepSampler(env=stub_env.make_stub, num_slots=5, actor_factory=StubActor).sample(
        5, pol, T=30, pegasus_seed=3, mode="evaluation")
    _same(serial, batched, exact=False)


def test_lockstep_batch_independent_of_slots_and_workers():
    pol = StubPolicy()
    runs = []
    for slots, workers in ((1, 0), (3, 0), (16, 0), (4, 2)):
        ls = LockstepSampler(env=stub_env.make_stub, num_slots=slots, num_workers=workers, actor_factory=StubActor,
                             worker_timeout=120)
        runs.append(ls.sample(11, pol, T=20, pegasus_seed=42))
        again = ls.sample(11, pol, T=20, pegasus_seed=42)       # a kept-alive sampler repeats itself
        _same(runs[-1], again)
        ls.close()
    for r in runs[1:]:
        _same(runs[0], r)
    other = LockstepSampler(env=stub_env.make_stub, num_slots=3, actor_factory=StubActor).sample(
        11, pol, T=20, pegasus_seed=43)
    assert not np.array_equal(other[0]["observations"], runs[0][0]["observations"])
    # episode_offset = a rank's share: episodes 4.. of the same rollout
    part = LockstepSampler(env=stub_env.make_stub, num_slots=3, actor_factory=StubActor).sample(
        7, pol, T=20, pegasus_seed=42, episode_offset=4)
    _same(runs[0][4:], part)


def test_lockstep_horizon_and_T_clamp():
    pol = StubPolicy()
    paths = LockstepSampler(env=stub_env.make_stub, num_slots=2, actor_factory=StubActor).sample(
        3, pol, T=7, pegasus_seed=1)
    assert all(len(p["rewards"]) <= 7 for p in paths)
    paths = LockstepSampler(env=stub_env.make_stub, num_slots=2, actor_factory=StubActor).sample(
        3, pol, T=1e6, pegasus_seed=1)
    assert max(len(p["rewards"]) for p in paths) == 30           # env.horizon wins (base_sampler.py:35)


def test_worker_failure_is_reported():
    ls = LockstepSampler(env=_Boom(), num_slots=2, num_workers=1, actor_factory=StubActor, worker_timeout=60)
    with pytest.raises(RuntimeError):
        ls.sample(2, StubPolicy(), T=5, pegasus_seed=0)


class _Boom:
    def __call__(self):
        raise RuntimeError("no simulator here")
