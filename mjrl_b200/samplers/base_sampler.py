"""Single-environment rollouts with the reference's exact semantics
(mjrl/samplers/base_sampler.py:10-92, evaluation_sampler.py:11-89).

This is the parity arm of the rollout path: one environment, one `policy.get_action` per step,
exploration noise from numpy's global MT19937 stream re-seeded per episode with
`pegasus_seed + ep`.  The throughput arm is `lockstep.LockstepSampler`, which steps every slot
with one launch per simulator step.
"""
import numpy as np

from ..utils import tensor_utils
from ..utils.get_environment import get_environment, seed_environment


def _rollout(N, policy, T, env, env_name, pegasus_seed, evaluation):
    if env is None:
        if env_name is None:
            raise ValueError("one of env / env_name must be given")
        env = get_environment(env_name)
    if pegasus_seed is not None:
        seed_environment(env, pegasus_seed)
    T = min(T, env.horizon)
    paths = []
    for ep in range(N):
        if pegasus_seed is not None:
            seed = pegasus_seed + ep
            seed_environment(env, seed)
            np.random.seed(seed)
        else:
            np.random.seed()
        obs, acts, rews, agent_infos, env_infos = [], [], [], [], []
        o = env.reset()
        done, t = False, 0
        while t < T and done != True:       # noqa: E712  (the reference's comparison: done may be numpy.bool_)
            a, agent_info = policy.get_action(o)
            if evaluation:
                a = agent_info['evaluation']
            next_o, r, done, env_info = env.step(a)
            obs.append(o); acts.append(a); rews.append(r)
            agent_infos.append(agent_info); env_infos.append(env_info)
            o = next_o
            t += 1
        paths.append(dict(observations=np.array(obs), actions=np.array(acts), rewards=np.array(rews),
                          agent_infos=tensor_utils.stack_tensor_dict_list(agent_infos),
                          env_infos=tensor_utils.stack_tensor_dict_list(env_infos),
                          terminated=done))
    return paths


def do_rollout(N, policy, T=1e6, env=None, env_name=None, pegasus_seed=None):
    """base_sampler.py:10-88."""
    return _rollout(N, policy, T, env, env_name, pegasus_seed, evaluation=False)


def do_rollout_star(args_list):
    return do_rollout(*args_list)


def do_evaluation_rollout(N, policy, T=1e6, env=None, env_name=None, pegasus_seed=None):
    """evaluation_sampler.py:11-85: same loop, acting with agent_info['evaluation'] (the mean)."""
    return _rollout(N, policy, T, env, env_name, pegasus_seed, evaluation=True)


def do_evaluation_rollout_star(args_list):
    return do_evaluation_rollout(*args_list)
