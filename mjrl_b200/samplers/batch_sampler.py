"""Sample until N TIMESTEPS are gathered (mjrl/samplers/batch_sampler.py:11-103)."""
import time as timer

import numpy as np

from . import base_sampler, trajectory_sampler


def sample_paths(N, policy, T=1e6, env=None, env_name=None, pegasus_seed=None, num_cpu='max',
                 paths_per_call=5, mode='sample'):
    if num_cpu == 1:
        return sample_paths_one_core(N, policy, T, env, env_name, pegasus_seed, mode)
    start_time = timer.time()
    print("####### Gathering Samples #######")
    sampled_so_far, paths_so_far, paths = 0, 0, []
    while sampled_so_far <= N:                 # :44 (<=, unlike the one-core loop)
        if pegasus_seed is not None:
            pegasus_seed += paths_so_far        # :50: cumulative, as the reference advances it
        new_paths = trajectory_sampler.sample_paths_parallel(paths_per_call, policy, T, env_name, pegasus_seed,
                                                             num_cpu, suppress_print=True, mode=mode)
        paths.extend(new_paths)
        paths_so_far += paths_per_call
        sampled_so_far += int(np.sum([len(p['rewards']) for p in new_paths]))
    print("======= Samples Gathered  ======= | >>>> Time taken = %f " % (timer.time() - start_time))
    print("................................. | >>>> # samples = %i # trajectories = %i " % (sampled_so_far, paths_so_far))
    return paths


def sample_paths_one_core(N, policy, T=1e6, env=None, env_name=None, pegasus_seed=None, mode='sample'):
    """:63-103: one environment, one rollout at a time, seeds pegasus_seed, +1, +2, ..."""
    if env is None:
        if env_name is None:
            raise ValueError("one of env / env_name must be given")
        from ..utils.get_environment import get_environment
        env = get_environment(env_name)
    start_time = timer.time()
    print("####### Gathering Samples #######")
    sampled_so_far, paths = 0, []
    seed = pegasus_seed if pegasus_seed is not None else 0
    while sampled_so_far < N:
        if mode == 'sample':
            this_path = base_sampler.do_rollout(1, policy, T, env, env_name, seed)
        elif mode == 'evaluation':
            this_path = base_sampler.do_evaluation_rollout(1, policy, T, env, env_name, seed)
        else:
            raise ValueError("Mode has to be either 'sample' for training time or 'evaluation' for test time performance")
        paths.append(this_path[0])
        seed += 1
        sampled_so_far += len(this_path[0]["rewards"])
    print("======= Samples Gathered  ======= | >>>> Time taken = %f " % (timer.time() - start_time))
    print("................................. | >>>> # samples = %i # trajectories = %i " % (sampled_so_far, len(paths)))
    return paths
