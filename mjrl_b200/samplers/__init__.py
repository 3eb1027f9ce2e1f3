"""Rollout side of the path (north_star (e)).

    base_sampler        the reference's single-environment loop, value-for-value (parity arm)
    lockstep            E slots per step on the GPU, paths filed straight into the device batch
    trajectory_sampler  sample_paths / sample_paths_parallel     (reference signatures)
    batch_sampler       sample until N timesteps                 (reference signatures)
"""


def sample_paths(N, policy, T, env_name, seed, num_cpu, sample_mode):
    """What `train_step` calls (batch_reinforce.py:74-79)."""
    if sample_mode == 'trajectories':
        from . import trajectory_sampler
        return trajectory_sampler.sample_paths_parallel(N, policy, T, env_name, seed, num_cpu)
    from . import batch_sampler
    return batch_sampler.sample_paths(N, policy, T, env_name=env_name, pegasus_seed=seed, num_cpu=num_cpu)
