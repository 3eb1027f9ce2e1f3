"""`sample_paths` / `sample_paths_parallel` with the reference's signatures
(mjrl/samplers/trajectory_sampler.py:10-65).

`num_cpu == 1` is the reference's in-process loop (parity arm, base_sampler.py).  Anything else is the
lock-step sampler: `num_cpu` worker processes own the environments, the GPU samples the actions of
all slots per step, and the returned paths are already filed on the device (`paths.batch`).

Kept from the reference on purpose: the call returns `num_cpu * ceil(N / num_cpu)` trajectories
(:37, every worker is asked for `paths_per_cpu`), and trajectory k uses seed `pegasus_seed + k`
(:41-45 with base_sampler.py:44-46) -- so the same call yields the same number of paths from the
same environment seeds.
"""
import multiprocessing as mp
import time as timer

import numpy as np

from . import base_sampler
from .lockstep import LockstepSampler

_samplers = {}


def sample_paths(N, policy, T=1e6, env=None, env_name=None, pegasus_seed=None, mode='sample'):
    """trajectory_sampler.py:10-17."""
    if mode == 'sample':
        return base_sampler.do_rollout(N, policy, T, env, env_name, pegasus_seed)
    if mode == 'evaluation':
        return base_sampler.do_evaluation_rollout(N, policy, T, env, env_name, pegasus_seed)
    raise ValueError("Mode has to be either 'sample' for training time or 'evaluation' for test time performance")


def lockstep_sampler(env_name, num_workers, num_slots=None):
    """Process-wide sampler per (environment, worker count): the workers and their environments
    outlive the call, unlike the reference's per-call `mp.Pool(maxtasksperchild=1)` (:71)."""
    key = (env_name, int(num_workers), num_slots)
    if key not in _samplers:
        slots = num_slots if num_slots is not None else max(64, 8 * int(num_workers))
        _samplers[key] = LockstepSampler(env_name=env_name, num_slots=slots, num_workers=num_workers)
    return _samplers[key]


def shutdown():
    for s in _samplers.values():
        s.close()
    _samplers.clear()


def sample_paths_parallel(N, policy, T=1e6, env_name=None, pegasus_seed=None, num_cpu='max',
                          max_process_time=300, max_timeouts=4, suppress_print=False, mode='sample',
                          noise='philox', num_slots=None):
    if num_cpu is None or num_cpu == 'max':
        num_cpu = mp.cpu_count()
    elif num_cpu == 1:
        return base_sampler.do_rollout(N, policy, T, None, env_name, pegasus_seed)   # :32-33 (mode ignored there too)
    else:
        num_cpu = min(mp.cpu_count(), num_cpu)
    paths_per_cpu = int(np.ceil(N / num_cpu))
    total = paths_per_cpu * num_cpu
    if not suppress_print:
        start_time = timer.time()
        print("####### Gathering Samples #######")
    sampler = lockstep_sampler(env_name, num_cpu, num_slots)
    sampler.worker_timeout = max_process_time
    paths = None
    for attempt in range(max_timeouts):      # :80-88: a timed-out gather is retried, max_timeouts times in all
        try:
            paths = sampler.sample(total, policy, T, pegasus_seed, mode=mode, noise=noise)
            break
        except TimeoutError as e:
            print(str(e))
            print("Timeout Error raised... Trying again")
    if not suppress_print:
        print("======= Samples Gathered  ======= | >>>> Time taken = %f " % (timer.time() - start_time))
    return paths
