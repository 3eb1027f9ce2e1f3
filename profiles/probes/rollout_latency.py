"""Sampler overhead per simulator step: lock-step `mjrl_rollout_act` round trips vs
one `get_action` per environment step.  The environment is a no-op (fixed observation, never done), so
what is timed is everything BUT the simulator: staging, H2D, forward + noise + filing, D2H, bookkeeping.
The CPU column times the reference's way of acting (gaussian_mlp.py:92-98: a batch-1 torch fp32 forward
plus np.random.randn) with a plain torch module of the same shape on this box's host."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from mjrl_b200.policies.gaussian_mlp import MLP                  # noqa: E402
from mjrl_b200.samplers.lockstep import LockstepSampler          # noqa: E402
from mjrl_b200.utils.gym_env import EnvSpec                      # noqa: E402

N_OBS, N_ACT, T = 11, 3, 200


class NullEnv:
    horizon = T

    def __init__(self):
        self.env = self
        self.o = np.linspace(-1, 1, N_OBS)

    def seed(self, s):
        pass

    def reset(self):
        return self.o

    def step(self, a):
        return self.o, 0.0, False, {}


def main():
    pol = MLP(EnvSpec(N_OBS, N_ACT, T, 1), hidden_sizes=(64, 64), seed=0)
    rows = []
    for E in (16, 256, 4096):
        ls = LockstepSampler(env=NullEnv, num_slots=E, collect_env_infos=False)
        ls.sample(E, pol, T=20, pegasus_seed=0)                      # warm-up
        t0 = time.perf_counter()
        paths = ls.sample(E, pol, T=T, pegasus_seed=0)
        dt = time.perf_counter() - t0
        # the device round trip alone
        from mjrl_b200.samplers.lockstep import DeviceActor
        actor = DeviceActor(pol, ls._blk, ls._pinned, E)
        ls._blk.dst_rows[:] = np.arange(E)
        for _ in range(20):
            actor(True, False)
        t1 = time.perf_counter()
        for _ in range(200):
            actor(True, False)
        rt = (time.perf_counter() - t1) / 200
        rows.append(dict(slots=E, us_per_lockstep_step=1e6 * dt / T, env_steps_per_s=E * T / dt,
                         us_device_round_trip=1e6 * rt, timesteps=int(paths.batch.N),
                         h2d_bytes_after_rollout=int(paths.batch.h2d_bytes)))
        ls.close()
    # one get_action per step through the device (parity arm) and the reference's CPU way
    o = np.linspace(-1, 1, N_OBS)
    for _ in range(20):
        pol.get_action(o)
    t0 = time.perf_counter()
    for _ in range(300):
        pol.get_action(o)
    us_single = 1e6 * (time.perf_counter() - t0) / 300
    torch.set_num_threads(1)
    net = torch.nn.Sequential(torch.nn.Linear(N_OBS, 64), torch.nn.Tanh(), torch.nn.Linear(64, 64), torch.nn.Tanh(),
                              torch.nn.Linear(64, N_ACT))
    ls_val = np.zeros(N_ACT)

    def cpu_act(ob):
        x = torch.from_numpy(np.float32(ob.reshape(1, -1)))
        mean = net(x).data.numpy().ravel()
        return mean + np.exp(ls_val) * np.random.randn(N_ACT)
    for _ in range(100):
        cpu_act(o)
    t0 = time.perf_counter()
    for _ in range(2000):
        cpu_act(o)
    us_cpu = 1e6 * (time.perf_counter() - t0) / 2000
    print(json.dumps(dict(lockstep=rows, us_per_get_action_device=us_single,
                          us_per_get_action_cpu_torch_batch1=us_cpu,
                          cpu_env_steps_per_s_per_core=1e6 / us_cpu)))


if __name__ == "__main__":
    main()
