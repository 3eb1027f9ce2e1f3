"""A tiny numpy environment with the GymEnv surface the samplers use (mjrl/utils/gym_env.py:11-62):
rotating linear dynamics, its own seeded RandomState, early termination for some start states.
Shared by the sampler tests and by tests/golden/make_golden.py (rollout fixture)."""
import numpy as np


class _Spec:
    def __init__(self, n, m, horizon):
        self.observation_dim, self.action_dim, self.horizon, self.num_agents = n, m, horizon, 1


class StubEnv:
    def __init__(self, n=5, m=2, horizon=30):
        rs = np.random.RandomState(20240229)
        q, _ = np.linalg.qr(rs.randn(n, n))
        self.A = 0.97 * q
        self.B = 0.3 * rs.randn(n, m)
        self.n, self.m = n, m
        self.horizon = horizon
        self.env = self                 # the `env.env` level the reference seeds through
        self.env_id = "stub-v0"
        self.spec = _Spec(n, m, horizon)
        self.rng = np.random.RandomState(0)
        self.x, self.t = np.zeros(n), 0

    def seed(self, seed):
        self.rng = np.random.RandomState(int(seed) % (2 ** 32))

    _seed = seed        # old-gym spelling; the reference's evaluation sampler only knows this one

    def reset(self):
        self.x = self.rng.uniform(-1.0, 1.0, self.n)
        self.t = 0
        return self.x.copy()

    def step(self, a):
        a = np.clip(np.asarray(a, dtype=np.float64).ravel(), -1.0, 1.0)
        self.x = self.A @ self.x + self.B @ a + 0.01 * self.rng.randn(self.n)
        self.t += 1
        r = float(-np.sum(self.x ** 2) - 0.01 * np.sum(a ** 2))
        done = bool(self.x[0] > 0.8)
        return self.x.copy(), r, done, dict(t=self.t, state=dict(x0=self.x[0], big=done))


def make_stub():
    return StubEnv()


def make_stub_long():
    return StubEnv(n=11, m=3, horizon=100)
