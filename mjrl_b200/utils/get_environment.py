"""Environment lookup for the samplers (mjrl/utils/get_environment.py:8-10).

The reference resolves `env_name` with `gym.make` inside every worker.  Simulators stay host-side
and out of scope here, so the lookup is a registry: a name maps to a picklable
zero-argument factory returning an object with the `GymEnv` surface the samplers use
(mjrl/utils/gym_env.py:11-62): `reset()`, `step(a) -> (obs, reward, done, info)`, `horizon`,
`env_id`, `spec`, and a seeding hook reachable as `env.env._seed`, `env.env.seed` or `env.seed`
(base_sampler.py:30-34).  Names that are not registered fall through to gym when it is installed.
"""
from .gym_env import EnvSpec

_REGISTRY = {}


def register_environment(env_name, factory):
    """`factory()` must build a fresh environment; it is called once per slot (and inside worker
    processes, so it has to be picklable: a module-level function or class)."""
    _REGISTRY[env_name] = factory


def registered(env_name):
    return env_name in _REGISTRY


def factory_for(env_name):
    if env_name in _REGISTRY:
        return _REGISTRY[env_name]
    return _GymFactory(env_name)


def get_environment(env_name=None):
    if env_name is None:
        raise ValueError("Need to specify environment name")
    return factory_for(env_name)()


class _GymFactory:
    def __init__(self, env_name):
        self.env_name = env_name

    def __call__(self):
        return GymEnv(self.env_name)


class GymEnv(object):
    """gym-backed environment with the attribute layout of mjrl/utils/gym_env.py:11-62."""

    def __init__(self, env_name):
        try:
            import gym
        except ImportError as e:
            raise ImportError("environment %r is not registered with mjrl_b200.utils.get_environment."
                              "register_environment and gym is not installed" % (env_name,)) from e
        env = gym.make(env_name)
        self.env = env
        self.env_id = env.spec.id
        self._horizon = getattr(env.spec, "timestep_limit", None) or env.spec.max_episode_steps
        inner = getattr(env, "env", env)
        self._action_dim = getattr(inner, "action_dim", None) or inner.action_space.shape[0]
        self._observation_dim = getattr(inner, "obs_dim", None) or inner.observation_space.shape[0]
        self._num_agents = getattr(inner, "num_agents", 1)
        self.spec = EnvSpec(self._observation_dim, self._action_dim, self._horizon, self._num_agents)

    action_dim = property(lambda self: self._action_dim)
    observation_dim = property(lambda self: self._observation_dim)
    horizon = property(lambda self: self._horizon)

    def reset(self):
        return self.env.reset()

    def step(self, action):
        return self.env.step(action)


def seed_environment(env, seed):
    """base_sampler.py:30-34 / :46-49: `env.env._seed(seed)`, else `env.env.seed(seed)`; bare
    environments without the wrapper level are accepted too."""
    for holder in (getattr(env, "env", None), env):
        for name in ("_seed", "seed"):
            fn = getattr(holder, name, None) if holder is not None else None
            if callable(fn):
                fn(seed)
                return
    raise AttributeError("environment has no seeding hook (env.env._seed / env.env.seed / env.seed)")
