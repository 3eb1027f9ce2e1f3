"""gym-free EnvSpec (the constructor argument of every policy and baseline).

Same four fields as mjrl/utils/gym_env.py:4-9; the reference module cannot be imported
without gym, which is why this one exists."""


class EnvSpec(object):
    def __init__(self, obs_dim, act_dim, horizon, num_agents=1):
        self.observation_dim = obs_dim
        self.action_dim = act_dim
        self.horizon = horizon
        self.num_agents = num_agents
