"""Drop-in for mjrl/policies/gaussian_linear.py (LinearPolicy + LinearModel): the
zero-hidden-layer case of the same CUDA passes."""
import numpy as np
import torch
import torch.nn as nn

from .gaussian_mlp import _GaussianPolicyBase


class LinearModel(nn.Module):
    """mean = fc0((x - in_shift)/(in_scale + 1e-8)) * out_scale + out_shift
    (mjrl/policies/gaussian_linear.py:141-175)."""

    def __init__(self, obs_dim, act_dim, in_shift=None, in_scale=None, out_shift=None,
                 out_scale=None):
        super().__init__()
        self.obs_dim, self.act_dim = obs_dim, act_dim
        self.set_transformations(in_shift, in_scale, out_shift, out_scale)
        self.fc0 = nn.Linear(obs_dim, act_dim)

    def set_transformations(self, in_shift=None, in_scale=None, out_shift=None, out_scale=None):
        self.transformations = dict(in_shift=in_shift, in_scale=in_scale, out_shift=out_shift,
                                    out_scale=out_scale)
        as_t = lambda v, fill, k: (torch.from_numpy(np.float32(v)) if v is not None
                                   else fill(k))
        self.in_shift = as_t(in_shift, torch.zeros, self.obs_dim)
        self.in_scale = as_t(in_scale, torch.ones, self.obs_dim)
        self.out_shift = as_t(out_shift, torch.zeros, self.act_dim)
        self.out_scale = as_t(out_scale, torch.ones, self.act_dim)

    def forward(self, x):
        h = (x - self.in_shift) / (self.in_scale + 1e-8)
        return self.fc0(h) * self.out_scale + self.out_shift


class LinearPolicy(_GaussianPolicyBase):
    def __init__(self, env_spec, min_log_std=-3, init_log_std=0, seed=None):
        self.n = env_spec.observation_dim
        self.m = env_spec.action_dim
        self.min_log_std = min_log_std
        if seed is not None:
            torch.manual_seed(seed)
            np.random.seed(seed)
        self.model = self._make_model()
        self._finish_init(init_log_std)

    def _make_model(self):
        return LinearModel(self.n, self.m)
