"""Drop-in for mjrl/algos/behavior_cloning_2.py on the CUDA path: behaviour cloning with an MSE
loss on the mean network (behavior_cloning_2.py:50-54), Adam over `policy.model.parameters()` only
(:43), and the policy's log_std set once from the action scale (:36-39).

The MSE loss is the Gaussian log-likelihood with log_std = 0 up to constants:
    mean_{t,j} (mu_tj - a_tj)^2  =  -(2 / (B m)) * sum_t [ LL_t + (m/2) log(2 pi) ]   at sigma = 1,
so the minibatch gradient is the policy gradient pass run on a working vector whose log_std block is
zero, scaled by -2/(B m) and applied to the first d - m entries; see algos/behavior_cloning.py for
the chain itself.
"""
import numpy as np
import torch

from .behavior_cloning import BC as _BC


class BC(_BC):
    mode = "mse"

    def _init_policy(self, out_scale):
        # set the variance of gaussian policy based on out_scale (:36-39); set_param_values' defaults
        # set new AND old from a float32 vector, so the reference's new/old aliasing applies (F14)
        params = self.policy.get_param_values()
        params[-self.policy.m:] = np.log(out_scale + 1e-12)
        self.policy.set_param_values(params)

    def _work_params(self, params):
        w = params.clone()
        w[-self.policy.m:] = 0.0
        return w

    def _d_opt(self, dv):
        return dv.d - dv.m

    def _grad_scale(self, dv):
        return -2.0 / (self.mb_size * dv.m)

    def _loss_dev(self, dv, params, obs, act):
        mean, _ = dv.forward(self._work_params(params), obs, None, want_mean=True, want_ll=False)
        return torch.mean((mean - act) ** 2)     # torch.nn.MSELoss over all N*m elements (:45,54)
