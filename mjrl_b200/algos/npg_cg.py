"""Drop-in for mjrl/algos/npg_cg.py on the CUDA path: analytic Fisher-vector product and a
conjugate-gradient loop that never leaves the device."""
import time as timer

import numpy as np
import torch

from ..utils.logger import DataLog
from .. import _lib
from .batch_reinforce import BatchREINFORCE, _dev32


class NPG(BatchREINFORCE):
    def __init__(self, env, policy, baseline, normalized_step_size=0.01, const_learn_rate=None,
                 FIM_invert_args={'iters': 10, 'damping': 1e-4}, hvp_sample_frac=1.0, seed=None,
                 save_logs=False, kl_dist=None):
        self.env = env
        self.policy = policy
        self.baseline = baseline
        self.alpha = const_learn_rate
        self.n_step_size = normalized_step_size if kl_dist is None else 2.0 * kl_dist
        self.seed = seed
        self.save_logs = save_logs
        self.FIM_invert_args = FIM_invert_args
        self.hvp_subsample = hvp_sample_frac
        self.running_score = None
        if save_logs: self.logger = DataLog()

    # ---- reference-compatible numerics API ----------------------------------------------------
    def HVP(self, observations, actions, vector, regu_coef=None):
        """npg_cg.py:55-74.  Valid where the reference calls it: new == old parameters."""
        regu_coef = self.FIM_invert_args['damping'] if regu_coef is None else regu_coef
        eng, obs, act, _ = self._arrays(observations, actions)
        if self.hvp_subsample is not None and self.hvp_subsample < 0.99:
            N = obs.shape[0]
            idx = np.random.choice(N, size=int(self.hvp_subsample * N))
            obs = obs[torch.from_numpy(idx).to(eng.device)]
        params = eng.dv.params("new")
        eng.cg_flags.zero_()
        out = torch.empty(eng.dv.d, dtype=torch.float64, device=eng.device)
        eng.fvp(params, obs, _dev32(vector, eng.device), out)
        eng.dist.allreduce(out)
        hv = (out / eng.dist.global_count(obs.shape[0], eng.device)).float().cpu().numpy()
        return hv + regu_coef * vector

    def build_Hvp_eval(self, inputs, regu_coef=None):
        def eval(v):
            return self.HVP(*(inputs + [v] + [regu_coef]))
        return eval

    # ---- shared body: gradient -> CG -> step size; returns everything still on device ----------
    def _npg_direction(self, eng, batch, w, g_scale_extra=1.0, demo=None, trace=None):
        d, m = eng.dv.d, eng.dv.m
        N_glob = self._global_n(eng, batch)
        p_old, curr = eng.dv.params("old"), eng.dv.params("new")
        old_ll = torch.empty(batch.N, dtype=torch.float32, device=eng.device)
        old_mean = torch.empty((batch.N, m), dtype=torch.float32, device=eng.device)
        # host-side: both vectors come from the host copy of the policy (no device reduce, no sync)
        same = np.array_equal(self.policy.get_param_values(), self.policy.get_old_param_values())
        ts = timer.time()
        if demo is None:
            if same:
                eng.gradient(curr, batch.obs, batch.act, w, None, cache=(old_ll, old_mean))
            else:
                old_ll, old_mean = eng.old_cache(p_old, batch.obs, batch.act)
                eng.gradient(curr, batch.obs, batch.act, w, old_ll)
        else:
            # DAPG: gradient weights differ from the surrogate weights (dapg.py:70,93,98)
            w_grad, d_obs, d_act, d_w = demo
            old_ll, old_mean = eng.old_cache(p_old, batch.obs, batch.act)
            eng.gradient(curr, batch.obs, batch.act, w_grad, None if same else old_ll)
            if d_obs is not None:
                d_old = None if same else eng.old_cache(p_old, d_obs, d_act)[0]
                eng.gradient(curr, d_obs, d_act, d_w, d_old, accumulate=True)
        eng.dist.allreduce(eng.grad)
        t_gLL = timer.time() - ts
        ts = timer.time()
        g_scale = g_scale_extra / N_glob
        subsample = None
        if self.hvp_subsample is not None and self.hvp_subsample < 0.99:
            def subsample():
                idx = np.random.choice(batch.N, size=int(self.hvp_subsample * batch.N))
                return batch.obs[torch.from_numpy(idx).to(eng.device)]
        eng.cg_solve(curr, batch.obs, g_scale, N_glob, self.FIM_invert_args['iters'],
                     self.FIM_invert_args['damping'], trace=trace, subsample=subsample)
        t_FIM = timer.time() - ts
        return dict(same=same, curr=curr, p_old=p_old, old_ll=old_ll, old_mean=old_mean, g_scale=g_scale,
                    N_glob=N_glob, t_gLL=t_gLL, t_FIM=t_FIM)

    def _commit(self, eng, st, new_dev, out, paths, stats, extra_log=None, surr_before_dev=None):
        d = eng.dv.d
        if out is None:
            # the sums of evaluate(raw=True) sit behind the gradient next to the step scalars: ONE contiguous read,
            # no device-side concatenation or division (the means are formed here, in the same fp64 arithmetic)
            t = _lib.to_host(eng.tail())                       # the update's single sync point
            host = np.concatenate([t[1:3] / st["N_glob"], t[0:1] / st["N_glob"], t[3:7]])
        else:
            sb = eng.grad[d:d + 1] / st["N_glob"] if surr_before_dev is None else surr_before_dev
            host = _lib.to_host(torch.cat([out, sb, eng.scal4]))   # the update's single sync point
        new_params = _lib.to_host(new_dev)
        eng.dist.check()     # multi-GPU: a peer exchange that timed out invalidates the update
        surr_after, kl_dist, surr_before = float(host[0]), float(host[1]), float(host[2])
        alpha, n_step_size = float(host[3]), float(host[5])
        self.policy.set_param_values(new_params, set_new=True, set_old=True)
        base_stats = self._finish_stats(stats, paths)
        if self.save_logs:
            self.logger.log_kv('alpha', alpha if extra_log is None else extra_log.get('alpha', alpha))
            self.logger.log_kv('delta', n_step_size)
            self.logger.log_kv('time_vpg', st["t_gLL"])
            self.logger.log_kv('time_npg', st["t_FIM"])
            self.logger.log_kv('kl_dist', kl_dist)
            self.logger.log_kv('surr_improvement', surr_after - surr_before)
            self.logger.log_kv('running_score', self.running_score)
            self._log_success(paths)
        self.last_update = dict(surr_before=surr_before, surr_after=surr_after, kl_dist=kl_dist,
                                alpha=alpha, delta=n_step_size, gdotx=float(host[4]))
        if extra_log:
            self.last_update.update(extra_log)
        return base_stats

    def train_from_paths(self, paths, cg_trace=None):
        """npg_cg.py:84-165."""
        eng, batch, w, stats = self._prepare(paths)
        st = self._npg_direction(eng, batch, w, trace=cg_trace)
        new_dev = eng.step(st["curr"], st["g_scale"], self.n_step_size, const_alpha=self.alpha)
        eng.evaluate(new_dev, batch.obs, batch.act, w, st["old_ll"], st["old_mean"],
                     st["p_old"][eng.dv.d - eng.dv.m:], st["N_glob"], raw=True)
        return self._commit(eng, st, new_dev, None, paths, stats)
