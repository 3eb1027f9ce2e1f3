"""Drop-in for mjrl/algos/batch_reinforce.py on the CUDA path.

Constructor, `train_step`, `train_from_paths`, `CPI_surrogate`, `kl_old_new`, `flat_vpg` and the
logged keys follow the reference; the numerics run through algos/engine.py (no autograd).
"""
import time as timer

import numpy as np
import torch

from .. import _lib
from ..batch import PathBatch, batch_for
from ..utils import process_samples
from ..utils.logger import DataLog
from .engine import Dist, UpdateEngine


def _dev32(x, device):
    return torch.from_numpy(np.ascontiguousarray(np.asarray(x, dtype=np.float32))).to(device)


class BatchREINFORCE:
    def __init__(self, env, policy, baseline, learn_rate=0.01, seed=None, save_logs=False):
        self.env = env
        self.policy = policy
        self.baseline = baseline
        self.alpha = learn_rate
        self.seed = seed
        self.save_logs = save_logs
        self.running_score = None
        if save_logs: self.logger = DataLog()

    # ---- engine plumbing ------------------------------------------------------------------
    def _engine(self):
        eng = getattr(self, "_eng", None)
        if eng is None or eng.policy is not self.policy:
            eng = self._eng = UpdateEngine(self.policy, getattr(self, "dist", None))
        return eng

    def __getstate__(self):
        d = dict(self.__dict__)
        d.pop("_eng", None)
        return d

    def _arrays(self, observations, actions, advantages=None):
        eng = self._engine()
        obs = _dev32(observations, eng.device).reshape(-1, self.policy.n)
        act = _dev32(actions, eng.device).reshape(-1, self.policy.m)
        w = _dev32(advantages, eng.device) if advantages is not None else None
        return eng, obs, act, w

    def _surr_kl(self, observations, actions, advantages):
        eng, obs, act, w = self._arrays(observations, actions, advantages)
        p_old, p_new = eng.dv.params("old"), eng.dv.params("new")
        old_ll, old_mean = eng.old_cache(p_old, obs, act)
        N = obs.shape[0]
        out = eng.evaluate(p_new, obs, act, w, old_ll, old_mean, p_old[eng.dv.d - eng.dv.m:], float(N))
        return out.cpu()

    # ---- reference-compatible numerics API (numpy in) --------------------------------------
    def CPI_surrogate(self, observations, actions, advantages):
        """mean(LR * adv), batch_reinforce.py:37-43.  Returns a 0-dim fp32 torch tensor."""
        return self._surr_kl(observations, actions, advantages)[0].float()

    def kl_old_new(self, observations, actions):
        """batch_reinforce.py:45-49."""
        return self._surr_kl(observations, actions, None)[1].float()

    def flat_vpg(self, observations, actions, advantages):
        """batch_reinforce.py:51-55 -> fp32 numpy [d]."""
        eng, obs, act, w = self._arrays(observations, actions, advantages)
        p_old, p_new = eng.dv.params("old"), eng.dv.params("new")
        old_ll = None
        if not torch.equal(p_old, p_new):
            old_ll, _ = eng.old_cache(p_old, obs, act)
        g = eng.gradient(p_new, obs, act, w, old_ll)
        eng.dist.allreduce(g)
        return (g[:eng.dv.d] / eng.dist.global_count(obs.shape[0], eng.device)).float().cpu().numpy()

    # ---- one iteration -------------------------------------------------------------------
    def train_step(self, N, sample_mode='trajectories', env_name=None, T=1e6, gamma=0.995,
                   gae_lambda=0.98, num_cpu='max', paths=None):
        """batch_reinforce.py:58-103.  `paths` may be supplied directly (the sampler needs gym +
        MuJoCo, which stay on host worker processes)."""
        if sample_mode != 'trajectories' and sample_mode != 'samples':
            raise ValueError("sample_mode must be either 'trajectories' or 'samples'")
        ts = timer.time()
        if paths is None:
            from ..samplers import sample_paths
            if env_name is None: env_name = self.env.env_id
            paths = sample_paths(N, self.policy, T, env_name, self.seed, num_cpu, sample_mode)
        if self.save_logs:
            self.logger.log_kv('time_sampling', timer.time() - ts)
        self.seed = self.seed + N if self.seed is not None else self.seed

        batch = paths if isinstance(paths, PathBatch) else batch_for(paths)
        process_samples.returns_on_device(batch, gamma)
        b = process_samples.baseline_values_on_device(batch, self.baseline)
        process_samples.advantages_on_device(batch, b, gamma, gae_lambda)
        eval_statistics = self.train_from_paths(batch)
        eval_statistics.append(N)
        if self.save_logs:
            ts = timer.time()
            error_before, error_after = self._fit_baseline(batch, paths, True)
            self.logger.log_kv('time_VF', timer.time() - ts)
            self.logger.log_kv('VF_error_before', error_before)
            self.logger.log_kv('VF_error_after', error_after)
        else:
            self._fit_baseline(batch, paths, False)
        return eval_statistics

    def _fit_baseline(self, batch, paths, return_errors):
        if hasattr(self.baseline, "fit_batch"):
            dist = self._engine().dist
            return self.baseline.fit_batch(batch, return_errors=return_errors,
                                           dist=dist if dist.world > 1 else None)
        batch.scatter("returns", batch.returns)
        return self.baseline.fit(batch.paths, return_errors=return_errors)

    # ---- shared front half of every train_from_paths -----------------------------------------
    def _prepare(self, paths):
        """Device batch with advantages, whitened fp32 weights, and the path-return stats
        (npg_cg.py:87-104)."""
        eng = self._engine()
        batch = paths if isinstance(paths, PathBatch) else batch_for(paths)
        batch.adv = batch.current("advantages", batch.adv)
        w = eng.whiten(batch.adv, eps=1e-6)
        stats = torch.empty(4, dtype=torch.float64, device=eng.device)
        eng.ctx.call("mjrl_path_return_stats", _ptr(batch.rew), _ptr(batch.offsets),
                     _c_int(batch.P), _ptr(stats), _stream())
        eng.launches += 2
        return eng, batch, w, stats

    @staticmethod
    def _global_n(eng, batch):
        """Global timestep count of a sharded batch (cached on the batch object)."""
        n = getattr(batch, "_n_global", None)
        if n is None:
            n = batch._n_global = eng.dist.global_count(batch.N, eng.device)
        return n

    def _finish_stats(self, stats, paths):
        base_stats = [float(v) for v in _lib.to_host(stats)]
        mean_return = base_stats[0]
        self.running_score = mean_return if self.running_score is None else \
            0.9 * self.running_score + 0.1 * mean_return
        if self.save_logs: self.log_rollout_statistics(base_stats)
        return base_stats

    def train_from_paths(self, paths):
        """batch_reinforce.py:106-164: new = curr + learn_rate * vpg_grad."""
        eng, batch, w, stats = self._prepare(paths)
        N_glob = self._global_n(eng, batch)
        d, m = eng.dv.d, eng.dv.m
        ts = timer.time()
        p_old, curr = eng.dv.params("old"), eng.dv.params("new")
        old_ll = torch.empty(batch.N, dtype=torch.float32, device=eng.device)
        old_mean = torch.empty((batch.N, m), dtype=torch.float32, device=eng.device)
        if np.array_equal(self.policy.get_param_values(), self.policy.get_old_param_values()):
            eng.gradient(curr, batch.obs, batch.act, w, None, cache=(old_ll, old_mean))
        else:
            old_ll, old_mean = eng.old_cache(p_old, batch.obs, batch.act)
            eng.gradient(curr, batch.obs, batch.act, w, old_ll)
        eng.dist.allreduce(eng.grad)
        # x := g, then the generic step kernel with a constant alpha
        eng.cg_x.copy_(eng.grad[:d] / N_glob)
        new_dev = eng.step(curr, 1.0 / N_glob, 0.0, const_alpha=float(self.alpha))
        out = eng.evaluate(new_dev, batch.obs, batch.act, w, old_ll, old_mean, p_old[d - m:], N_glob)
        host = _lib.to_host(torch.cat([out, eng.grad[d:d + 1] / N_glob]))
        t_gLL = timer.time() - ts
        new_params = _lib.to_host(new_dev)
        eng.dist.check()     # multi-GPU: a peer exchange that timed out invalidates the update
        surr_after, kl_dist, surr_before = float(host[0]), float(host[1]), float(host[2])
        self.policy.set_param_values(new_params, set_new=True, set_old=True)
        base_stats = self._finish_stats(stats, paths)
        if self.save_logs:
            self.logger.log_kv('alpha', self.alpha)
            self.logger.log_kv('time_vpg', t_gLL)
            self.logger.log_kv('kl_dist', kl_dist)
            self.logger.log_kv('surr_improvement', surr_after - surr_before)
            self.logger.log_kv('running_score', self.running_score)
            self._log_success(paths)
        self.last_update = dict(surr_before=surr_before, surr_after=surr_after, kl_dist=kl_dist,
                                alpha=self.alpha)
        return base_stats

    def _log_success(self, paths):
        if isinstance(paths, PathBatch): paths = paths.paths
        try:
            self.env.env.env.evaluate_success(paths, self.logger)
        except Exception:
            try:
                self.logger.log_kv('success_rate', self.env.env.env.evaluate_success(paths))
            except Exception:
                pass

    def log_rollout_statistics(self, stats_or_paths):
        if isinstance(stats_or_paths, list) and stats_or_paths and isinstance(stats_or_paths[0], dict):
            R = [sum(p["rewards"]) for p in stats_or_paths]
            stats_or_paths = [np.mean(R), np.std(R), np.amin(R), np.amax(R)]
        mean_return, std_return, min_return, max_return = stats_or_paths
        self.logger.log_kv('stoc_pol_mean', mean_return)
        self.logger.log_kv('stoc_pol_std', std_return)
        self.logger.log_kv('stoc_pol_max', max_return)
        self.logger.log_kv('stoc_pol_min', min_return)


# small aliases so the body above reads like the C-ABI call it makes
from .._lib import c_int as _c_int, ptr as _ptr, stream_ptr as _stream  # noqa: E402
