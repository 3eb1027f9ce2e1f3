"""Drop-in for mjrl/algos/ppo_clip.py on the CUDA path.

Same constructor and `train_from_paths` as the reference (ppo_clip.py:22-120).  The full-batch
evaluations (`CPI_surrogate`, `kl_old_new`, :82,99-100) are the fused evaluation pass; every
minibatch step (:87-96) runs on the device as
    gather rows -> forward (LL) -> clipped-surrogate weights -> policy gradient -> Adam
with the policy kernels of csrc/policy.cu and csrc/ppo.cu.  The chain of `epochs * int(N/mb)`
steps is serial like `MLPBaseline.fit`; the with-replacement minibatch indices come
from `np.random.choice` exactly like :89, so a seeded run visits the reference's rows.
`self.optimizer`'s role (Adam moments + step count, kept across iterations) is played by device
tensors that are dropped on pickle and re-created from host copies.
"""
import time as timer

import numpy as np
import torch

from .._lib import c_double, c_float, c_i64, c_int, ptr, stream_ptr
from ..utils.logger import DataLog
from .batch_reinforce import BatchREINFORCE


class PPO(BatchREINFORCE):
    def __init__(self, env, policy, baseline, clip_coef=0.2, epochs=10, mb_size=64, learn_rate=3e-4,
                 seed=0, save_logs=False):
        self.env = env
        self.policy = policy
        self.baseline = baseline
        self.learn_rate = learn_rate
        self.seed = seed
        self.save_logs = save_logs
        self.clip_coef = clip_coef
        self.epochs = epochs
        self.mb_size = mb_size
        self.running_score = None
        if save_logs: self.logger = DataLog()
        # torch.optim.Adam(self.policy.trainable_params, lr=learn_rate) state (ppo_clip.py:45)
        self.adam_step = 0
        self._adam_host = None      # (m, v) numpy copies for pickling
        self._adam_dev = None

    def __getstate__(self):
        d = super().__getstate__()
        if self._adam_dev is not None:
            d["_adam_host"] = tuple(t.cpu().numpy() for t in self._adam_dev)
        d["_adam_dev"] = None
        return d

    def _adam(self, eng):
        if self._adam_dev is None or self._adam_dev[0].device != eng.device:
            if self._adam_host is not None:
                self._adam_dev = tuple(torch.from_numpy(h).to(eng.device) for h in self._adam_host)
            else:
                z = lambda: torch.zeros(eng.dv.d, dtype=torch.float32, device=eng.device)
                self._adam_dev = (z(), z())
        return self._adam_dev

    def minibatch_indices(self, num_samples):
        """ppo_clip.py:87-89: one np.random.choice(num_samples, size=mb_size) per step."""
        steps = self.epochs * int(num_samples / self.mb_size)
        if steps == 0:
            return np.zeros((0, self.mb_size), dtype=np.int32)
        return np.stack([np.random.choice(num_samples, size=self.mb_size) for _ in range(steps)]).astype(np.int32)

    def train_from_paths(self, paths, minibatches=None):
        eng, batch, w, stats = self._prepare(paths)
        world = eng.dist.world
        d, m, n = eng.dv.d, eng.dv.m, eng.dv.n
        N = batch.N
        N_glob = self._global_n(eng, batch) if world > 1 else float(N)
        ctx, dev = eng.ctx, eng.device
        ts = timer.time()
        p_old = eng.dv.params("old")
        params = eng.dv.params("new").clone()
        # Reference quirk (see policies/gaussian_mlp.py set_param_values): once new/old alias, the
        # old NETWORK weights follow every in-place Adam step and only old_log_std stays fixed.
        aliased = bool(getattr(self.policy, "_old_aliases_new", False))
        p_old_eff = p_old.clone()
        old_ll, old_mean = eng.old_cache(p_old, batch.obs, batch.act)
        surr_before_dev = eng.evaluate(params, batch.obs, batch.act, w, old_ll, old_mean, p_old[d - m:],
                                       N_glob)[0:1].clone()
        # The minibatch chain samples the GLOBAL batch (ppo_clip.py:89) and is serial: with several ranks it
        # runs as REPLICAS -- every rank gathers all rows (rank order), rank 0 draws the minibatches, and
        # every rank runs the identical deterministic chain; the full-batch passes stay sharded.
        c_obs, c_act, c_w, c_oll, Nc = batch.obs, batch.act, w, old_ll, N
        if world > 1:
            packed, sizes = eng.dist.allgather_rows(
                torch.cat([batch.obs, batch.act, w[:, None], old_ll[:, None]], dim=1))
            Nc = int(sum(sizes))
            c_obs, c_act = packed[:, :n].contiguous(), packed[:, n:n + m].contiguous()
            c_w, c_oll = packed[:, n + m].contiguous(), packed[:, n + m + 1].contiguous()
        if minibatches is not None:
            idx = np.asarray(minibatches, dtype=np.int32)
            d_idx = torch.from_numpy(np.ascontiguousarray(idx)).to(dev)
        else:
            steps = self.epochs * int(Nc / self.mb_size)
            if world == 1 or eng.dist.rank == 0:
                idx = self.minibatch_indices(Nc)
                d_idx = torch.from_numpy(np.ascontiguousarray(idx)).to(dev)
            else:
                d_idx = torch.empty((steps, self.mb_size), dtype=torch.int32, device=dev)
            if world > 1 and steps > 0:
                eng.dist.broadcast0(d_idx)
            idx = d_idx      # only its shape is used below
        B = idx.shape[1] if idx.shape[0] else self.mb_size
        am, av = self._adam(eng)
        f32 = lambda *s: torch.empty(s, dtype=torch.float32, device=dev)
        mb_obs, mb_act, mb_adv, mb_oll, mb_ll, mb_w = f32(B, n), f32(B, m), f32(B), f32(B), f32(B), f32(B)
        gsum = torch.empty(d + 1, dtype=torch.float64, device=dev)
        st = stream_ptr()
        for s in range(idx.shape[0]):
            rows = ptr(d_idx[s])
            ctx.call("mjrl_gather_rows_f32", ptr(c_obs), rows, c_int(B), c_int(n), ptr(mb_obs), st)
            ctx.call("mjrl_gather_rows_f32", ptr(c_act), rows, c_int(B), c_int(m), ptr(mb_act), st)
            ctx.call("mjrl_gather_rows_f32", ptr(c_w), rows, c_int(B), c_int(1), ptr(mb_adv), st)
            if aliased:
                p_old_eff[:d - m].copy_(params[:d - m])
                ctx.call("mjrl_policy_forward", eng.dv.desc_ref(), ptr(p_old_eff), ptr(mb_obs), ptr(mb_act),
                         c_i64(B), ptr(None), ptr(mb_oll), st)
            else:
                ctx.call("mjrl_gather_rows_f32", ptr(c_oll), rows, c_int(B), c_int(1), ptr(mb_oll), st)
            ctx.call("mjrl_policy_forward", eng.dv.desc_ref(), ptr(params), ptr(mb_obs), ptr(mb_act),
                     c_i64(B), ptr(None), ptr(mb_ll), st)
            ctx.call("mjrl_ppo_clip_weights", ptr(mb_ll), ptr(mb_oll), ptr(mb_adv), c_int(B),
                     c_float(self.clip_coef), ptr(mb_w), ptr(None), st)
            ctx.call("mjrl_policy_grad", eng.dv.desc_ref(), ptr(params), ptr(mb_obs), ptr(mb_act), ptr(mb_w),
                     ptr(mb_oll), c_i64(B), ptr(gsum), ptr(gsum[d:]), ptr(None), ptr(None), st)
            self.adam_step += 1
            # loss = -PPO_surrogate (ppo_clip.py:94): gradient = -(1/B) sum_t w_t LR_t dLL_t
            ctx.call("mjrl_adam_step", c_i64(d), ptr(params), ptr(gsum), c_double(-1.0 / B), ptr(am), ptr(av),
                     c_i64(self.adam_step), c_double(self.learn_rate), c_double(0.9), c_double(0.999),
                     c_double(1e-8), c_double(0.0), st)
        if aliased:
            p_old_eff[:d - m].copy_(params[:d - m])
            old_ll, old_mean = eng.old_cache(p_old_eff, batch.obs, batch.act)
        # ppo_clip.py:100-103: set_param_values(set_new=True) clamps log_std at min_log_std BEFORE surr_after and the
        # KL are computed; Adam can push log_std below the floor
        ctx.call("mjrl_policy_clamp_log_std", eng.dv.desc_ref(), ptr(params), c_float(self.policy.min_log_std), st)
        out = eng.evaluate(params, batch.obs, batch.act, w, old_ll, old_mean, p_old[d - m:], N_glob)
        host = torch.cat([out, surr_before_dev]).cpu().numpy()
        params_after_opt = params.cpu().numpy()
        eng.dist.check()
        surr_after, kl_dist, surr_before = float(host[0]), float(host[1]), float(host[2])
        self.policy.set_param_values(params_after_opt, set_new=True, set_old=True)
        t_opt = timer.time() - ts
        base_stats = self._finish_stats(stats, paths)
        if self.save_logs:
            self.logger.log_kv('t_opt', t_opt)
            self.logger.log_kv('kl_dist', kl_dist)
            self.logger.log_kv('surr_improvement', surr_after - surr_before)
            self.logger.log_kv('running_score', self.running_score)
            self._log_success(paths)
        self.last_update = dict(surr_before=surr_before, surr_after=surr_after, kl_dist=kl_dist,
                                steps=int(idx.shape[0]))
        return base_stats
