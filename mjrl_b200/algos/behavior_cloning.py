"""Drop-in for mjrl/algos/behavior_cloning.py (and, through `mode="mse"`, behavior_cloning_2.py) on
the CUDA path.

Same constructor, `loss` and `train` as the reference (behavior_cloning.py:11-68).  The in/out
transforms are column means / standard deviations of the expert data computed in fp64 numpy like
:25-29 (a few thousand demo timesteps: host glue), then pushed into `policy.model` /
`policy.old_model` through `set_transformations` exactly like :32-33.  The chain of
`epochs * int(N / batch_size)` Adam steps (:52-61) runs on the device through ONE C-ABI call per
epoch (`mjrl_policy_adam_chain`): gather rows -> policy gradient pass -> Adam, no host work or
synchronisation between steps.  The with-replacement minibatch rows come from `np.random.choice`
on numpy's global stream exactly like :56, so a seeded run visits the reference's rows.

`self.optimizer`'s role (Adam moments + step count, kept across `train()` calls) is played by
device tensors that are dropped on pickle and re-created from host copies.  An `optimizer`
argument is accepted when it is a `torch.optim.Adam` (its lr / betas / eps / weight_decay are
read); anything else raises, there is no CPU fallback.
"""
import time as timer

import numpy as np
import torch

from .._lib import c_double, c_i64, c_int, ptr, stream_ptr
from ..utils.logger import DataLog


class BC:
    mode = "mle"    # behavior_cloning.py: loss = -mean LL over trainable_params (network + log_std)

    def __init__(self, expert_paths, policy, epochs=5, batch_size=64, lr=1e-3, optimizer=None):
        self.policy = policy
        self.expert_paths = expert_paths
        self.epochs = epochs
        self.mb_size = batch_size
        self.logger = DataLog()

        # get transformations (:25-29)
        observations = np.concatenate([path["observations"] for path in expert_paths])
        actions = np.concatenate([path["actions"] for path in expert_paths])
        in_shift, in_scale = np.mean(observations, axis=0), np.std(observations, axis=0)
        out_shift, out_scale = np.mean(actions, axis=0), np.std(actions, axis=0)

        # set scalings in the target policy (:32-33)
        self.policy.model.set_transformations(in_shift, in_scale, out_shift, out_scale)
        self.policy.old_model.set_transformations(in_shift, in_scale, out_shift, out_scale)
        self._init_policy(out_scale)

        self.hyper = dict(lr=lr, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0)
        if optimizer is not None:
            if not isinstance(optimizer, torch.optim.Adam) or optimizer.param_groups[0].get("amsgrad", False):
                raise NotImplementedError("the CUDA path runs torch.optim.Adam (no amsgrad) only")
            g = optimizer.param_groups[0]
            self.hyper = dict(lr=g["lr"], betas=tuple(g["betas"]), eps=g["eps"], weight_decay=g["weight_decay"])
        self.adam_step = 0
        self._adam_host = None      # (m, v) numpy copies for pickling
        self._adam_dev = None
        self._data = None

    def _init_policy(self, out_scale):
        pass

    # ---- state -------------------------------------------------------------------------
    def __getstate__(self):
        d = dict(self.__dict__)
        if self._adam_dev is not None:
            d["_adam_host"] = tuple(t.cpu().numpy() for t in self._adam_dev)
        d["_adam_dev"] = None
        d["_data"] = None
        return d

    def _adam(self, dv):
        if self._adam_dev is None or self._adam_dev[0].device != dv.device:
            if self._adam_host is not None:
                self._adam_dev = tuple(torch.from_numpy(h).to(dv.device) for h in self._adam_host)
            else:
                z = lambda: torch.zeros(dv.d, dtype=torch.float32, device=dv.device)
                self._adam_dev = (z(), z())
        return self._adam_dev

    def _expert(self, dv):
        """Expert obs / act as device fp32 (the casts of gaussian_mlp.py:103-104), uploaded once."""
        if self._data is None or self._data[0].device != dv.device:
            up = lambda key: torch.from_numpy(np.ascontiguousarray(
                np.concatenate([p[key] for p in self.expert_paths]), dtype=np.float32)).to(dv.device)
            self._data = (up("observations"), up("actions"))
        return self._data

    # ---- loss --------------------------------------------------------------------------
    def _work_params(self, params):
        """Parameter vector the device passes run on: the policy's own for the likelihood loss."""
        return params

    def _loss_dev(self, dv, params, obs, act):
        mean, ll = dv.forward(self._work_params(params), obs, act, want_mean=False, want_ll=True)
        return -torch.mean(ll)      # :38-41

    def loss(self, obs, act):
        """:38-41 on arbitrary host arrays; returns a host torch scalar like the reference."""
        dv = self.policy.device_view()
        f32 = lambda a, k: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).reshape(-1, k).to(dv.device)
        return self._loss_dev(dv, dv.params("new"), f32(obs, dv.n), f32(act, dv.m)).cpu()

    # ---- train -------------------------------------------------------------------------
    def minibatch_indices(self, num_samples):
        """:55-56: one np.random.choice(num_samples, size=mb_size) per step of ONE epoch."""
        steps = int(num_samples / self.mb_size)
        if steps == 0:
            return np.zeros((0, self.mb_size), dtype=np.int32)
        return np.stack([np.random.choice(num_samples, size=self.mb_size) for _ in range(steps)]).astype(np.int32)

    def _d_opt(self, dv):
        return dv.d

    def _grad_scale(self, dv):
        return -1.0 / self.mb_size       # loss = -(1/B) sum_t LL_t

    def train(self, minibatches=None):
        """`minibatches` (tests): int array [epochs * int(N/mb), mb] overriding the global numpy stream."""
        dv = self.policy.device_view()
        ctx, dev = dv.ctx, dv.device
        obs, act = self._expert(dv)
        num_samples = obs.shape[0]
        B, n, m, d = self.mb_size, dv.n, dv.m, dv.d
        params_before_opt = self.policy.get_param_values()
        params = self._work_params(dv.params("new")).clone()
        am, av = self._adam(dv)
        f32 = lambda *s: torch.empty(s, dtype=torch.float32, device=dev)
        mb_obs, mb_act = f32(B, n), f32(B, m)
        ones = torch.ones(B, dtype=torch.float32, device=dev)
        gsum = torch.empty(d, dtype=torch.float64, device=dev)
        per_epoch = int(num_samples / B)
        ts = timer.time()
        h = self.hyper
        for ep in range(self.epochs):
            self.logger.log_kv('epoch', ep)
            loss_val = float(self._loss_dev(dv, params, obs, act).item())
            self.logger.log_kv('loss', loss_val)
            self.logger.log_kv('time', (timer.time() - ts))
            idx = (self.minibatch_indices(num_samples) if minibatches is None else
                   np.asarray(minibatches[ep * per_epoch:(ep + 1) * per_epoch], dtype=np.int32))
            if idx.shape[0] == 0:
                continue
            d_idx = torch.from_numpy(np.ascontiguousarray(idx)).to(dev)
            ctx.call("mjrl_policy_adam_chain", dv.desc_ref(), ptr(params), ptr(obs), ptr(act), ptr(d_idx),
                     c_int(idx.shape[0]), c_int(B), c_i64(self._d_opt(dv)), c_double(self._grad_scale(dv)),
                     ptr(am), ptr(av), c_i64(self.adam_step), c_double(h["lr"]), c_double(h["betas"][0]),
                     c_double(h["betas"][1]), c_double(h["eps"]), c_double(h["weight_decay"]),
                     ptr(mb_obs), ptr(mb_act), ptr(ones), ptr(gsum), stream_ptr())
            ctx.launches += 5 * idx.shape[0] - 1
            self.adam_step += idx.shape[0]
        params_after_opt = params_before_opt.copy()
        k = self._d_opt(dv)
        params_after_opt[:k] = params[:k].cpu().numpy()
        self.policy.set_param_values(params_after_opt, set_new=True, set_old=True)
        self.logger.log_kv('epoch', self.epochs)
        # the reference evaluates this loss AFTER set_param_values, i.e. on the clamped log_std (:62-66)
        loss_val = float(self._loss_dev(dv, dv.params("new"), obs, act).item())
        self.logger.log_kv('loss', loss_val)
        self.logger.log_kv('time', (timer.time() - ts))
