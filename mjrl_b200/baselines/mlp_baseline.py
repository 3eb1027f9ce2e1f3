"""Drop-in for mjrl/baselines/mlp_baseline.py on the CUDA path.

Same constructor, `fit(paths, return_errors=False)` and `predict(path)` as the reference
(mlp_baseline.py:15-115).  `self.model` stays a host `nn.Sequential` with the reference's
module names (`fc_0`, `relu_0`, ...), so pickles, `deepcopy` and seed-for-seed initialisation
match; the arithmetic runs in csrc/mlpbase.cu: `predict` as a persistent tile kernel over the
flat buffer, `fit` as ONE persistent CTA that walks the whole serial chain of Adam steps
(`epochs x (floor(N/bs) - 1)` dependent minibatch steps).

The minibatch order comes from `np.random.permutation` on the host exactly like :84, so with
the same numpy seed the GPU chain visits the same rows as the reference.  Adam's moments and
step count persist across `fit` calls like `self.optimizer`'s state does in the reference.
Unlike the reference, importing this module does not touch CUDA_VISIBLE_DEVICES.
"""
import numpy as np
import torch
import torch.nn as nn

from .. import _lib
from .._lib import c_double, c_i64, c_int, ptr, stream_ptr
from ..batch import PathBatch, batch_for


def _allgather_rows(dist, feat, y32):
    """Concatenate every rank's (feat [N_r, F], y32 [N_r]) in rank order on every rank."""
    td, dev = dist.td, feat.device
    n_local = torch.tensor([feat.shape[0]], dtype=torch.int64, device=dev)
    sizes = [torch.zeros_like(n_local) for _ in range(dist.world)]
    td.all_gather(sizes, n_local, group=dist.group)
    sizes = [int(s.item()) for s in sizes]
    n_max, F = max(sizes), feat.shape[1]
    packed = torch.zeros((n_max, F + 1), dtype=torch.float32, device=dev)
    packed[:feat.shape[0], :F] = feat
    packed[:feat.shape[0], F] = y32
    parts = [torch.empty_like(packed) for _ in range(dist.world)]
    td.all_gather(parts, packed, group=dist.group)
    allrows = torch.cat([p[:k] for p, k in zip(parts, sizes)])
    return allrows[:, :F].contiguous(), allrows[:, F].contiguous(), int(sum(sizes))


def _broadcast_schedule(dist, base, n_rows, dev):
    """Rank 0 draws the minibatch schedule from ITS numpy stream; everyone receives it."""
    td = dist.td
    per_epoch = int(n_rows / base.batch_size) - 1
    nsteps = max(per_epoch, 0) * base.epochs
    if nsteps == 0:
        return np.zeros(0, dtype=np.int32), 0
    if dist.rank == 0:
        perm, _ = base.minibatch_schedule(n_rows)
        d_perm = torch.from_numpy(perm).to(dev)
    else:
        d_perm = torch.empty(nsteps * base.batch_size, dtype=torch.int32, device=dev)
    src = td.get_global_rank(dist.group, 0) if dist.group is not None else 0
    td.broadcast(d_perm, src=src, group=dist.group)
    return d_perm, nsteps


class MLPBaseline:
    def __init__(self, env_spec, obs_dim=None, learn_rate=1e-3, reg_coef=0.0,
                 batch_size=64, epochs=1, use_gpu=False):
        self.n = obs_dim if obs_dim is not None else env_spec.observation_dim
        self.batch_size = batch_size
        self.epochs = epochs
        self.reg_coef = reg_coef
        self.learn_rate = learn_rate
        self.use_gpu = use_gpu      # kept for signature compatibility; the CUDA path is the only path

        self.model = nn.Sequential()
        self.model.add_module('fc_0', nn.Linear(self.n + 4, 128))
        self.model.add_module('relu_0', nn.ReLU())
        self.model.add_module('fc_1', nn.Linear(128, 128))
        self.model.add_module('relu_1', nn.ReLU())
        self.model.add_module('fc_2', nn.Linear(128, 1))
        self.adam_betas = (0.9, 0.999)
        self.adam_eps = 1e-8
        self.adam_step = 0
        self._dev = None            # device state: params, m, v (dropped on pickle)
        self._host_m = None
        self._host_v = None
        self.last_fit_steps = 0

    # ---- flat parameter vector <-> host model --------------------------------------------
    def get_param_values(self):
        return torch.cat([p.data.reshape(-1) for p in self.model.parameters()]).numpy().astype(np.float32)

    def set_param_values(self, flat):
        flat = torch.from_numpy(np.ascontiguousarray(flat, dtype=np.float32))
        i = 0
        for p in self.model.parameters():
            k = p.numel()
            p.data.copy_(flat[i:i + k].reshape(p.shape))
            i += k
        if self._dev is not None:
            self._dev["params"].copy_(flat)

    def _state(self, device):
        if self._dev is None or self._dev["params"].device != device:
            d = int(sum(p.numel() for p in self.model.parameters()))
            z = lambda h: (torch.zeros(d, dtype=torch.float32, device=device) if h is None
                           else torch.from_numpy(h).to(device))
            self._dev = dict(params=torch.from_numpy(self.get_param_values()).to(device),
                             m=z(self._host_m), v=z(self._host_v))
        return self._dev

    def _sync_host(self):
        """Device parameters -> self.model (one D2H of 75 KB)."""
        if self._dev is None:
            return
        flat = self._dev["params"].cpu()
        i = 0
        for p in self.model.parameters():
            k = p.numel()
            p.data.copy_(flat[i:i + k].reshape(p.shape))
            i += k

    # ---- device-level API -------------------------------------------------------------------
    def predict_batch(self, batch: PathBatch):
        st = self._state(batch.device)
        out = batch.alloc_f64()
        batch.ctx.call("mjrl_mlpbase_predict", ptr(batch.obs), ptr(batch.offsets), c_int(batch.P),
                       c_i64(batch.N), c_int(batch.n), ptr(st["params"]), ptr(out), stream_ptr())
        return out

    def _rel_error(self, batch, dist=None):
        pred = self.predict_batch(batch)
        sums = torch.empty(2, dtype=torch.float64, device=batch.device)
        batch.ctx.call("mjrl_mlpbase_sqerr", ptr(pred), ptr(batch.returns), c_i64(batch.N), ptr(sums),
                       stream_ptr())
        if dist is not None and dist.world > 1:
            dist.allreduce(sums)
        return sums

    def minibatch_schedule(self, num_samples):
        """Row indices of every Adam step, drawn like mlp_baseline.py:83-86 (one global-numpy
        permutation per epoch, the last full batch and the tail skipped)."""
        per_epoch = int(num_samples / self.batch_size) - 1
        chunks = []
        for _ in range(self.epochs):
            rand_idx = np.random.permutation(num_samples)
            if per_epoch > 0:
                chunks.append(rand_idx[:per_epoch * self.batch_size])
        if not chunks:
            return np.zeros(0, dtype=np.int32), 0
        return np.concatenate(chunks).astype(np.int32), per_epoch * self.epochs

    def fit_batch(self, batch: PathBatch, return_errors=False, perm=None, sync=True, dist=None):
        """One `fit` on the flat buffer.  Multi-GPU (`dist.world > 1`): the chain of Adam steps
        is serial and cannot be sharded, so this is REPLICAS ONLY -- every rank
        all-gathers the (features, returns) rows of all ranks, rank 0 draws the permutation and
        broadcasts it, and every rank runs the identical deterministic chain."""
        if batch.returns is None:
            batch.returns = batch.gather("returns")
        st = self._state(batch.device)
        dev = batch.device
        F = batch.n + 4
        feat = torch.empty((batch.N, F), dtype=torch.float32, device=dev)
        y32 = torch.empty(batch.N, dtype=torch.float32, device=dev)
        batch.ctx.call("mjrl_mlpbase_features", ptr(batch.obs), ptr(batch.returns), ptr(batch.offsets),
                       c_int(batch.P), c_i64(batch.N), c_int(batch.n), ptr(feat), ptr(y32), stream_ptr())
        multi = dist is not None and dist.world > 1
        before = self._rel_error(batch, dist) if return_errors else None
        n_rows = batch.N
        if multi:
            feat, y32, n_rows = _allgather_rows(dist, feat, y32)
        if perm is None:
            if multi:
                perm, nsteps = _broadcast_schedule(dist, self, n_rows, dev)
            else:
                perm, nsteps = self.minibatch_schedule(n_rows)
        else:
            perm = np.ascontiguousarray(perm, dtype=np.int32).reshape(-1)
            nsteps = perm.size // self.batch_size
        if nsteps > 0:
            d_perm = perm if torch.is_tensor(perm) else torch.from_numpy(perm).to(dev, non_blocking=True)
            batch.ctx.call("mjrl_mlpbase_fit", ptr(feat), ptr(y32), c_i64(n_rows), c_int(batch.n),
                           ptr(st["params"]), ptr(st["m"]), ptr(st["v"]), c_i64(self.adam_step),
                           ptr(d_perm), c_i64(nsteps), c_int(self.batch_size),
                           c_double(self.learn_rate), c_double(self.adam_betas[0]),
                           c_double(self.adam_betas[1]), c_double(self.adam_eps),
                           c_double(self.reg_coef), stream_ptr())
            self.adam_step += nsteps
        self.last_fit_steps = nsteps
        errs = None
        if return_errors:
            after = self._rel_error(batch, dist)
            b, a = before.cpu().numpy(), after.cpu().numpy()
            errs = (float(b[0] / (b[1] + 1e-8)), float(a[0] / (a[1] + 1e-8)))
        if sync:
            self._sync_host()
        return errs

    # ---- reference-compatible API -------------------------------------------------------------
    def fit(self, paths, return_errors=False):
        batch = batch_for(paths)
        if "returns" in paths[0]:
            batch.returns = batch.current("returns", batch.returns)
        return self.fit_batch(batch, return_errors=return_errors)

    def predict(self, path):
        T = len(path["rewards"])
        obs = np.asarray(path["observations"]).reshape(T, -1)
        b = PathBatch(obs, None, np.zeros(T), np.array([0, T]), np.array([0], dtype=np.uint8))
        return self.predict_batch(b).cpu().numpy().astype(np.float32)

    def __getstate__(self):
        d = dict(self.__dict__)
        if self._dev is not None:
            self._sync_host()
            d["_host_m"] = self._dev["m"].cpu().numpy()
            d["_host_v"] = self._dev["v"].cpu().numpy()
        d["_dev"] = None
        return d
