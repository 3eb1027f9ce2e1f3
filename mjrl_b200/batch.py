"""The flat, segment-indexed timestep buffer that every kernel works on.

The reference passes `paths`, a python list of dicts (mjrl/samplers/base_sampler.py:76-83),
and concatenates it again in every consumer (npg_cg.py:87-89, linear_baseline.py:22-23, ...).
Here the list is flattened ONCE into pinned host staging, copied to the GPU in one H2D per
array, and kept resident for the whole update:

    obs   fp32 [N, n]     act  fp32 [N, m]      (the fp32 casts of gaussian_mlp.py:103-104)
    rew   fp64 [N]        returns / baseline / adv  fp64 [N]
    path_offsets int32 [P+1]        terminated uint8 [P]
"""
import ctypes

import numpy as np
import torch

from . import _lib
from ._lib import c_double, c_i64, c_int, ptr, stream_ptr


def _pinned(shape, dtype):
    return torch.empty(shape, dtype=dtype, pin_memory=True)


# the action block of a large pinned fp32 upload travels on its own stream behind the observations (MJRL_B200_OVERLAP_ACT=0: off)
import os as _os
OVERLAP_ACT = _os.environ.get("MJRL_B200_OVERLAP_ACT", "1") != "0"
_copy_streams = {}


def _copy_stream(device):
    key = (device.index, )
    if key not in _copy_streams:
        _copy_streams[key] = torch.cuda.Stream(device=device)
    return _copy_streams[key]


class PathBatch:
    def __init__(self, obs, act, rew, path_offsets, terminated, device=None, paths=None):
        """`obs` [N,n], `act` [N,m], `rew` [N] are HOST arrays (numpy fp64/fp32 or pinned
        torch tensors); path_offsets int [P+1]; terminated bool/uint8 [P]."""
        self.ctx = _lib.context(device)
        self.device = torch.device("cuda", self.ctx.device)
        self.paths = paths
        off = np.ascontiguousarray(np.asarray(path_offsets), dtype=np.int32)
        self.P = int(off.shape[0] - 1)
        self.N = int(off[-1])
        self.h_offsets = off
        self.n = int(obs.shape[1]) if obs.ndim == 2 else int(np.prod(obs.shape[1:]))
        self.m = int(act.shape[1]) if act is not None else 0
        dev = self.device
        self.offsets = torch.from_numpy(off).to(dev, non_blocking=False)
        self.terminated = torch.from_numpy(
            np.ascontiguousarray(np.asarray(terminated), dtype=np.uint8)).to(dev)
        # copy order = order of first use: rewards (returns scan), observations (baseline predictions, GAE, moments), then the
        # actions, which nothing reads before the policy gradient pass: a pinned fp32 action block goes over a copy stream
        # behind the observations, and `act` makes the consuming stream wait for it on first use, so ~0.2 ms of a 1M-timestep
        # upload overlaps with the scans instead of preceding them
        self.rew = self._upload_f64(rew)
        self.obs = self._upload_f32(obs, (self.N, self.n))
        self._act = None
        self._act_event = None
        if act is not None:
            h = self._host_tensor(act)
            if OVERLAP_ACT and h.dtype == torch.float32 and h.is_pinned() and self.N * self.m >= (1 << 20):
                self._upload_act_behind(h.reshape(self.N, self.m))
            else:
                self._act = self._upload_f32(act, (self.N, self.m))
        self.returns = None
        self.baseline = None
        self.adv = None
        self.h2d_bytes = (self._nbytes(obs) + (self._nbytes(act) if act is not None else 0)
                          + self._nbytes(rew) + off.nbytes + self.P)

    # ---- actions: possibly still in flight on the copy stream -----------------------------
    @property
    def act(self):
        ev = self._act_event
        if ev is not None:
            torch.cuda.current_stream(self.device).wait_event(ev)     # the consumer's stream, once
            self._act_event = None
        return self._act

    @act.setter
    def act(self, value):
        self._act, self._act_event = value, None

    def _upload_act_behind(self, h):
        dev = self.device
        cur = torch.cuda.current_stream(dev)
        side = _copy_stream(dev)
        d = torch.empty(h.shape, dtype=torch.float32, device=dev)     # allocated on the consumer's stream
        after_obs = torch.cuda.Event()
        after_obs.record(cur)                 # behind the observation copy (and anything that used this memory before)
        side.wait_event(after_obs)
        with torch.cuda.stream(side):
            d.copy_(h, non_blocking=True)
            done = torch.cuda.Event()
            done.record(side)
        d.record_stream(side)
        self._act, self._act_event = d, done

    @staticmethod
    def _nbytes(a):
        return int(a.numel() * a.element_size()) if torch.is_tensor(a) else int(a.nbytes)

    def _host_tensor(self, a):
        if torch.is_tensor(a):
            return a
        return torch.from_numpy(np.ascontiguousarray(a))

    def _upload_f32(self, a, shape):
        h = self._host_tensor(a).reshape(shape)
        d = h.to(self.device, non_blocking=True)
        if d.dtype == torch.float32:
            return d
        if d.dtype != torch.float64:
            raise TypeError("observations/actions must be float32 or float64")
        out = torch.empty(shape, dtype=torch.float32, device=self.device)
        self.ctx.call("mjrl_cast_f64_to_f32", ptr(d), c_i64(d.numel()), ptr(out), stream_ptr())
        return out

    def _upload_f64(self, a):
        h = self._host_tensor(a).reshape(-1)
        if h.dtype != torch.float64:
            h = h.double()
        return h.to(self.device, non_blocking=True)

    # ---- construction from the reference's list-of-dicts -----------------------------
    @classmethod
    def from_paths(cls, paths, device=None, with_actions=True):
        lens = np.fromiter((len(p["rewards"]) for p in paths), dtype=np.int64, count=len(paths))
        off = np.zeros(len(paths) + 1, dtype=np.int64)
        np.cumsum(lens, out=off[1:])
        N = int(off[-1])
        o0 = np.asarray(paths[0]["observations"])
        n = int(np.prod(o0.shape[1:]))
        obs = _pinned((N, n), torch.float64).numpy()
        rew = _pinned((N,), torch.float64).numpy()
        act = None
        if with_actions and "actions" in paths[0]:
            m = int(np.asarray(paths[0]["actions"]).shape[1])
            act = _pinned((N, m), torch.float64).numpy()
        for i, p in enumerate(paths):
            s, e = off[i], off[i + 1]
            obs[s:e] = np.asarray(p["observations"]).reshape(e - s, n)
            rew[s:e] = p["rewards"]
            if act is not None:
                act[s:e] = p["actions"]
        term = np.fromiter((bool(p.get("terminated", False)) for p in paths), dtype=np.uint8,
                           count=len(paths))
        return cls(obs, act, rew, off, term, device=device, paths=paths)

    @classmethod
    def from_device(cls, obs_dev, act_dev, rew, path_offsets, terminated, paths=None):
        """Batch whose observation / action rows are already resident (filed there step by step by
        the lock-step sampler, samplers/lockstep.py): only the reward column is uploaded."""
        self = cls.__new__(cls)
        self.device = obs_dev.device
        self.ctx = _lib.context(self.device.index)
        self.paths = paths
        off = np.ascontiguousarray(np.asarray(path_offsets), dtype=np.int32)
        self.P, self.N, self.h_offsets = int(off.shape[0] - 1), int(off[-1]), off
        self.n, self.m = int(obs_dev.shape[1]), int(act_dev.shape[1])
        if obs_dev.shape[0] != self.N or act_dev.shape[0] != self.N:
            raise ValueError("device rows do not match the path offsets")
        self.offsets = torch.from_numpy(off).to(self.device)
        term8 = np.ascontiguousarray(np.asarray(terminated), dtype=np.uint8)
        self.terminated = torch.from_numpy(term8).to(self.device)
        self.obs, self.act = obs_dev, act_dev
        self.rew = self._upload_f64(rew)
        self.returns = self.baseline = self.adv = None
        self.h2d_bytes = self._nbytes(self.rew) + off.nbytes + self.P
        return self

    # ---- results back into the path dicts (reference contract: fp64 numpy [T]) --------
    def scatter(self, key, tensor):
        if self.paths is None:
            return
        h = tensor.detach().cpu().numpy()
        off = self.h_offsets
        views = []
        for i, p in enumerate(self.paths):
            p[key] = h[off[i]:off[i + 1]]
            views.append(p[key])
        if not hasattr(self, "_views"):
            self._views = {}
        self._views[key] = views

    def current(self, key, dev):
        """Device copy of a per-path field for a consumer that the reference would hand the path dicts to: the
        resident tensor `dev` when every dict still holds exactly the array this batch scattered there, else the
        dicts' arrays re-uploaded -- the reference re-reads `path[key]` on every call (npg_cg.py:87-91,
        linear_baseline.py:22-23), so a caller that REPLACES advantages / returns between compute_advantages and
        train_from_paths / fit is honoured.  (Editing a scattered array in place is not detected: drop_cache().)"""
        if self.paths is None or key not in self.paths[0]:
            return dev
        views = getattr(self, "_views", {}).get(key)
        if dev is not None and views is not None and len(views) == len(self.paths) and \
                all(p.get(key) is v for p, v in zip(self.paths, views)):
            return dev
        return self.gather(key)

    def gather(self, key):
        """Upload a per-path fp64 field of the dicts (e.g. 'returns' set by someone else)."""
        h = np.concatenate([np.asarray(p[key], dtype=np.float64) for p in self.paths])
        return torch.from_numpy(h).to(self.device)

    def alloc_f64(self):
        return torch.empty(self.N, dtype=torch.float64, device=self.device)


_cache = {"key": None, "batch": None}


def _signature(paths):
    return (id(paths), len(paths), id(paths[0]["observations"]), id(paths[-1]["rewards"]))


def batch_for(paths, device=None):
    """PathBatch of `paths`, built once per list (1-entry cache keyed on object identity) so
    compute_returns -> compute_advantages -> train_from_paths -> fit share one upload."""
    resident = getattr(paths, "batch", None)     # lock-step rollouts arrive already filed on the device
    if resident is not None and resident.P == len(paths):
        return resident
    key = _signature(paths)
    if _cache["key"] == key:
        return _cache["batch"]
    b = PathBatch.from_paths(paths, device=device)
    _cache["key"], _cache["batch"] = key, b
    return b


def drop_cache():
    _cache["key"], _cache["batch"] = None, None
