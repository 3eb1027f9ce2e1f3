"""Drop-in for mjrl/baselines/linear_baseline.py on the CUDA path.

Same constructor, `fit(paths, return_errors=False)` and `predict(path)` as the reference.
The feature build is fused with the fp64 Gram accumulation (csrc/linbase.cu); the
regularised solve, the NaN retry ladder (:30-38) and the error ratios (:25-28,40-44) run on
the device.  `_coeffs` stays a host numpy fp64 [n+4] (or None), so the object pickles and
deep-copies like the reference's.
"""
import copy

import numpy as np
import torch

from .. import _lib
from .._lib import c_double, c_i64, c_int, ptr, stream_ptr
from ..batch import PathBatch, batch_for


class LinearBaseline:
    _gram_fn, _predict_fn = "mjrl_linbase_gram", "mjrl_linbase_predict"

    @staticmethod
    def _num_feat(n):
        return n + 4

    def __init__(self, env_spec, reg_coeff=1e-5):
        self.n = env_spec.observation_dim
        self._reg_coeff = reg_coeff
        self._coeffs = None
        self._dev_coeffs = None   # device mirror, dropped on pickle
        self.last_fit_info = None

    # ---- device-level API --------------------------------------------------------
    def _coeffs_on(self, device):
        if self._coeffs is None:
            return None
        if self._dev_coeffs is None or self._dev_coeffs.device != device:
            self._dev_coeffs = torch.from_numpy(np.asarray(self._coeffs, dtype=np.float64)).to(device)
        return self._dev_coeffs

    def predict_batch(self, batch: PathBatch):
        out = batch.alloc_f64()
        batch.ctx.call(self._predict_fn, ptr(batch.obs), ptr(batch.offsets), c_int(batch.P),
                       c_i64(batch.N), c_int(batch.n), ptr(self._coeffs_on(batch.device)),
                       ptr(out), stream_ptr())
        return out

    def fit_batch(self, batch: PathBatch, return_errors=False, sync=True, allreduce=None, dist=None):
        """Gram -> (optional all-reduce over ranks) -> Cholesky.  Returns (err_before, err_after)
        when asked.  With sync=False the coefficients stay on the device until `coeffs` is read."""
        if batch.returns is None:
            batch.returns = batch.gather("returns")
        F = self._num_feat(batch.n)
        gram = torch.empty(F * (F + 1) + 1, dtype=torch.float64, device=batch.device)
        batch.ctx.call(self._gram_fn, ptr(batch.obs), ptr(batch.returns), ptr(batch.offsets),
                       c_int(batch.P), c_i64(batch.N), c_int(batch.n), ptr(gram), stream_ptr())
        if allreduce is None and dist is not None and dist.world > 1:
            allreduce = dist.allreduce
        if allreduce is not None:
            allreduce(gram)
        prev = self._coeffs_on(batch.device)
        both = torch.empty(F + 4, dtype=torch.float64, device=batch.device)    # coefficients | solve info: one read-back
        new, info = both[:F], both[F:]
        batch.ctx.call("mjrl_linbase_solve", ptr(gram), c_int(F), c_double(copy.deepcopy(self._reg_coeff)),
                       ptr(prev), ptr(new), ptr(info), stream_ptr())
        errs = None
        if sync or return_errors:
            hb = _lib.to_host(both)
            self._coeffs, h = hb[:F].copy(), hb[F:]
            self.last_fit_info = dict(reg_coeff=h[0], retries=int(h[1]))
            if return_errors:
                errs = (float(h[2]), float(h[3]))
        else:
            self._coeffs = _LazyCoeffs(new)
        self._dev_coeffs = new   # after _coeffs: assigning _coeffs drops the device mirror
        return errs

    # ---- reference-compatible API ---------------------------------------------------
    def fit(self, paths, return_errors=False):
        batch = batch_for(paths)
        if "returns" in paths[0]:
            batch.returns = batch.current("returns", batch.returns)
        return self.fit_batch(batch, return_errors=return_errors)

    def predict(self, path):
        T = len(path["rewards"])
        if self._coeffs is None:
            return np.zeros(T)
        obs = np.asarray(path["observations"]).reshape(T, -1)
        b = PathBatch(obs, None, np.zeros(T), np.array([0, T]), np.array([0], dtype=np.uint8))
        return self.predict_batch(b).cpu().numpy()

    def __getstate__(self):
        d = dict(self.__dict__)
        d["_dev_coeffs"] = None
        if isinstance(d["_coeffs"], _LazyCoeffs):
            d["_coeffs"] = d["_coeffs"].numpy()
        return d

    def __setattr__(self, k, v):
        if k == "_coeffs":
            object.__setattr__(self, "_dev_coeffs", None)
        object.__setattr__(self, k, v)


class _LazyCoeffs:
    """Device-resident coefficients that materialise on the host on first numpy use."""

    def __init__(self, dev):
        self.dev = dev
        self._h = None

    def numpy(self):
        if self._h is None:
            self._h = self.dev.cpu().numpy()
        return self._h

    def __array__(self, dtype=None, copy=None):
        a = self.numpy()
        return a.astype(dtype) if dtype is not None else a
