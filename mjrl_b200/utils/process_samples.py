"""Drop-in for mjrl/utils/process_samples.py on the CUDA path.

Same names, argument meaning and in-place effects as the reference: each path dict gains
`returns`, `baseline`, `advantages` as fp64 numpy [T].  The arithmetic runs in the
segmented-scan kernels of csrc/scans.cu over the flat buffer (mjrl_b200/batch.py).
"""
import numpy as np
import torch

from .. import _lib
from .._lib import c_double, c_i64, c_int, ptr, stream_ptr
from ..batch import PathBatch, batch_for


# ---- device-level API (flat buffers; what train_step uses) -----------------------------

def returns_on_device(batch: PathBatch, gamma):
    """process_samples.py:3-5 over the whole batch; result stays on the GPU."""
    batch.returns = batch.alloc_f64()
    batch.ctx.call("mjrl_returns_scan", ptr(batch.rew), ptr(batch.offsets), c_int(batch.P),
                   c_i64(batch.N), c_double(gamma), ptr(batch.returns), stream_ptr())
    return batch.returns


def advantages_on_device(batch: PathBatch, baseline_values, gamma, gae_lambda=None,
                         normalize=False):
    """process_samples.py:7-35 given the baseline predictions [N] (fp64, device)."""
    batch.baseline = baseline_values
    batch.adv = batch.alloc_f64()
    ctx = batch.ctx
    if gae_lambda is None or gae_lambda < 0.0 or gae_lambda > 1.0:
        if batch.returns is None:
            raise ValueError("plain advantages need returns: call compute_returns first")
        ctx.call("mjrl_adv_plain", ptr(batch.returns), ptr(baseline_values), c_i64(batch.N),
                 ptr(batch.adv), stream_ptr())
    else:
        ctx.call("mjrl_gae_scan", ptr(batch.rew), ptr(baseline_values), ptr(batch.offsets),
                 ptr(batch.terminated), c_int(batch.P), c_i64(batch.N), c_double(gamma),
                 c_double(gae_lambda), ptr(batch.adv), stream_ptr())
    if normalize:
        mom = torch.empty(4, dtype=torch.float64, device=batch.device)
        ctx.call("mjrl_moments", ptr(batch.adv), c_i64(batch.N), ptr(mom), stream_ptr())
        ctx.call("mjrl_normalize_inplace", ptr(batch.adv), c_i64(batch.N), ptr(mom),
                 c_double(1e-8), stream_ptr())
    return batch.adv


def baseline_values_on_device(batch: PathBatch, baseline):
    """baseline.predict over the whole flat buffer (device fp64 [N])."""
    if hasattr(baseline, "predict_batch"):
        return baseline.predict_batch(batch)
    # foreign baseline object (e.g. the reference's QuadraticBaseline): per-path host calls
    vals = np.concatenate([np.asarray(baseline.predict(p), dtype=np.float64).reshape(-1)
                           for p in batch.paths])
    return torch.from_numpy(vals).to(batch.device)


# ---- reference-compatible API -------------------------------------------------------

def compute_returns(paths, gamma):
    batch = batch_for(paths)
    batch.scatter("returns", returns_on_device(batch, gamma))


def compute_advantages(paths, baseline, gamma, gae_lambda=None, normalize=False):
    batch = batch_for(paths)
    if batch.returns is None and "returns" in paths[0]:
        batch.returns = batch.gather("returns")
    b = baseline_values_on_device(batch, baseline)
    adv = advantages_on_device(batch, b, gamma, gae_lambda, normalize)
    batch.scatter("baseline", b)
    batch.scatter("advantages", adv)


def discount_sum(x, gamma, terminal=0.0):
    """process_samples.py:37-43 for one sequence (host in, host out; runs the scan kernel)."""
    x = np.asarray(x, dtype=np.float64)
    if x.size == 0:
        return x.copy()
    ctx = _lib.context()
    dev = torch.device("cuda", ctx.device)
    xs = x.copy()
    if terminal != 0.0:
        xs[-1] = xs[-1] + gamma * terminal
    d = torch.from_numpy(xs).to(dev)
    off = torch.tensor([0, x.size], dtype=torch.int32, device=dev)
    out = torch.empty_like(d)
    ctx.call("mjrl_returns_scan", ptr(d), ptr(off), c_int(1), c_i64(x.size), c_double(gamma),
             ptr(out), stream_ptr())
    return out.cpu().numpy()
