"""Drop-in for mjrl/algos/trpo.py on the CUDA path: NPG direction + KL backtracking."""
import numpy as np
import torch

from ..utils.logger import DataLog
from .. import _lib
from .npg_cg import NPG


class TRPO(NPG):
    def __init__(self, env, policy, baseline, kl_dist=0.01,
                 FIM_invert_args={'iters': 10, 'damping': 1e-4}, hvp_sample_frac=1.0, seed=None,
                 save_logs=False, normalized_step_size=0.01):
        self.env = env
        self.policy = policy
        self.baseline = baseline
        self.kl_dist = kl_dist if kl_dist is not None else 0.5 * normalized_step_size
        self.seed = seed
        self.save_logs = save_logs
        self.FIM_invert_args = FIM_invert_args
        self.hvp_subsample = hvp_sample_frac
        self.running_score = None
        self.alpha = None
        self.n_step_size = 2.0 * self.kl_dist
        if save_logs: self.logger = DataLog()

    def train_from_paths(self, paths, direction_override=None, verbose=True):
        """trpo.py:54-145.  The backtracking loop (:105-124) is data dependent, so each trial
        reads its two scalars back (one small D2H); everything else stays on the device.
        `direction_override` feeds a given npg direction into the line search (test hook)."""
        eng, batch, w, stats = self._prepare(paths)
        st = self._npg_direction(eng, batch, w)
        if direction_override is not None:
            eng.cg_x.copy_(torch.from_numpy(np.asarray(direction_override, dtype=np.float64)).to(eng.device))
        d, m = eng.dv.d, eng.dv.m
        n_step_size = 2.0 * self.kl_dist
        old_ls = st["p_old"][d - m:]
        surr_before = float((eng.grad[d] / st["N_glob"]).item())
        scale, trials = 1.0, []
        for k in range(100):
            new_dev = eng.step(st["curr"], st["g_scale"], n_step_size, alpha_scale=scale)
            out = eng.evaluate(new_dev, batch.obs, batch.act, w, st["old_ll"], st["old_mean"], old_ls,
                               st["N_glob"])
            out = _lib.to_host(out)
            surr_after, kl = float(out[0]), float(np.float32(out[1]))
            ok = bool(kl < self.kl_dist)
            trials.append((kl, surr_after, ok))
            if ok:
                break
            scale = 0.9 * scale
            if verbose:
                print("Step size too high. Backtracking. | kl = %f | surr diff = %f" %
                      (kl, surr_after - surr_before))
            if k == 99:
                scale = 0.0
        new_dev = eng.step(st["curr"], st["g_scale"], n_step_size, alpha_scale=scale)
        out = eng.evaluate(new_dev, batch.obs, batch.act, w, st["old_ll"], st["old_mean"], old_ls,
                           st["N_glob"])
        alpha0 = float(eng.scal4[0].item())
        return self._commit(eng, st, new_dev, out, paths, stats,
                            extra_log=dict(alpha=alpha0 * scale, alpha0=alpha0, trials=trials))
