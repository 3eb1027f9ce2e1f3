"""Drop-in for mjrl/algos/dapg.py on the CUDA path: NPG whose gradient is augmented with a
behaviour-cloning term on demonstration data."""
import numpy as np
import torch

from ..batch import PathBatch
from ..utils.logger import DataLog
from .npg_cg import NPG


class DAPG(NPG):
    def __init__(self, env, policy, baseline, demo_paths=None, normalized_step_size=0.01,
                 FIM_invert_args={'iters': 10, 'damping': 1e-4}, hvp_sample_frac=1.0, seed=None,
                 save_logs=False, kl_dist=None, lam_0=1.0, lam_1=0.95):
        self.env = env
        self.policy = policy
        self.baseline = baseline
        self.kl_dist = kl_dist if kl_dist is not None else 0.5 * normalized_step_size
        self.seed = seed
        self.save_logs = save_logs
        self.FIM_invert_args = FIM_invert_args
        self.hvp_subsample = hvp_sample_frac
        self.running_score = None
        self.demo_paths = demo_paths
        self.lam_0 = lam_0
        self.lam_1 = lam_1
        self.iter_count = 0.0
        self.alpha = None
        self.n_step_size = 2.0 * self.kl_dist
        if save_logs: self.logger = DataLog()

    def __getstate__(self):
        d = super().__getstate__()
        d.pop("_demo_batch", None)
        return d

    def _demos(self, eng):
        b = getattr(self, "_demo_batch", None)
        if b is None or b[0] is not self.demo_paths:
            db = PathBatch.from_paths(self.demo_paths, device=eng.ctx.device)
            b = self._demo_batch = (self.demo_paths, db)
        return b[1]

    def train_from_paths(self, paths):
        """dapg.py:54-141.  all_adv = 1e-2*[adv/(std(adv)+1e-8) ; lam_0*lam_1^iter]; the gradient
        is sample_coef * mean over [samples ; demos] = (sum over both)/N_samples; CG, surrogate
        and KL use the samples only."""
        eng, batch, w, stats = self._prepare(paths)
        use_demo = self.demo_paths is not None and self.lam_0 > 0.0
        demo = None
        if use_demo:
            # std of the whitened advantages = sigma/(sigma + 1e-6) (dapg.py:60,70)
            mom = eng.mom.cpu().numpy()
            sigma = float(np.sqrt(mom[2] / mom[0]))
            std_white = sigma / (sigma + 1e-6)
            w_grad = eng.whiten(batch.adv, eps=1e-6, scale=1e-2 / (std_white + 1e-8))
            d_obs = d_act = d_w = None
            if eng.dist.rank == 0:   # demonstrations are replicated work: rank 0 adds them once
                db = self._demos(eng)
                d_obs, d_act = db.obs, db.act
                d_w = torch.full((db.N,), 1e-2 * self.lam_0 * (self.lam_1 ** self.iter_count),
                                 dtype=torch.float32, device=eng.device)
            self.iter_count += 1
            demo = (w_grad, d_obs, d_act, d_w)
        # surr_before uses the plain whitened advantages (dapg.py:93)
        st = self._npg_direction(eng, batch, w, demo=demo)
        sb = None
        if use_demo and st["same"]:      # LR == 1: surrogate = mean of the whitened advantages
            sb = (w.double().sum().reshape(1))
            eng.dist.allreduce(sb)
            sb = sb / st["N_glob"]
        elif use_demo:
            sb = eng.evaluate(st["curr"], batch.obs, batch.act, w, st["old_ll"], st["old_mean"],
                              st["p_old"][eng.dv.d - eng.dv.m:], st["N_glob"])[0:1].clone()
        n_step_size = 2.0 * self.kl_dist
        new_dev = eng.step(st["curr"], st["g_scale"], n_step_size)
        out = eng.evaluate(new_dev, batch.obs, batch.act, w, st["old_ll"], st["old_mean"],
                           st["p_old"][eng.dv.d - eng.dv.m:], st["N_glob"])
        return self._commit(eng, st, new_dev, out, paths, stats, surr_before_dev=sb)
