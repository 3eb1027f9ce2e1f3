"""Drop-in for mjrl/baselines/quadratic_baseline.py on the CUDA path.

Same constructor, `fit(paths, return_errors=False)` and `predict(path)` as the reference
(quadratic_baseline.py:4-68).  Features [o, o_i o_j (i <= j), 1, al..al^4] with o = clip(obs,-10,10)/10
are built inside the fp64 tensor-core Gram kernel (csrc/linbase.cu, `mjrl_quadbase_gram`); the ridge
solve with the NaN ladder and the error ratios are LinearBaseline's (`mjrl_linbase_solve`).
obs_dim <= 16 (n + n(n+1)/2 + 5 = 157 features).
"""
from .linear_baseline import LinearBaseline


class QuadraticBaseline(LinearBaseline):
    _gram_fn, _predict_fn = "mjrl_quadbase_gram", "mjrl_quadbase_predict"

    @staticmethod
    def _num_feat(n):
        return n + n * (n + 1) // 2 + 5

    def __init__(self, env_spec, reg_coeff=1e-5, obs_dim=None):
        super().__init__(env_spec, reg_coeff=reg_coeff)
        self.n = obs_dim if obs_dim is not None else env_spec.observation_dim
