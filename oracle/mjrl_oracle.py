"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the mjrl policy-update hot path.

This file restates, in numpy + CPU torch, the algorithm of the reference
(bennevans/mjrl, pure Python) for the path
    returns -> GAE -> baseline predict/fit -> vanilla gradient -> CG on the
    Hessian-of-KL product -> NPG step / TRPO line search / DAPG.
Every function cites the reference file:line it follows (paths relative to the
reference checkout).  It is written functionally over explicit parameter lists,
not as a copy of the reference's classes.

Parity pin: the reference has no golden vectors, so the
oracle is pinned against the LIVE reference imported from /root/reference in
the build container; `tests/golden/make_golden.py` records both the inputs and
the reference's outputs as fixtures under `tests/golden/`, and
`tests/test_oracle_golden.py` re-checks this file against them on any box.
Versions the pin was taken under: numpy 2.3.5, torch 2.11.0 (the reference
pins numpy 1.14 / torch 0.3.1; see DESIGN.md "numpy-version drift").

Two precision flavours live here:
  * the *faithful* fp32 path (torch autograd, fp32 numpy CG) -- this is what
    the reference computes and what the CPU baseline times;
  * an *fp64 analytic* path (`truth_*`) used as the "truth" leg of the CG
    noise-floor protocol.
"""
import copy
import math

import numpy as np
import torch

# --------------------------------------------------------------------------
# (a) returns and advantages
# --------------------------------------------------------------------------


def discount_sum(x, gamma, terminal=0.0):
    """y[t] = x[t] + gamma*y[t+1], y[T] = terminal; strictly sequential fp64.

    Follows mjrl/utils/process_samples.py:37-43 (python-float accumulator,
    multiply then add, i.e. two roundings per step)."""
    out = np.empty(len(x), dtype=np.float64)
    acc = terminal
    for t in range(len(x) - 1, -1, -1):
        acc = x[t] + gamma * acc
        out[t] = acc
    return out


def compute_returns(paths, gamma):
    """mjrl/utils/process_samples.py:3-5."""
    for p in paths:
        p["returns"] = discount_sum(p["rewards"], gamma)


def _normalize_advantages(paths):
    # mjrl/utils/process_samples.py:14-19 and :30-35 (eps 1e-8, population std)
    flat = np.concatenate([p["advantages"] for p in paths])
    mu, sd = flat.mean(), flat.std()
    for p in paths:
        p["advantages"] = (p["advantages"] - mu) / (sd + 1e-8)


def compute_advantages(paths, baseline, gamma, gae_lambda=None, normalize=False):
    """mjrl/utils/process_samples.py:7-35.

    Plain mode when gae_lambda is None or outside [0,1] (:10-13); otherwise GAE
    (:21-29) where a non-terminated path bootstraps with its OWN last baseline
    value and a terminated path with 0."""
    plain = gae_lambda is None or gae_lambda < 0.0 or gae_lambda > 1.0
    for p in paths:
        b = p["baseline"] = baseline.predict(p)
        if plain:
            p["advantages"] = p["returns"] - b
            continue
        if b.ndim == 1:
            tail = 0.0 if p["terminated"] else b[-1]
            b_ext = np.append(b, tail)
        else:  # multi-agent baselines, :26-27
            tail = np.zeros(b.shape[1]) if p["terminated"] else b[-1]
            b_ext = np.vstack((b, tail))
        td = p["rewards"] + gamma * b_ext[1:] - b_ext[:-1]
        p["advantages"] = discount_sum(td, gamma * gae_lambda)
    if normalize:
        _normalize_advantages(paths)


# --------------------------------------------------------------------------
# (b) baselines
# --------------------------------------------------------------------------


def linear_features(path):
    """[clip(obs,-10,10), t/1000, (t/1000)^2, (t/1000)^3, 1] -- n+4 columns.

    mjrl/baselines/linear_baseline.py:10-18; t is the index within the path."""
    o = np.clip(path["observations"], -10, 10)
    if o.ndim > 2:
        o = o.reshape(o.shape[0], -1)
    T = len(path["rewards"])
    al = np.arange(T).reshape(-1, 1) / 1000.0
    return np.concatenate([o, al, al ** 2, al ** 3, np.ones((T, 1))], axis=1)


class LinearBaseline:
    """Ridge regression on `linear_features`; mjrl/baselines/linear_baseline.py:4-49."""

    def __init__(self, env_spec=None, reg_coeff=1e-5):
        self._reg_coeff = reg_coeff
        self._coeffs = None

    def fit(self, paths, return_errors=False):
        X = np.concatenate([linear_features(p) for p in paths])
        y = np.concatenate([p["returns"] for p in paths])
        if return_errors:  # :25-28
            pred = X.dot(self._coeffs) if self._coeffs is not None else np.zeros(y.shape)
            err_before = np.sum((y - pred) ** 2) / np.sum(y ** 2)
        lam = copy.deepcopy(self._reg_coeff)
        for _ in range(10):  # :30-38, NaN -> x10 regulariser, up to 10 tries
            self._coeffs = np.linalg.lstsq(
                X.T.dot(X) + lam * np.identity(X.shape[1]), X.T.dot(y), rcond=None)[0]
            if not np.any(np.isnan(self._coeffs)):
                break
            lam *= 10
        if return_errors:  # :40-44
            pred = X.dot(self._coeffs)
            err_after = np.sum((y - pred) ** 2) / np.sum(y ** 2)
            return err_before, err_after

    def predict(self, path):
        if self._coeffs is None:  # :46-48
            return np.zeros(len(path["rewards"]))
        return linear_features(path).dot(self._coeffs)


def quadratic_features(paths):
    """[o (n), o_i*o_j for i <= j row-major (n(n+1)/2), 1, al, al^2, al^3, al^4] with
    o = clip(obs,-10,10)/10, al = t/1000.  mjrl/baselines/quadratic_baseline.py:10-37."""
    o = np.concatenate([p["observations"] for p in paths])
    o = np.clip(o, -10, 10) / 10.0
    if o.ndim > 2:
        o = o.reshape(o.shape[0], -1)
    N, n = o.shape
    F = np.ones((N, int(n + n * (n + 1) / 2 + 1 + 4)))
    F[:, :n] = o
    k = n
    for i in range(n):
        for j in range(i, n):
            F[:, k] = o[:, i] * o[:, j]
            k += 1
    row = 0
    for p in paths:
        T = len(p["rewards"])
        al = np.arange(T) / 1000.0
        for j in range(4):
            F[row:row + T, -4 + j] = al ** (j + 1)
        row += T
    return F


class QuadraticBaseline:
    """Ridge regression on `quadratic_features`; mjrl/baselines/quadratic_baseline.py:4-68
    (same normal-equation lstsq and NaN ladder as LinearBaseline, :51-58)."""

    def __init__(self, env_spec=None, reg_coeff=1e-5, obs_dim=None):
        self._reg_coeff = reg_coeff
        self._coeffs = None

    def fit(self, paths, return_errors=False):
        X = quadratic_features(paths)
        y = np.concatenate([p["returns"] for p in paths])
        if return_errors:  # :45-48
            pred = X.dot(self._coeffs) if self._coeffs is not None else np.zeros(y.shape)
            err_before = np.sum((y - pred) ** 2) / np.sum(y ** 2)
        lam = copy.deepcopy(self._reg_coeff)
        for _ in range(10):  # :51-58
            self._coeffs = np.linalg.lstsq(
                X.T.dot(X) + lam * np.identity(X.shape[1]), X.T.dot(y), rcond=None)[0]
            if not np.any(np.isnan(self._coeffs)):
                break
            lam *= 10
        if return_errors:  # :60-64
            pred = X.dot(self._coeffs)
            err_after = np.sum((y - pred) ** 2) / np.sum(y ** 2)
            return err_before, err_after

    def predict(self, path):
        if self._coeffs is None:  # :66-68
            return np.zeros(len(path["rewards"]))
        return quadratic_features([path]).dot(self._coeffs)


def mlp_baseline_features(paths):
    """[clip(obs,-10,10)/10, al, al^2, al^3, al^4], no bias column.

    mjrl/baselines/mlp_baseline.py:37-56."""
    o = np.concatenate([p["observations"] for p in paths])
    o = np.clip(o, -10, 10) / 10.0
    if o.ndim > 2:
        o = o.reshape(o.shape[0], -1)
    N, n = o.shape
    F = np.ones((N, n + 4))
    F[:, :n] = o
    row = 0
    for p in paths:
        T = len(p["rewards"])
        al = np.arange(T) / 1000.0
        for j in range(4):
            F[row:row + T, -4 + j] = al ** (j + 1)
        row += T
    return F


class MLPBaseline:
    """(n+4)-128-128-1 ReLU value net, Adam + MSE; mjrl/baselines/mlp_baseline.py:15-115.

    `perm_source` lets a test inject the minibatch permutations so the GPU
    path can be driven with the very same index order (the reference draws
    them from numpy's global stream, :84)."""

    def __init__(self, env_spec, obs_dim=None, learn_rate=1e-3, reg_coef=0.0,
                 batch_size=64, epochs=1, perm_source=None):
        self.n = obs_dim if obs_dim is not None else env_spec.observation_dim
        self.batch_size, self.epochs, self.reg_coef = batch_size, epochs, reg_coef
        self.model = torch.nn.Sequential(
            torch.nn.Linear(self.n + 4, 128), torch.nn.ReLU(),
            torch.nn.Linear(128, 128), torch.nn.ReLU(),
            torch.nn.Linear(128, 1))
        self.optimizer = torch.optim.Adam(self.model.parameters(), lr=learn_rate,
                                          weight_decay=reg_coef)
        self.perm_source = perm_source or (lambda n: np.random.permutation(n))

    def _rel_err(self, X, y):
        with torch.no_grad():
            pred = self.model(X).numpy().ravel()
        e = y.numpy().ravel() - pred
        return np.sum(e ** 2) / (np.sum(y.numpy() ** 2) + 1e-8)

    def fit(self, paths, return_errors=False):
        X = torch.from_numpy(mlp_baseline_features(paths).astype("float32"))
        y = torch.from_numpy(
            np.concatenate([p["returns"] for p in paths]).reshape(-1, 1).astype("float32"))
        N = y.shape[0]
        if return_errors:
            before = self._rel_err(X, y)
        for _ in range(self.epochs):
            perm = self.perm_source(N)
            for mb in range(int(N / self.batch_size) - 1):  # :85 -- last batch skipped
                idx = torch.from_numpy(
                    np.asarray(perm[mb * self.batch_size:(mb + 1) * self.batch_size], dtype=np.int64))
                self.optimizer.zero_grad()
                loss = torch.nn.functional.mse_loss(self.model(X[idx]), y[idx])
                loss.backward()
                self.optimizer.step()
        if return_errors:
            return before, self._rel_err(X, y)

    def predict(self, path):
        X = torch.from_numpy(mlp_baseline_features([path]).astype("float32"))
        with torch.no_grad():
            return self.model(X).numpy().ravel()


# --------------------------------------------------------------------------
# (c) Gaussian policies (two-hidden-layer tanh MLP, or affine), fp32 torch
# --------------------------------------------------------------------------


def param_shapes(n, m, hidden):
    """Flat-vector layout: fc0.weight (H1,n), fc0.bias, fc1.weight (H2,H1), fc1.bias,
    fc2.weight (m,H2), fc2.bias, log_std (m) -- row-major (out,in).

    mjrl/policies/gaussian_mlp.py:38,51-53,156-158; for the affine policy only
    fc0 (m,n) + bias + log_std, mjrl/policies/gaussian_linear.py:36,153."""
    if hidden:
        h1, h2 = hidden
        return [(h1, n), (h1,), (h2, h1), (h2,), (m, h2), (m,), (m,)]
    return [(m, n), (m,), (m,)]


def unflatten(flat, shapes, requires_grad=False):
    out, k = [], 0
    for s in shapes:
        sz = int(np.prod(s))
        t = torch.from_numpy(np.asarray(flat[k:k + sz]).reshape(s)).float()
        out.append(t.clone().requires_grad_(requires_grad))
        k += sz
    return out


def flatten(tensors):
    return np.concatenate([t.contiguous().view(-1).detach().numpy() for t in tensors])


class GaussianPolicy:
    """Holds the `new` (trainable) and `old` parameter sets of mjrl's MLP /
    LinearPolicy and evaluates them exactly as the reference does.

    mjrl/policies/gaussian_mlp.py:8-140 and gaussian_linear.py:8-138."""

    def __init__(self, n, m, hidden, flat_params, min_log_std=-3.0,
                 in_shift=None, in_scale=None, out_shift=None, out_scale=None):
        self.n, self.m, self.hidden = n, m, tuple(hidden) if hidden else ()
        self.shapes = param_shapes(n, m, self.hidden)
        self.sizes = [int(np.prod(s)) for s in self.shapes]
        self.d = int(sum(self.sizes))
        self.min_log_std = min_log_std
        f32 = lambda v, dflt, k: (torch.from_numpy(np.float32(v)) if v is not None
                                  else dflt(k))
        self.in_shift = f32(in_shift, torch.zeros, n)
        self.in_scale = f32(in_scale, torch.ones, n)
        self.out_shift = f32(out_shift, torch.zeros, m)
        self.out_scale = f32(out_scale, torch.ones, m)
        self.new = None
        self.old = None
        self.set_param_values(np.asarray(flat_params, dtype=np.float32))
        self.aliased = False   # a freshly constructed reference policy clones old from new (:42-46)

    # gaussian_mlp.py:61-64
    def get_param_values(self):
        return flatten(self.new).copy()

    # gaussian_mlp.py:66-88 -- fp32 cast, then clamp log_std at min_log_std
    def set_param_values(self, flat, set_new=True, set_old=True):
        # Reference quirk (gaussian_mlp.py:66-88): `param.data = torch.from_numpy(vals).float()` does
        # NOT copy when `new_params` is already float32, so after a call with set_new=set_old=True on
        # a float32 array the new and old tensors share ONE buffer.  It only matters to optimisers
        # that update the new parameters in place (PPO, ppo_clip.py:96): from then on old == new.
        if set_new and set_old:
            self.aliased = np.asarray(flat).dtype == np.float32
        elif set_new or set_old:
            self.aliased = False
        if set_new:
            self.new = unflatten(flat, self.shapes, requires_grad=True)
            self.new[-1] = torch.clamp(self.new[-1].detach(), self.min_log_std).requires_grad_(True)
            self.log_std_val = np.float64(self.new[-1].detach().numpy().ravel())
        if set_old:
            self.old = unflatten(flat, self.shapes)
            self.old[-1] = torch.clamp(self.old[-1], self.min_log_std)

    def sync_aliased_old(self):
        """What the shared buffer does in the reference once `aliased`: the old NETWORK tensors equal
        the new ones.  log_std is excluded: the clamp at :75-76 / :87-88 re-assigns `.data` to a fresh
        tensor on both sides, so old_log_std keeps the value it had when it was set."""
        self.old = [t.detach().clone() for t in self.new[:-1]] + [self.old[-1]]

    def mean(self, obs, params):
        """MuNet.forward gaussian_mlp.py:176-182 / LinearModel.forward gaussian_linear.py:171-175."""
        x = (obs - self.in_shift) / (self.in_scale + 1e-8)
        if self.hidden:
            W0, b0, W1, b1, W2, b2 = params[:6]
            x = torch.tanh(torch.nn.functional.linear(x, W0, b0))
            x = torch.tanh(torch.nn.functional.linear(x, W1, b1))
            x = torch.nn.functional.linear(x, W2, b2)
        else:
            x = torch.nn.functional.linear(x, params[0], params[1])
        return x * self.out_scale + self.out_shift

    def mean_LL(self, obs, act, params):
        """gaussian_mlp.py:100-110: fp32 casts, z=(a-mu)/exp(log_std)."""
        o = torch.from_numpy(obs).float()
        a = torch.from_numpy(act).float()
        mu = self.mean(o, params)
        log_std = params[-1]
        z = (a - mu) / torch.exp(log_std)
        LL = -0.5 * torch.sum(z ** 2, dim=1) + -torch.sum(log_std) + -0.5 * self.m * np.log(2 * np.pi)
        return mu, LL

    def dist_info(self, obs, act, which):
        params = self.new if which == "new" else self.old
        mu, LL = self.mean_LL(obs, act, params)
        return LL, mu, params[-1]

    @staticmethod
    def mean_kl(new_info, old_info):
        """KL(old || new) averaged over timesteps; gaussian_mlp.py:130-140."""
        _, mu_n, ls_n = new_info
        _, mu_o, ls_o = old_info
        sd_o, sd_n = torch.exp(ls_o), torch.exp(ls_n)
        num = (mu_o - mu_n) ** 2 + sd_o ** 2 - sd_n ** 2
        den = 2 * sd_n ** 2 + 1e-8
        return torch.mean(torch.sum(num / den + ls_n - ls_o, dim=1))

    def get_action(self, observation):
        """gaussian_mlp.py:92-98; noise from numpy's GLOBAL stream."""
        o = torch.from_numpy(np.float32(observation.reshape(1, -1)))
        with torch.no_grad():
            mean = self.mean(o, self.new).numpy().ravel()
        noise = np.exp(self.log_std_val) * np.random.randn(self.m)
        return [mean + noise, {"mean": mean, "log_std": self.log_std_val, "evaluation": mean}]


def init_policy_params(n, m, hidden, seed, init_log_std=0.0):
    """Initial flat parameter vector with the reference's RNG consumption order.

    mjrl/policies/gaussian_mlp.py:27-57: seed torch+numpy, build the model
    (torch default nn.Linear init), scale the LAST layer's weight and bias by
    1e-2, log_std = init_log_std; then a second network is built for `old`
    (its draws are discarded, :42-46) and `randn(n)` is drawn (:57)."""
    if seed is not None:
        torch.manual_seed(seed)
        np.random.seed(seed)
    dims = [n] + list(hidden) + [m]
    layers = [torch.nn.Linear(dims[i], dims[i + 1]) for i in range(len(dims) - 1)]
    with torch.no_grad():
        layers[-1].weight.mul_(1e-2)
        layers[-1].bias.mul_(1e-2)
    _ = [torch.nn.Linear(dims[i], dims[i + 1]) for i in range(len(dims) - 1)]  # old_model draws
    _ = torch.randn(n)  # obs_var placeholder draw
    parts = []
    for l in layers:
        parts += [l.weight.detach().numpy().ravel(), l.bias.detach().numpy().ravel()]
    parts.append(np.full(m, init_log_std, dtype=np.float32))
    return np.concatenate(parts).astype(np.float32)


# --------------------------------------------------------------------------
# (c)(d) surrogate, gradient, Hessian-of-KL product, CG
# --------------------------------------------------------------------------


def cpi_surrogate(pol, obs, act, adv):
    """mean(exp(LL_new - LL_old) * adv); mjrl/algos/batch_reinforce.py:37-43."""
    a = torch.from_numpy(adv).float()
    LL_o = pol.dist_info(obs, act, "old")[0]
    LL_n = pol.dist_info(obs, act, "new")[0]
    return torch.mean(torch.exp(LL_n - LL_o) * a)


def kl_old_new(pol, obs, act):
    """mjrl/algos/batch_reinforce.py:45-49."""
    return GaussianPolicy.mean_kl(pol.dist_info(obs, act, "new"), pol.dist_info(obs, act, "old"))


def flat_vpg(pol, obs, act, adv):
    """Autograd gradient of the surrogate, flattened fp32; batch_reinforce.py:51-55."""
    g = torch.autograd.grad(cpi_surrogate(pol, obs, act, adv), pol.new)
    return flatten(g)


def hvp(pol, obs, act, vector, regu_coef, sample_frac=1.0):
    """Hessian of mean_kl(new, old) times `vector` by double backprop, plus
    regu_coef*vector; mjrl/algos/npg_cg.py:55-74."""
    vec = torch.from_numpy(vector).float()
    if sample_frac is not None and sample_frac < 0.99:  # :58-62
        N = obs.shape[0]
        idx = np.random.choice(N, size=int(sample_frac * N))
        obs, act = obs[idx], act[idx]
    kl = kl_old_new(pol, obs, act)
    g1 = torch.autograd.grad(kl, pol.new, create_graph=True)
    flat_g = torch.cat([g.contiguous().view(-1) for g in g1])
    h = torch.sum(flat_g * vec)
    g2 = torch.autograd.grad(h, pol.new)
    return flatten(g2) + regu_coef * vector


def cg_solve(f_Ax, b, x_0=None, cg_iters=10, residual_tol=1e-10, trace=None):
    """Textbook CG from x=0 (x_0 is IGNORED); mjrl/utils/cg_solve.py:3-22.

    dtype follows `b` (fp32 in the reference's use).  `trace`, if a list,
    receives a copy of x after every iteration (test instrumentation only)."""
    x = np.zeros_like(b)
    r = b.copy()
    p = r.copy()
    rdotr = r.dot(r)
    for _ in range(cg_iters):
        z = f_Ax(p)
        v = rdotr / p.dot(z)
        x += v * p
        r -= v * z
        new_rdotr = r.dot(r)
        mu = new_rdotr / rdotr
        p = r + mu * p
        rdotr = new_rdotr
        if trace is not None:
            trace.append(x.copy())
        if rdotr < residual_tol:
            break
    return x


# --------------------------------------------------------------------------
# (d) one update of each algorithm (the body of train_from_paths)
# --------------------------------------------------------------------------


def concat_paths(paths):
    """obs/act/adv concatenated over paths, advantages whitened with eps 1e-6.

    mjrl/algos/npg_cg.py:87-91 (same lines in batch_reinforce.py:109-113,
    trpo.py:57-61, dapg.py:57-60, ppo_clip.py:60-64)."""
    obs = np.concatenate([p["observations"] for p in paths])
    act = np.concatenate([p["actions"] for p in paths])
    adv = np.concatenate([p["advantages"] for p in paths])
    adv = (adv - np.mean(adv)) / (np.std(adv) + 1e-6)
    return obs, act, adv


def path_return_stats(paths):
    """[mean, std, min, max] of per-path reward sums; npg_cg.py:97-102."""
    R = [sum(p["rewards"]) for p in paths]
    return [np.mean(R), np.std(R), np.amin(R), np.amax(R)]


def _scalar(t):
    return t.detach().numpy().ravel()[0]


def reinforce_update(pol, paths, learn_rate):
    """mjrl/algos/batch_reinforce.py:106-145."""
    obs, act, adv = concat_paths(paths)
    out = {"stats": path_return_stats(paths)}
    out["surr_before"] = _scalar(cpi_surrogate(pol, obs, act, adv))
    g = out["vpg_grad"] = flat_vpg(pol, obs, act, adv)
    new = pol.get_param_values() + learn_rate * g
    pol.set_param_values(new, set_new=True, set_old=False)
    out["surr_after"] = _scalar(cpi_surrogate(pol, obs, act, adv))
    out["kl_dist"] = _scalar(kl_old_new(pol, obs, act))
    pol.set_param_values(new, set_new=True, set_old=True)
    out["new_params"] = new
    return out


def npg_update(pol, paths, normalized_step_size=0.01, const_learn_rate=None,
               cg_iters=10, damping=1e-4, hvp_sample_frac=1.0, cg_trace=None):
    """mjrl/algos/npg_cg.py:84-144."""
    obs, act, adv = concat_paths(paths)
    out = {"stats": path_return_stats(paths), "adv_white": adv}
    out["surr_before"] = _scalar(cpi_surrogate(pol, obs, act, adv))
    g = out["vpg_grad"] = flat_vpg(pol, obs, act, adv)
    op = lambda v: hvp(pol, obs, act, v, damping, hvp_sample_frac)
    x = out["npg_grad"] = cg_solve(op, g, x_0=g.copy(), cg_iters=cg_iters, trace=cg_trace)
    if const_learn_rate is not None:  # :130-132
        alpha = const_learn_rate
        delta = (alpha ** 2) * np.dot(g.T, x)
    else:  # :133-135
        delta = normalized_step_size
        alpha = np.sqrt(np.abs(normalized_step_size / (np.dot(g.T, x) + 1e-20)))
    new = pol.get_param_values() + alpha * x
    pol.set_param_values(new, set_new=True, set_old=False)
    out["surr_after"] = _scalar(cpi_surrogate(pol, obs, act, adv))
    out["kl_dist"] = _scalar(kl_old_new(pol, obs, act))
    pol.set_param_values(new, set_new=True, set_old=True)
    out.update(alpha=alpha, delta=delta, new_params=new)
    return out


def trpo_line_search(pol, obs, act, adv, curr, direction, alpha, kl_target, max_trials=100):
    """Backtracking loop of mjrl/algos/trpo.py:105-124.

    Returns (alpha_final, trials=[(kl, surr, accepted)], kl_final, surr_final,
    new_params).  The policy is left committed (new == old == new_params)."""
    trials = []
    for k in range(max_trials):
        new = curr + alpha * direction
        pol.set_param_values(new, set_new=True, set_old=False)
        kl = _scalar(kl_old_new(pol, obs, act))
        surr = _scalar(cpi_surrogate(pol, obs, act, adv))
        ok = bool(kl < kl_target)
        trials.append((kl, surr, ok))
        if ok:
            break
        alpha = 0.9 * alpha
        if k == max_trials - 1:
            alpha = 0.0
    new = curr + alpha * direction
    pol.set_param_values(new, set_new=True, set_old=False)
    kl = _scalar(kl_old_new(pol, obs, act))
    surr = _scalar(cpi_surrogate(pol, obs, act, adv))
    pol.set_param_values(new, set_new=True, set_old=True)
    return alpha, trials, kl, surr, new


def trpo_update(pol, paths, kl_dist=0.01, cg_iters=10, damping=1e-4, hvp_sample_frac=1.0,
                direction_override=None):
    """mjrl/algos/trpo.py:54-124.  `direction_override` (test hook, SURVEY H2
    stage 1) feeds a given npg direction into the step-size rule + line search."""
    obs, act, adv = concat_paths(paths)
    out = {"stats": path_return_stats(paths)}
    out["surr_before"] = _scalar(cpi_surrogate(pol, obs, act, adv))
    g = out["vpg_grad"] = flat_vpg(pol, obs, act, adv)
    if direction_override is None:
        op = lambda v: hvp(pol, obs, act, v, damping, hvp_sample_frac)
        x = cg_solve(op, g, x_0=g.copy(), cg_iters=cg_iters)
    else:
        x = np.asarray(direction_override, dtype=np.float32)
    out["npg_grad"] = x
    delta = 2.0 * kl_dist
    alpha0 = np.sqrt(np.abs(delta / (np.dot(g.T, x) + 1e-20)))
    curr = pol.get_param_values()
    alpha, trials, kl, surr, new = trpo_line_search(pol, obs, act, adv, curr, x, alpha0, kl_dist)
    out.update(alpha0=alpha0, alpha=alpha, delta=delta, trials=trials, kl_dist=kl,
               surr_after=surr, new_params=new)
    return out


def dapg_update(pol, paths, demo_paths, kl_dist, lam_0=1.0, lam_1=0.95, iter_count=0.0,
                cg_iters=10, damping=1e-4, hvp_sample_frac=1.0):
    """mjrl/algos/dapg.py:54-121.  Returns the incremented iter_count in out."""
    obs, act, adv = concat_paths(paths)
    if demo_paths is not None and lam_0 > 0.0:  # :62-70
        d_obs = np.concatenate([p["observations"] for p in demo_paths])
        d_act = np.concatenate([p["actions"] for p in demo_paths])
        d_adv = lam_0 * (lam_1 ** iter_count) * np.ones(d_obs.shape[0])
        iter_count += 1
        all_obs = np.concatenate([obs, d_obs])
        all_act = np.concatenate([act, d_act])
        all_adv = 1e-2 * np.concatenate([adv / (np.std(adv) + 1e-8), d_adv])
    else:
        all_obs, all_act, all_adv = obs, act, adv
    out = {"stats": path_return_stats(paths), "iter_count": iter_count}
    out["surr_before"] = _scalar(cpi_surrogate(pol, obs, act, adv))
    coef = all_adv.shape[0] / adv.shape[0]
    g = out["vpg_grad"] = coef * flat_vpg(pol, all_obs, all_act, all_adv)
    op = lambda v: hvp(pol, obs, act, v, damping, hvp_sample_frac)
    x = out["npg_grad"] = cg_solve(op, g, x_0=g.copy(), cg_iters=cg_iters)
    delta = 2.0 * kl_dist
    alpha = np.sqrt(np.abs(delta / (np.dot(g.T, x) + 1e-20)))
    new = pol.get_param_values() + alpha * x
    pol.set_param_values(new, set_new=True, set_old=False)
    out["surr_after"] = _scalar(cpi_surrogate(pol, obs, act, adv))
    out["kl_dist"] = _scalar(kl_old_new(pol, obs, act))
    pol.set_param_values(new, set_new=True, set_old=True)
    out.update(alpha=alpha, delta=delta, new_params=new)
    return out


def ppo_surrogate(pol, obs, act, adv, clip_coef):
    """mjrl/algos/ppo_clip.py:47-54."""
    adv_var = torch.from_numpy(np.asarray(adv)).float()
    old = pol.dist_info(obs, act, "old")
    new = pol.dist_info(obs, act, "new")
    LR = torch.exp(new[0] - old[0])
    LR_clip = torch.clamp(LR, min=1 - clip_coef, max=1 + clip_coef)
    return torch.mean(torch.min(LR * adv_var, LR_clip * adv_var))


class PPOState:
    """The optimizer the reference builds once in PPO.__init__ (ppo_clip.py:45): Adam over the
    policy's trainable tensors, state kept across iterations."""

    def __init__(self, pol, learn_rate=3e-4):
        self.pol = pol
        self.optimizer = torch.optim.Adam(pol.new, lr=learn_rate)


def ppo_update(pol, paths, state=None, clip_coef=0.2, epochs=10, mb_size=64, learn_rate=3e-4,
               choice_source=None):
    """mjrl/algos/ppo_clip.py:57-120.  `choice_source(num_samples, mb_size)` lets a test inject the
    minibatch indices; default np.random.choice on the global stream like :89."""
    obs, act, adv = concat_paths(paths)
    out = {"stats": path_return_stats(paths)}
    out["surr_before"] = _scalar(cpi_surrogate(pol, obs, act, adv))
    state = state or PPOState(pol, learn_rate)
    # the optimizer must hold the CURRENT trainable tensors (set_param_values re-creates them here,
    # the reference mutates .data in place); moments are carried over by position
    if any(a is not b for a, b in zip(state.optimizer.param_groups[0]["params"], pol.new)):
        old_params = state.optimizer.param_groups[0]["params"]
        st = [state.optimizer.state.get(p, None) for p in old_params]
        state.optimizer = torch.optim.Adam(pol.new, lr=state.optimizer.param_groups[0]["lr"])
        for p, s_ in zip(pol.new, st):
            if s_ is not None:
                state.optimizer.state[p] = s_
    choice = choice_source or (lambda n, k: np.random.choice(n, size=k))
    num_samples = obs.shape[0]
    for _ in range(epochs):
        for _mb in range(int(num_samples / mb_size)):
            idx = choice(num_samples, mb_size)
            state.optimizer.zero_grad()
            if pol.aliased:   # old network tensors share new's storage: they track every Adam step
                pol.sync_aliased_old()
            loss = -ppo_surrogate(pol, obs[idx], act[idx], adv[idx], clip_coef)
            loss.backward()
            state.optimizer.step()
    params_after_opt = pol.get_param_values()
    if pol.aliased:
        pol.sync_aliased_old()
    out["surr_after"] = _scalar(cpi_surrogate(pol, obs, act, adv))
    out["kl_dist"] = _scalar(kl_old_new(pol, obs, act))
    pol.set_param_values(params_after_opt, set_new=True, set_old=True)
    out["new_params"] = pol.get_param_values()
    out["state"] = state
    return out


def bc_transformations(expert_paths):
    """mjrl/algos/behavior_cloning.py:25-29 (same lines in behavior_cloning_2.py): fp64 column
    mean / std of the expert observations and actions, used as the policy's in/out transforms."""
    observations = np.concatenate([p["observations"] for p in expert_paths])
    actions = np.concatenate([p["actions"] for p in expert_paths])
    return dict(in_shift=np.mean(observations, axis=0), in_scale=np.std(observations, axis=0),
                out_shift=np.mean(actions, axis=0), out_scale=np.std(actions, axis=0))


class BCState:
    """What BC.__init__ does to the policy and the optimizer it builds once
    (behavior_cloning.py:24-36; behavior_cloning_2.py:24-45).  mode "mle" = behavior_cloning.py
    (Adam over trainable_params, loss -mean LL); mode "mse" = behavior_cloning_2.py (log_std set
    to log(out_scale + 1e-12) through set_param_values, Adam over the NETWORK only, MSE loss)."""

    def __init__(self, pol, expert_paths, mode="mle", lr=1e-3):
        assert mode in ("mle", "mse")
        self.pol, self.mode = pol, mode
        tr = bc_transformations(expert_paths)
        for k, v in tr.items():   # MuNet.set_transformations gaussian_mlp.py:159-174: fp32 cast
            setattr(pol, k, torch.from_numpy(np.float32(v)))
        if mode == "mse":
            params = pol.get_param_values()
            params[-pol.m:] = np.log(tr["out_scale"] + 1e-12)
            pol.set_param_values(params)
        self.optimizer = torch.optim.Adam(pol.new if mode == "mle" else pol.new[:-1], lr=lr)


def bc_loss(pol, obs, act, mode):
    """behavior_cloning.py:38-41 (-mean LL of the new parameters) / behavior_cloning_2.py:50-54
    (torch MSELoss between model(obs) and the expert actions, both cast to fp32)."""
    if mode == "mle":
        return -torch.mean(pol.dist_info(obs, act, "new")[0])
    o = torch.from_numpy(obs).float()
    a = torch.from_numpy(act).float()
    return torch.nn.functional.mse_loss(pol.mean(o, pol.new), a)


def bc_train(state, expert_paths, epochs=5, mb_size=64, choice_source=None):
    """BC.train, behavior_cloning.py:43-68 / behavior_cloning_2.py:56-77: per epoch one full-data
    loss (logged), then int(N/mb) Adam steps on with-replacement minibatches drawn by
    np.random.choice from the GLOBAL stream; finally set_param_values(new, True, True)."""
    pol, mode = state.pol, state.mode
    obs = np.concatenate([p["observations"] for p in expert_paths])
    act = np.concatenate([p["actions"] for p in expert_paths])
    choice = choice_source or (lambda n, k: np.random.choice(n, size=k))
    num_samples = obs.shape[0]
    losses = []
    for _ in range(epochs):
        losses.append(_scalar(bc_loss(pol, obs, act, mode)))
        for _mb in range(int(num_samples / mb_size)):
            idx = choice(num_samples, mb_size)
            state.optimizer.zero_grad()
            loss = bc_loss(pol, obs[idx], act[idx], mode)
            loss.backward()
            state.optimizer.step()
    params_after_opt = pol.get_param_values()
    # set_param_values re-creates the tensor list here (the reference re-points .data of the same
    # leaves): keep the optimizer on the live tensors, moments carried over by position
    old_leaves = state.optimizer.param_groups[0]["params"]
    st = [state.optimizer.state.get(p, None) for p in old_leaves]
    pol.set_param_values(params_after_opt, set_new=True, set_old=True)
    leaves = pol.new if mode == "mle" else pol.new[:-1]
    state.optimizer = torch.optim.Adam(leaves, lr=state.optimizer.param_groups[0]["lr"])
    for p, s_ in zip(leaves, st):
        if s_ is not None:
            state.optimizer.state[p] = s_
    losses.append(_scalar(bc_loss(pol, obs, act, mode)))
    return {"losses": np.asarray(losses), "new_params": pol.get_param_values()}


def full_npg_iteration(pol, baseline, paths, gamma, gae_lambda, **npg_kw):
    """What BatchREINFORCE.train_step does after sampling, batch_reinforce.py:86-101:
    returns, advantages (with the PREVIOUS fit), policy update, then baseline.fit."""
    compute_returns(paths, gamma)
    compute_advantages(paths, baseline, gamma, gae_lambda)
    out = npg_update(pol, paths, **npg_kw)
    baseline.fit(paths)
    return out


# --------------------------------------------------------------------------
# fp64 analytic "truth" leg: same maths, no autograd, float64
# --------------------------------------------------------------------------


class TruthPolicy:
    """fp64 analytic forward / gradient / Gauss-Newton product for the same
    network (gaussian_mlp.py:100-140,176-182).  At new == old the Hessian of
    mean_kl equals J^T diag(2/(2 sigma^2+1e-8)) J / N on the network block and
    diag(4s(2s-eps)/(2s+eps)^2), s = sigma^2, on the log_std block (npg_cg.py:55-74)."""

    def __init__(self, pol: GaussianPolicy):
        self.n, self.m, self.hidden = pol.n, pol.m, pol.hidden
        self.shapes, self.sizes = pol.shapes, pol.sizes
        self.in_shift = pol.in_shift.numpy().astype(np.float64)
        self.in_scale = pol.in_scale.numpy().astype(np.float64)
        self.out_shift = pol.out_shift.numpy().astype(np.float64)
        self.out_scale = pol.out_scale.numpy().astype(np.float64)
        self.params = [p.detach().numpy().astype(np.float64) for p in pol.new]
        self.old = [p.detach().numpy().astype(np.float64) for p in pol.old]

    def _split(self, flat):
        out, k = [], 0
        for s, sz in zip(self.shapes, self.sizes):
            out.append(np.asarray(flat[k:k + sz], dtype=np.float64).reshape(s))
            k += sz
        return out

    def forward(self, obs, params=None):
        P = self.params if params is None else params
        x = (np.float32(obs).astype(np.float64) - self.in_shift) / (self.in_scale + 1e-8)
        if self.hidden:
            h1 = np.tanh(x @ P[0].T + P[1])
            h2 = np.tanh(h1 @ P[2].T + P[3])
            y = h2 @ P[4].T + P[5]
        else:
            h1 = h2 = None
            y = x @ P[0].T + P[1]
        return x, h1, h2, y * self.out_scale + self.out_shift

    def log_lik(self, obs, act, params=None):
        P = self.params if params is None else params
        mu = self.forward(obs, P)[3]
        a = np.float32(act).astype(np.float64)
        z = (a - mu) / np.exp(P[-1])
        return -0.5 * np.sum(z ** 2, axis=1) - np.sum(P[-1]) - 0.5 * self.m * np.log(2 * np.pi), mu

    def _backprop(self, x, h1, h2, dy, weights):
        """sum_t of parameter gradients given dL/dy [N,m] (y = pre-out_scale)."""
        if self.hidden:
            gW2, gb2 = dy.T @ h2, dy.sum(0)
            dz2 = (dy @ weights[4]) * (1 - h2 ** 2)
            gW1, gb1 = dz2.T @ h1, dz2.sum(0)
            dz1 = (dz2 @ weights[2]) * (1 - h1 ** 2)
            gW0, gb0 = dz1.T @ x, dz1.sum(0)
            return [gW0, gb0, gW1, gb1, gW2, gb2]
        return [dy.T @ x, dy.sum(0)]

    def vpg(self, obs, act, adv):
        """d/dtheta mean(exp(LL_new-LL_old)*adv), analytic (batch_reinforce.py:37-55)."""
        N = obs.shape[0]
        LLn, mu = self.log_lik(obs, act, self.params)
        LLo, _ = self.log_lik(obs, act, self.old)
        w = np.exp(LLn - LLo) * np.float32(adv).astype(np.float64) / N
        x, h1, h2, _ = self.forward(obs)
        a = np.float32(act).astype(np.float64)
        sig2 = np.exp(2 * self.params[-1])
        dmu = w[:, None] * (a - mu) / sig2
        g = self._backprop(x, h1, h2, dmu * self.out_scale, self.params)
        gls = (w[:, None] * (((a - mu) ** 2) / sig2 - 1.0)).sum(0)
        return np.concatenate([t.ravel() for t in g] + [gls])

    def fvp(self, obs, vec, damping=0.0):
        N = obs.shape[0]
        V = self._split(vec)
        P = self.params
        x, h1, h2, _ = self.forward(obs)
        if self.hidden:
            dz1 = x @ V[0].T + V[1]
            dh1 = (1 - h1 ** 2) * dz1
            dz2 = h1 @ V[2].T + dh1 @ P[2].T + V[3]
            dh2 = (1 - h2 ** 2) * dz2
            dy = h2 @ V[4].T + dh2 @ P[4].T + V[5]
        else:
            dy = x @ V[0].T + V[1]
        sig2 = np.exp(2 * P[-1])
        M = 2.0 / (2.0 * sig2 + 1e-8)
        u = dy * self.out_scale * M * self.out_scale / N
        g = self._backprop(x, h1, h2, u, P)
        eps = 1e-8
        hls = 4 * sig2 * (2 * sig2 - eps) / (2 * sig2 + eps) ** 2
        out = np.concatenate([t.ravel() for t in g] + [hls * V[-1]])
        return out + damping * np.asarray(vec, dtype=np.float64)

    def surrogate_kl(self, obs, act, adv):
        LLn, mun = self.log_lik(obs, act, self.params)
        LLo, muo = self.log_lik(obs, act, self.old)
        surr = np.mean(np.exp(LLn - LLo) * np.float32(adv).astype(np.float64))
        so, sn = np.exp(self.old[-1]), np.exp(self.params[-1])
        kl = np.mean(np.sum(((muo - mun) ** 2 + so ** 2 - sn ** 2) / (2 * sn ** 2 + 1e-8)
                            + self.params[-1] - self.old[-1], axis=1))
        return surr, kl


def truth_cg(tp: TruthPolicy, obs, b, cg_iters=10, damping=1e-4, trace=None):
    """fp64 CG on the fp64 analytic operator (the x* of SURVEY.md H1)."""
    return cg_solve(lambda v: tp.fvp(obs, v, damping), np.asarray(b, dtype=np.float64),
                    cg_iters=cg_iters, trace=trace)
