"""Synthetic trajectory batches of the shapes BASELINE.json names.

One generator feeds the oracle, the CPU baseline and the CUDA path alike
(SURVEY.md section 8(d)): fixed seeds, fp32-representable observations
N(0, 3^2) and rewards N(1, 1) stored as fp64, on-policy actions, 90 % full
horizon non-terminated paths and 10 % early-terminated paths.

The output follows the `paths` contract of the reference's sampler
(mjrl/samplers/base_sampler.py:76-83): a list of dicts with `observations`
[T, n] fp64, `actions` [T, m] fp64, `rewards` [T] fp64 and `terminated`.
"""
import numpy as np

# name -> (obs_dim, act_dim, hidden_sizes, horizon, total timesteps)
SHAPES = {
    "C1-pointmass": (6, 2, (), 25, 1_000),
    "C2-swimmer": (12, 4, (32, 32), 500, 10_000),
    "C3-hopper": (11, 3, (64, 64), 1000, 1_000_000),
    "C4-ant": (111, 8, (64, 64), 1000, 4_000_000),
    "C5-adroit": (45, 24, (64, 64), 200, 2_000_000),
    "C5-adroit-32": (45, 24, (32, 32), 200, 2_000_000),
}


def path_lengths(N, horizon, rng):
    """90 % full-horizon (terminated=False), 10 % UniformInt[2, horizon] (terminated=True);
    the last path is truncated so the lengths sum to N exactly."""
    lens, term, total = [], [], 0
    while total < N:
        if rng.random() < 0.9 or horizon <= 2:
            T, done = horizon, False
        else:
            T, done = int(rng.integers(2, horizon + 1)), True
        T = min(T, N - total)
        lens.append(T)
        term.append(done)
        total += T
    return np.asarray(lens, dtype=np.int64), np.asarray(term, dtype=bool)


def mean_forward_f32(obs, flat, n, m, hidden, in_shift=None, in_scale=None,
                     out_shift=None, out_scale=None):
    """fp32 mean of the Gaussian policy, used ONLY to draw on-policy synthetic
    actions (same maths as mjrl/policies/gaussian_mlp.py:176-182)."""
    f = np.float32
    x = obs.astype(f)
    if in_shift is not None:
        x = (x - f(in_shift)) / (f(in_scale) + f(1e-8))
    k = 0
    dims = [n] + list(hidden) + [m]
    for li in range(len(dims) - 1):
        o, i = dims[li + 1], dims[li]
        W = flat[k:k + o * i].reshape(o, i).astype(f); k += o * i
        b = flat[k:k + o].astype(f); k += o
        x = x @ W.T + b
        if li < len(dims) - 2:
            x = np.tanh(x)
    if out_scale is not None:
        x = x * f(out_scale) + f(out_shift)
    return x.astype(f), flat[k:k + m].astype(f)


def make_flat_batch(N, n, m, horizon, seed, flat_params, hidden,
                    transforms=None, chunk=1 << 18):
    """Flat buffers: obs [N,n] fp64, act [N,m] fp64, rew [N] fp64,
    path_offsets int32 [P+1], terminated uint8 [P]."""
    rng = np.random.default_rng(seed)
    lens, term = path_lengths(N, horizon, rng)
    offsets = np.zeros(len(lens) + 1, dtype=np.int32)
    offsets[1:] = np.cumsum(lens)
    obs = (3.0 * rng.standard_normal((N, n), dtype=np.float32)).astype(np.float64)
    rew = (1.0 + rng.standard_normal(N, dtype=np.float32)).astype(np.float64)
    act = np.empty((N, m), dtype=np.float64)
    tr = transforms or {}
    for s in range(0, N, chunk):
        e = min(N, s + chunk)
        mu, log_std = mean_forward_f32(obs[s:e], flat_params, n, m, hidden, **tr)
        eps = rng.standard_normal((e - s, m), dtype=np.float32)
        ls = np.maximum(log_std, np.float32(-3.0))
        act[s:e] = (mu + np.exp(ls) * eps).astype(np.float32).astype(np.float64)
    return dict(observations=obs, actions=act, rewards=rew, path_offsets=offsets,
                terminated=term.astype(np.uint8))


def flat_to_paths(batch):
    """Views of the flat buffers as the reference's list-of-dicts `paths`."""
    off = batch["path_offsets"]
    paths = []
    for i in range(len(off) - 1):
        s, e = int(off[i]), int(off[i + 1])
        paths.append(dict(observations=batch["observations"][s:e],
                          actions=batch["actions"][s:e],
                          rewards=batch["rewards"][s:e],
                          terminated=bool(batch["terminated"][i])))
    return paths


def mid_training_variant(flat_params, n, m, seed):
    """'Mid-training' policy of SURVEY.md 8(d): perturbed parameters, log_std
    around -0.5, and non-identity input/output transforms."""
    rng = np.random.default_rng(seed + 7919)
    p = flat_params.astype(np.float32) + np.float32(0.1) * rng.standard_normal(
        flat_params.shape, dtype=np.float32)
    p[-m:] = np.float32(-0.5) + np.float32(0.1) * rng.standard_normal(m, dtype=np.float32)
    tr = dict(in_shift=rng.standard_normal(n).astype(np.float32),
              in_scale=rng.uniform(0.5, 1.5, n).astype(np.float32),
              out_shift=rng.standard_normal(m).astype(np.float32),
              out_scale=rng.uniform(0.5, 1.5, m).astype(np.float32))
    return p.astype(np.float32), tr
