"""GPU parity: segmented scans and LinearBaseline vs the oracle / reference goldens.

Tolerance (north_star): returns, advantages and baseline predictions within 1e-6 relative
(fp64).  We assert 1e-9 on the scans (they are fp64 re-associations of the same sums)."""
import copy

import numpy as np
import pytest
import torch

from conftest import batch_of, rel_l2
from mjrl_b200.utils import synthetic as syn

pytestmark = pytest.mark.gpu


def _mods():
    from mjrl_b200.utils import process_samples as ps
    from mjrl_b200.baselines.linear_baseline import LinearBaseline
    from mjrl_b200.utils.gym_env import EnvSpec
    from mjrl_b200.batch import drop_cache
    return ps, LinearBaseline, EnvSpec, drop_cache


def test_returns_gae_baseline_match_reference_goldens(golden):
    g = golden
    ps, LinearBaseline, EnvSpec, drop_cache = _mods()
    spec = EnvSpec(g["n"], g["m"], g["horizon"], 1)
    base = LinearBaseline(spec)
    p0 = syn.flat_to_paths(batch_of(g, "b0_"))
    p1 = syn.flat_to_paths(batch_of(g, "b1_"))
    ps.compute_returns(p0, g["gamma"])
    r0 = np.concatenate([p["returns"] for p in p0])
    assert r0.dtype == np.float64
    assert rel_l2(r0, g["b0_returns"]) < 1e-12
    assert np.max(np.abs(r0 - g["b0_returns"]) / np.maximum(1.0, np.abs(g["b0_returns"]))) < 1e-12
    ps.compute_advantages(p0, base, g["gamma"], g["gae_lambda"])   # zero baseline (F8)
    a0 = np.concatenate([p["advantages"] for p in p0])
    assert rel_l2(a0, g["b0_adv_zero_baseline"]) < 1e-12
    eb, ea = base.fit(p0, return_errors=True)
    # predictions with the fitted coefficients, not the coefficients, are the contract
    from oracle import mjrl_oracle as O
    X0 = np.concatenate([O.linear_features(p) for p in p0])
    assert rel_l2(X0 @ np.asarray(base._coeffs), X0 @ g["fit0_coeffs"]) < 1e-8
    assert abs(eb - g["fit0_errors"][0]) < 1e-9 and abs(ea - g["fit0_errors"][1]) < 1e-7
    ps.compute_returns(p1, g["gamma"])
    ps.compute_advantages(p1, base, g["gamma"], g["gae_lambda"])
    b1 = np.concatenate([p["baseline"] for p in p1])
    a1 = np.concatenate([p["advantages"] for p in p1])
    assert rel_l2(b1, g["b1_baseline"]) < 1e-6
    assert rel_l2(a1, g["b1_adv"]) < 1e-6
    pp = copy.deepcopy(p1)
    drop_cache()
    ps.compute_advantages(pp, base, g["gamma"], None, normalize=True)
    assert rel_l2(np.concatenate([p["advantages"] for p in pp]), g["b1_adv_plain_norm"]) < 1e-6
    eb, ea = base.fit(p1, return_errors=True)
    X1 = np.concatenate([O.linear_features(p) for p in p1])
    assert rel_l2(X1 @ np.asarray(base._coeffs), X1 @ g["fit1_coeffs"]) < 1e-8
    assert abs(eb - g["fit1_errors"][0]) < 1e-6 * max(1, g["fit1_errors"][0])
    assert abs(ea - g["fit1_errors"][1]) < 1e-6


@pytest.mark.parametrize("lens", [[1], [2, 1, 3], [1023, 1024, 1025], [5000, 7, 2049], [1] * 300])
@pytest.mark.parametrize("gamma,lam", [(0.995, 0.97), (1.0, 1.0), (0.9, 0.0)])
def test_ragged_paths_scan_vs_oracle(lens, gamma, lam):
    from oracle import mjrl_oracle as O
    ps, LinearBaseline, EnvSpec, drop_cache = _mods()
    rng = np.random.default_rng(sum(lens))
    n = 5
    paths = []
    for i, T in enumerate(lens):
        paths.append(dict(observations=rng.standard_normal((T, n)).astype(np.float32).astype(np.float64) * 4,
                          actions=rng.standard_normal((T, 2)),
                          rewards=rng.standard_normal(T), terminated=bool(i % 2)))
    ref = copy.deepcopy(paths)
    ob = O.LinearBaseline()
    ob._coeffs = rng.standard_normal(n + 4)
    gb = LinearBaseline(EnvSpec(n, 2, 1000, 1))
    gb._coeffs = ob._coeffs.copy()
    O.compute_returns(ref, gamma)
    O.compute_advantages(ref, ob, gamma, lam)
    ps.compute_returns(paths, gamma)
    ps.compute_advantages(paths, gb, gamma, lam)
    for k in ("returns", "baseline", "advantages"):
        a = np.concatenate([p[k] for p in paths])
        b = np.concatenate([p[k] for p in ref])
        assert a.shape == b.shape
        assert rel_l2(a, b) < 1e-10, k
    # per-path predict and discount_sum entry points
    np.testing.assert_allclose(gb.predict(paths[0]), ob.predict(ref[0]), rtol=1e-10, atol=1e-10)
    x = rng.standard_normal(37)
    np.testing.assert_allclose(ps.discount_sum(x, gamma), O.discount_sum(x, gamma), rtol=1e-12, atol=1e-12)


def test_segment_indexing_bit_exact():
    """Constant rewards of 1 with gamma=1: returns[t] = T - t exactly (integers in fp64), so
    any mistake in path segmentation shows up as a bit difference."""
    ps, LinearBaseline, EnvSpec, drop_cache = _mods()
    rng = np.random.default_rng(5)
    lens = rng.integers(1, 3000, size=200)
    paths = [dict(observations=np.zeros((T, 3)), actions=np.zeros((T, 1)), rewards=np.ones(T),
                  terminated=False) for T in lens]
    ps.compute_returns(paths, 1.0)
    for p, T in zip(paths, lens):
        np.testing.assert_array_equal(p["returns"], np.arange(T, 0, -1, dtype=np.float64))


def test_linbase_nan_retry_and_large_feature_count():
    """NaN returns -> 10 tries, NaN coefficients (linear_baseline.py:30-38); Ant-sized F=115."""
    from oracle import mjrl_oracle as O
    ps, LinearBaseline, EnvSpec, drop_cache = _mods()
    rng = np.random.default_rng(11)
    n, T, P = 111, 300, 12
    paths = [dict(observations=3 * rng.standard_normal((T, n)).astype(np.float32).astype(np.float64),
                  actions=np.zeros((T, 8)), rewards=rng.standard_normal(T), terminated=False)
             for _ in range(P)]
    ref = copy.deepcopy(paths)
    O.compute_returns(ref, 0.99)
    ps.compute_returns(paths, 0.99)
    gb, ob = LinearBaseline(EnvSpec(n, 8, T, 1)), O.LinearBaseline()
    eg = gb.fit(paths, return_errors=True)
    eo = ob.fit(ref, return_errors=True)
    X = np.concatenate([O.linear_features(p) for p in ref])
    assert rel_l2(X @ np.asarray(gb._coeffs), X @ ob._coeffs) < 1e-7
    assert abs(eg[0] - eo[0]) < 1e-9 and abs(eg[1] - eo[1]) < 1e-7
    bad = copy.deepcopy(paths)
    drop_cache()
    bad[0]["returns"] = bad[0]["returns"].copy()
    bad[0]["returns"][3] = np.nan
    gb2 = LinearBaseline(EnvSpec(n, 8, T, 1))
    gb2.fit(bad)
    assert np.all(np.isnan(np.asarray(gb2._coeffs)))
    assert gb2.last_fit_info["retries"] == 9


@pytest.mark.parametrize("n,lens", [
    (11, [1000] * 7 + [37]),          # Hopper: split variant, bulk tiles + ragged last tile (rows*n % 4 != 0)
    (11, [25] * 90 + [3]),            # short paths: several boundaries inside every tile
    (8, [300, 513, 2]),               # n % 4 == 0: 16-byte row reads, one 8x8 tile row + target column
    (6, [257, 255]),                  # n % 2 == 0
    (17, [700, 1, 1, 400]),           # 3x3 tiles split variant, single-timestep paths
    (3, [5]),                         # fewer timesteps than one MMA group pair
    (39, [400, 333]),                 # warp-owned tiles (MAXT = 4)
    (61, [300, 211]),                 # MAXT = 8
])
def test_linbase_fit_predict_shapes_vs_oracle(n, lens):
    """Gram (fp64 tensor-core tiles) + solve + predict vs the oracle on shapes that exercise every
    staging variant.  1e-6 relative is the gate (north_star); measured ~1e-12."""
    from oracle import mjrl_oracle as O
    ps, LinearBaseline, EnvSpec, drop_cache = _mods()
    drop_cache()
    rng = np.random.default_rng(n * 1000 + len(lens))
    paths = [dict(observations=(4 * rng.standard_normal((T, n))).astype(np.float32).astype(np.float64),
                  actions=np.zeros((T, 2)), rewards=rng.standard_normal(T), terminated=False)
             for T in lens]
    ref = copy.deepcopy(paths)
    O.compute_returns(ref, 0.98)
    ps.compute_returns(paths, 0.98)
    gb, ob = LinearBaseline(EnvSpec(n, 2, max(lens), 1)), O.LinearBaseline()
    eg = gb.fit(paths, return_errors=True)
    eo = ob.fit(ref, return_errors=True)
    X = np.concatenate([O.linear_features(p) for p in ref])
    want = X @ ob._coeffs
    got_fit = X @ np.asarray(gb._coeffs)
    scale = max(1.0, float(np.abs(want).max()))
    if sum(lens) > n + 8:   # well-posed fits only: tiny batches are dominated by the ridge term
        assert np.max(np.abs(got_fit - want)) / scale < 1e-6
        assert abs(eg[1] - eo[1]) < 1e-6
    # predict with the ORACLE's coefficients: isolates the predict kernel
    gb._coeffs = ob._coeffs.copy()
    drop_cache()
    got = np.concatenate([gb.predict(p) for p in paths])
    assert np.max(np.abs(got - want)) / scale < 1e-9


def test_linbase_kernels_unaligned_obs_base():
    """obs pointer not 16-byte aligned (a row slice of a larger buffer): the kernels must take the
    plain-load staging path instead of bulk copies and give the same Gram / predictions."""
    from mjrl_b200 import _lib
    from mjrl_b200._lib import c_i64, c_int, ptr, stream_ptr
    ctx = _lib.context()
    dev = torch.device("cuda", ctx.device)
    rng = np.random.default_rng(5)
    n, lens = 11, [700, 450, 131]
    N, F = sum(lens), n + 4
    big = torch.from_numpy(rng.standard_normal((N + 1, n)).astype(np.float32) * 6).to(dev)
    obs_u = big[1:]                                  # data_ptr offset 44 B
    obs_a = obs_u.clone()
    assert obs_u.data_ptr() % 16 != 0 and obs_a.data_ptr() % 16 == 0
    ret = torch.from_numpy(rng.standard_normal(N)).to(dev)
    off = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(dev)
    coef = torch.from_numpy(rng.standard_normal(F)).to(dev)
    out = {}
    for name, obs in (("u", obs_u), ("a", obs_a)):
        gram = torch.empty(F * (F + 1) + 1, dtype=torch.float64, device=dev)
        pred = torch.empty(N, dtype=torch.float64, device=dev)
        ctx.call("mjrl_linbase_gram", ptr(obs), ptr(ret), ptr(off), c_int(len(lens)), c_i64(N), c_int(n),
                 ptr(gram), stream_ptr())
        ctx.call("mjrl_linbase_predict", ptr(obs), ptr(off), c_int(len(lens)), c_i64(N), c_int(n), ptr(coef),
                 ptr(pred), stream_ptr())
        out[name] = (gram.cpu().numpy(), pred.cpu().numpy())
    np.testing.assert_array_equal(out["u"][1], out["a"][1])
    np.testing.assert_allclose(out["u"][0], out["a"][0], rtol=1e-12, atol=1e-9)
    # and against numpy
    X = np.concatenate([obs_a.cpu().numpy().clip(-10, 10).astype(np.float64),
                        np.concatenate([np.stack([(np.arange(T) / 1000.0) ** k for k in (1, 2, 3)] +
                                                 [np.ones(T)], 1) for T in lens])], 1)
    np.testing.assert_allclose(out["a"][1], X @ coef.cpu().numpy(), rtol=1e-10, atol=1e-10)
    G = out["a"][0][:F * (F + 1)].reshape(F, F + 1)
    np.testing.assert_allclose(G[:, :F], X.T @ X, rtol=1e-10, atol=1e-8)
    np.testing.assert_allclose(G[:, F], X.T @ ret.cpu().numpy(), rtol=1e-10, atol=1e-8)


def test_staged_moments_equal_single_call_and_combine_across_shards():
    """mjrl_moments_stage (the sharded whitening of algos/engine.py): without all-reduces the two stages give
    mjrl_moments' bits; with the {N, sum} and centred-square sums of two shards added between / after the
    stages they give the moments of the concatenated batch (numpy mean / std of npg_cg.py:91)."""
    import torch
    from mjrl_b200 import _lib
    from mjrl_b200._lib import c_i64, c_int, ptr, stream_ptr
    ctx = _lib.context()
    dev = torch.device("cuda", ctx.device)
    rng = np.random.default_rng(12)
    x = rng.standard_normal(70001) * 3.0 + 0.7
    d_x = torch.from_numpy(x).to(dev)
    one = torch.zeros(4, dtype=torch.float64, device=dev)
    two = torch.zeros(4, dtype=torch.float64, device=dev)
    ctx.call("mjrl_moments", ptr(d_x), c_i64(x.size), ptr(one), stream_ptr())
    ctx.call("mjrl_moments_stage", ptr(d_x), c_i64(x.size), ptr(two), c_int(0), stream_ptr())
    ctx.call("mjrl_moments_stage", ptr(d_x), c_i64(x.size), ptr(two), c_int(1), stream_ptr())
    np.testing.assert_array_equal(one.cpu().numpy(), two.cpu().numpy())
    # two shards of unequal size
    cut = 26000
    shards = [torch.from_numpy(x[:cut].copy()).to(dev), torch.from_numpy(x[cut:].copy()).to(dev)]
    moms = [torch.zeros(4, dtype=torch.float64, device=dev) for _ in shards]
    for s, m in zip(shards, moms):
        ctx.call("mjrl_moments_stage", ptr(s), c_i64(s.numel()), ptr(m), c_int(0), stream_ptr())
    tot = moms[0][0:2] + moms[1][0:2]                 # what Dist.allreduce does to the slice
    for m in moms:
        m[0:2] = tot
    for s, m in zip(shards, moms):
        ctx.call("mjrl_moments_stage", ptr(s), c_i64(s.numel()), ptr(m), c_int(1), stream_ptr())
    ss = float((moms[0][2] + moms[1][2]).item())
    N, mean = float(moms[0][0].item()), float(moms[0][3].item())
    assert N == x.size and abs(mean - x.mean()) < 1e-13 * max(1.0, abs(x.mean()))
    assert abs(np.sqrt(ss / N) - x.std()) < 1e-12 * x.std()
