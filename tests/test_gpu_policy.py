"""GPU parity of the fused Gaussian-policy passes and the NPG/TRPO/DAPG updates against the
reference's outputs (tests/golden) and the oracle.

Tolerances (north_star): policy gradient, FVP and step direction within 1e-4 relative (fp32);
identical line-search decisions.  The 10-iteration CG direction follows the protocol of
SURVEY.md H1: it must be no further from the fp64 truth than max(1e-4, the reference's own
distance)."""
import numpy as np
import pytest

from conftest import batch_of, rel_l2
from mjrl_b200.utils import synthetic as syn

pytestmark = pytest.mark.gpu
TOL = 1e-4


def make_policy(g, params=None):
    from mjrl_b200.policies.gaussian_mlp import MLP
    from mjrl_b200.policies.gaussian_linear import LinearPolicy
    from mjrl_b200.utils.gym_env import EnvSpec
    spec = EnvSpec(g["n"], g["m"], g["horizon"], 1)
    pol = MLP(spec, hidden_sizes=g["hidden"], seed=123) if g["hidden"] else LinearPolicy(spec, seed=123)
    np.testing.assert_array_equal(pol.get_param_values(), g["init_params"])
    if g["transforms"]:
        pol.model.set_transformations(**g["transforms"])
        pol.old_model.set_transformations(**g["transforms"])
    pol.set_param_values((g["params"] if params is None else params).copy(), True, True)
    return pol


def prepared_paths(g):
    p1 = syn.flat_to_paths(batch_of(g, "b1_"))
    off = g["b1_path_offsets"]
    for i, p in enumerate(p1):
        p["advantages"] = g["b1_adv"][off[i]:off[i + 1]]
        p["returns"] = g["b1_returns"][off[i]:off[i + 1]]
    return p1


def agent(cls, g, pol, **kw):
    from mjrl_b200.baselines.linear_baseline import LinearBaseline
    from mjrl_b200.utils.gym_env import EnvSpec
    return cls(None, pol, LinearBaseline(EnvSpec(g["n"], g["m"], g["horizon"], 1)), **kw)


def test_forward_surrogate_gradient_fvp(golden):
    from mjrl_b200.algos.npg_cg import NPG
    g = golden
    b1 = batch_of(g, "b1_")
    obs, act, adv = b1["observations"], b1["actions"], g["b1_adv_white"]
    pol = make_policy(g)
    ag = agent(NPG, g, pol)
    LL = pol.log_likelihood(obs, act)
    assert rel_l2(LL, g["LL"]) < 1e-5
    assert rel_l2(pol.new_dist_info(obs, act)[1].numpy(), g["mean"]) < 1e-5
    assert abs(float(ag.CPI_surrogate(obs, act, adv)) - g["surr_before"]) < 1e-6
    assert abs(float(ag.kl_old_new(obs, act))) < 1e-7
    grad = ag.flat_vpg(obs, act, adv)
    assert grad.dtype == np.float32 and grad.shape == g["vpg_grad"].shape
    assert rel_l2(grad, g["vpg_grad"]) < TOL
    hv = ag.HVP(obs, act, g["vpg_grad"], 1e-4)
    assert rel_l2(hv, g["hvp_of_grad"]) < TOL
    assert rel_l2(ag.HVP(obs, act, g["vpg_grad"], 0.0), g["hvp_of_grad_nodamp"]) < TOL
    # new != old: likelihood ratio, KL and the LR-weighted gradient
    pol.set_param_values(g["pert_params"].copy(), set_new=True, set_old=False)
    assert abs(float(ag.CPI_surrogate(obs, act, adv)) - g["pert_surr"]) < 1e-5 * max(1, abs(g["pert_surr"]))
    assert abs(float(ag.kl_old_new(obs, act)) - g["pert_kl"]) < 1e-4 * g["pert_kl"] + 1e-8
    assert rel_l2(ag.flat_vpg(obs, act, adv), g["pert_vpg"]) < TOL
    assert rel_l2(pol.log_likelihood(obs, act), g["pert_LL"]) < 1e-5


@pytest.mark.parametrize("variant", ["jac", "act", "tc"])
def test_npg_update_and_cg_protocol(golden, variant, monkeypatch):
    """Which product the CG solve runs, whatever the batch size: "tc" the tensor-memory product (hidden 64 x 64
    shapes; the gradient pass then writes h1|h2 straight into its tile order), "jac" the Jacobian-cached rows
    kernel (act_dim <= 4 shapes), "act" the activation-cached one.  All must pass the same gates."""
    from mjrl_b200.algos import engine
    from mjrl_b200.algos.npg_cg import NPG
    from oracle import mjrl_oracle as O
    monkeypatch.setattr(engine, "JAC_MIN_N", 0 if variant == "jac" else 1 << 30)
    monkeypatch.setattr(engine, "TC_MIN_N", 0 if variant == "tc" else 1 << 30)
    g = golden
    p1 = prepared_paths(g)
    pol = make_policy(g)
    ag = agent(NPG, g, pol, normalized_step_size=0.01)
    trace = []
    stats = ag.train_from_paths(p1, cg_trace=trace)
    R = [sum(p["rewards"]) for p in p1]
    np.testing.assert_allclose(stats, [np.mean(R), np.std(R), np.amin(R), np.amax(R)], rtol=1e-10)
    # (ii) CG iterates k = 1..5 within 1e-4 of the reference
    for k in (1, 2, 3, 5):
        if k <= len(trace):
            assert rel_l2(trace[k - 1], g["cg_k%d" % k]) < TOL, k
    # (iii) k = 10: no further from the fp64 truth than the reference (or 1e-4)
    b1 = batch_of(g, "b1_")
    opol = O.GaussianPolicy(g["n"], g["m"], g["hidden"], g["params"], **(g["transforms"] or {}))
    tp = O.TruthPolicy(opol)
    xs = O.truth_cg(tp, b1["observations"], tp.vpg(b1["observations"], b1["actions"], g["b1_adv_white"]))
    d_ref = rel_l2(g["cg_k10"], xs)
    d_new = rel_l2(trace[-1], xs)
    assert d_new <= max(TOL, 1.05 * d_ref), (d_new, d_ref)
    # committed parameters: our step is no further from the all-fp64 step than the reference's
    gt = tp.vpg(b1["observations"], b1["actions"], g["b1_adv_white"])
    step_true = np.sqrt(abs(0.01 / (gt @ xs + 1e-20))) * xs
    step_ref = g["npg_new_params"].astype(np.float64) - g["params"]
    step_new = pol.get_param_values().astype(np.float64) - g["params"]
    s_ref, s_new = rel_l2(step_ref, step_true), rel_l2(step_new, step_true)
    assert s_new <= max(TOL, 1.05 * s_ref) + 1e-6, (s_new, s_ref)
    np.testing.assert_array_equal(pol.get_param_values(), pol.get_old_param_values())
    assert abs(ag.last_update["kl_dist"] - 0.005) < 0.004


def test_trpo_line_search_decisions(golden):
    from mjrl_b200.algos.trpo import TRPO
    g = golden
    p1 = prepared_paths(g)
    # stage 1: the reference's own direction through OUR step rule + line search
    pol = make_policy(g)
    ag = agent(TRPO, g, pol, kl_dist=0.01)
    ag.train_from_paths(p1, direction_override=g["cg_k10"], verbose=False)
    trials = ag.last_update["trials"]
    margin = min(abs(t[0] - 0.01) / 0.01 for t in trials)
    if margin > 1e-3:   # decisions are only meaningful away from the threshold
        assert sum(1 for t in trials if not t[2]) == int(g["trpo_backtracks"])
        step_ref = g["trpo_new_params"].astype(np.float64) - g["params"]
        step_new = pol.get_param_values().astype(np.float64) - g["params"]
        assert rel_l2(step_new, step_ref) < TOL
    # stage 2: end to end
    pol = make_policy(g)
    ag = agent(TRPO, g, pol, kl_dist=0.01)
    ag.train_from_paths(p1, verbose=False)
    assert ag.last_update["kl_dist"] < 0.01
    assert len(ag.last_update["trials"]) <= 100


def test_dapg_and_reinforce_updates(golden):
    from mjrl_b200.algos.dapg import DAPG
    from mjrl_b200.algos.batch_reinforce import BatchREINFORCE
    from oracle import mjrl_oracle as O
    g = golden
    p1 = prepared_paths(g)
    pol = make_policy(g)
    ag = agent(BatchREINFORCE, g, pol, learn_rate=0.01)
    ag.train_from_paths(p1)
    step_ref = g["reinforce_new_params"].astype(np.float64) - g["params"]
    assert rel_l2(pol.get_param_values().astype(np.float64) - g["params"], step_ref) < TOL

    pol = make_policy(g)
    demo = syn.flat_to_paths(batch_of(g, "demo_"))
    ag = agent(DAPG, g, pol, demo_paths=demo, kl_dist=0.01)
    ag.train_from_paths(p1)
    assert ag.iter_count == 1.0
    # all-fp64 truth of the same update (dapg.py:60-70,97-112); ours must be no further from it
    # than the reference's own fp32 result (the 10-iter direction is chaotic at ~1e-2)
    opol = O.GaussianPolicy(g["n"], g["m"], g["hidden"], g["params"], **(g["transforms"] or {}))
    tp = O.TruthPolicy(opol)
    obs, act, adv = O.concat_paths(prepared_paths(g))
    d_obs = np.concatenate([p["observations"] for p in demo])
    d_act = np.concatenate([p["actions"] for p in demo])
    all_adv = 1e-2 * np.concatenate([adv / (np.std(adv) + 1e-8), np.ones(d_obs.shape[0])])
    gt = tp.vpg(np.concatenate([obs, d_obs]), np.concatenate([act, d_act]), all_adv) * (all_adv.shape[0] / adv.shape[0])
    xs = O.truth_cg(tp, obs, gt)
    step_true = np.sqrt(abs(0.02 / (gt @ xs + 1e-20))) * xs
    step_ref = g["dapg_new_params"].astype(np.float64) - g["params"]
    step_new = pol.get_param_values().astype(np.float64) - g["params"]
    s_ref, s_new = rel_l2(step_ref, step_true), rel_l2(step_new, step_true)
    assert s_new <= max(TOL, 1.05 * s_ref) + 1e-6, (s_new, s_ref)


def test_batched_actions_philox():
    from mjrl_b200.policies.gaussian_mlp import MLP
    from mjrl_b200.utils.gym_env import EnvSpec
    pol = MLP(EnvSpec(11, 3, 1000, 1), hidden_sizes=(64, 64), seed=7)
    pol.set_param_values(np.concatenate([pol.get_param_values()[:-3], [-0.3, 0.1, 0.4]]).astype(np.float32))
    rng = np.random.default_rng(0)
    obs = rng.standard_normal((20000, 11)).astype(np.float32)
    a1, m1 = pol.get_action_batch(obs, seed=5, offset=0)
    a2, m2 = pol.get_action_batch(obs, seed=5, offset=0)
    a3, _ = pol.get_action_batch(obs, seed=6, offset=0)
    a1, m1, a2, a3 = a1.cpu().numpy(), m1.cpu().numpy(), a2.cpu().numpy(), a3.cpu().numpy()
    np.testing.assert_array_equal(a1, a2)               # counter-based: reproducible
    assert np.abs(a1 - a3).max() > 0.1                   # keyed by seed
    eps = (a1 - m1) / np.exp([-0.3, 0.1, 0.4])
    assert abs(eps.mean()) < 0.02 and abs(eps.std() - 1) < 0.02
    assert abs(np.corrcoef(eps[:, 0], eps[:, 1])[0, 1]) < 0.03
    # exact parity on the mean with the single-observation API (gaussian_mlp.py:92-98)
    act, info = pol.get_action(obs[3].astype(np.float64))
    np.testing.assert_allclose(info["mean"], m1[3], rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("n,m,hidden,N", [(11, 3, (64, 64), 1000), (11, 3, (64, 64), 777), (17, 6, (32, 32), 300),
                                           (8, 2, (64, 32), 129), (3, 1, (16, 16), 64)])
def test_fvp_variants_agree_and_cache_is_bit_exact(n, m, hidden, N):
    """FVP kernel variants over shapes that take the rows-in-registers kernel (hidden 64 / 32 / mixed,
    ragged N, several obs widths) against the oracle's fp64 Gauss-Newton product, and the activation
    cache: filling must not change the product, using it must agree to fp32 rounding and be deterministic."""
    import torch
    from mjrl_b200.policies.gaussian_mlp import MLP
    from mjrl_b200.utils.gym_env import EnvSpec
    from mjrl_b200.algos.engine import UpdateEngine
    from oracle import mjrl_oracle as O
    rng = np.random.default_rng(n * 100 + N)
    pol = MLP(EnvSpec(n, m, 100, 1), hidden_sizes=hidden, seed=5)
    flat = pol.get_param_values()
    flat = flat + 0.3 * rng.standard_normal(flat.shape).astype(np.float32)
    pol.set_param_values(flat.copy(), True, True)
    eng = UpdateEngine(pol)
    dev = eng.device
    obs = torch.from_numpy(rng.standard_normal((N, n)).astype(np.float32)).to(dev)
    vec = torch.from_numpy(rng.standard_normal(flat.shape[0]).astype(np.float32) * 0.1).to(dev)
    params = eng.dv.params("new")
    outs = []
    for mode in (0, 1, 2, 2):
        out = torch.zeros(flat.shape[0], dtype=torch.float64, device=dev)
        eng.cg_flags.zero_()
        eng.fvp(params, obs, vec, out, cache_mode=mode)
        outs.append(out.cpu().numpy())
    np.testing.assert_array_equal(outs[1], outs[0])        # filling the cache does not change the product
    np.testing.assert_array_equal(outs[3], outs[2])        # the cached product is deterministic
    # cached variant: same activations bit for bit, its own (fixed) summation order
    assert rel_l2(outs[2], outs[0]) < 2e-6
    # fp64 truth: sum_t J^T diag(2/(2 sigma^2 + 1e-8)) J v on the network block, N h(sigma) v on log_std
    tp = O.TruthPolicy(O.GaussianPolicy(n, m, hidden, flat.copy()))
    truth = N * tp.fvp(obs.cpu().numpy(), vec.cpu().numpy().astype(np.float64))
    assert rel_l2(outs[0], truth) < 1e-5


@pytest.mark.parametrize("n,m,hidden,N,mid", [(11, 3, (64, 64), 1000, False), (11, 3, (64, 64), 777, True),
                                               (12, 4, (32, 32), 500, True), (8, 2, (64, 32), 129, False),
                                               (3, 1, (16, 16), 64, False), (17, 4, (32, 32), 4097, True),
                                               (14, 3, (64, 64), 2049, True)])
def test_fvp_jacobian_cached_variant(n, m, hidden, N, mid):
    """cache_mode 3 (csrc/fvp_jac.cu, act_dim <= 4): the rows G1 = d mu_j / d z1 are filled once from the
    cached h1|h2, then the product runs without tangent / backward-data MMAs.  Same product as mode 0 to
    fp32 rounding, deterministic, exact against the fp64 Gauss-Newton truth, several vectors per fill, with
    in/out transforms, ragged N, mixed hidden sizes; act_dim > 4 has no such variant."""
    import ctypes
    import torch
    from mjrl_b200 import _lib
    from mjrl_b200.policies.gaussian_mlp import MLP
    from mjrl_b200.utils.gym_env import EnvSpec
    from mjrl_b200.algos.engine import UpdateEngine
    from mjrl_b200.utils import synthetic as syn
    from oracle import mjrl_oracle as O
    rng = np.random.default_rng(n * 100 + N)
    pol = MLP(EnvSpec(n, m, 100, 1), hidden_sizes=hidden, seed=5)
    flat = pol.get_param_values()
    tr = None
    if mid:
        flat, tr = syn.mid_training_variant(flat, n, m, 77)
        pol.model.set_transformations(**tr); pol.old_model.set_transformations(**tr)
    else:
        flat = flat + 0.3 * rng.standard_normal(flat.shape).astype(np.float32)
    pol.set_param_values(flat.copy(), True, True)
    eng = UpdateEngine(pol)
    dev = eng.device
    assert _lib.load().mjrl_policy_fvp_jac_ok(eng.ctx.h, eng.dv.desc_ref()) == 1
    obs = torch.from_numpy((3 * rng.standard_normal((N, n))).astype(np.float32)).to(dev)
    params = eng.dv.params("new")
    tp = O.TruthPolicy(O.GaussianPolicy(n, m, hidden, flat.copy(), **(tr or {})))

    def product(vec, mode):
        out = torch.zeros(flat.shape[0], dtype=torch.float64, device=dev)
        eng.cg_flags.zero_()
        eng.fvp(params, obs, vec, out, cache_mode=mode)
        return out.cpu().numpy()

    vecs = [torch.from_numpy(rng.standard_normal(flat.shape[0]).astype(np.float32) * s).to(dev) for s in (0.1, 3.0, 1e-3)]
    ref0 = product(vecs[0], 1)                     # fills h1|h2
    assert eng._jac_prepare(params, obs)           # fills G1 (once: the second call is a no-op)
    assert eng._jac_prepare(params, obs)
    for k, vec in enumerate(vecs):
        ref = ref0 if k == 0 else product(vec, 2)
        a, b = product(vec, 3), product(vec, 3)
        np.testing.assert_array_equal(a, b)
        assert rel_l2(a, ref) < 3e-6, (k, rel_l2(a, ref))
        truth = N * tp.fvp(obs.cpu().numpy(), vec.cpu().numpy().astype(np.float64))
        assert rel_l2(a, truth) < 1e-5, (k, rel_l2(a, truth))
        # the log_std block is the closed form of the reduce kernel in every mode
        np.testing.assert_array_equal(a[-m:], ref[-m:])


def test_fvp_jacobian_variant_not_offered_for_wide_actions():
    from mjrl_b200 import _lib
    from mjrl_b200.policies.gaussian_mlp import MLP
    from mjrl_b200.utils.gym_env import EnvSpec
    from mjrl_b200.algos.engine import UpdateEngine
    eng = UpdateEngine(MLP(EnvSpec(17, 6, 100, 1), hidden_sizes=(32, 32), seed=5))
    assert _lib.load().mjrl_policy_fvp_jac_ok(eng.ctx.h, eng.dv.desc_ref()) == 0
    import torch
    obs = torch.zeros((64, 17), device=eng.device)
    assert eng._jac_prepare(eng.dv.params("new"), obs) is False


def test_cg_solve_reference_signature_over_hvp_eval(golden):
    """utils.cg_solve (mjrl/utils/cg_solve.py:3-22) with the operator npg_cg.py:122-124 builds:
    `cg_solve(agent.build_Hvp_eval([obs, act], regu_coef), vpg_grad, x_0=vpg_grad.copy(), cg_iters=k)`
    against the reference's iterates (k <= 5 within 1e-4; k = 10 by the H1 protocol)."""
    from mjrl_b200.algos.npg_cg import NPG
    from mjrl_b200.utils.cg_solve import cg_solve
    from oracle import mjrl_oracle as O
    g = golden
    b1 = batch_of(g, "b1_")
    obs, act = b1["observations"], b1["actions"]
    ag = agent(NPG, g, make_policy(g))
    hvp = ag.build_Hvp_eval([obs, act], regu_coef=1e-4)
    grad = g["vpg_grad"]
    for k in (1, 2, 3, 5):
        x = cg_solve(hvp, grad, x_0=grad.copy(), cg_iters=k)
        assert x.dtype == np.float32 and rel_l2(x, g["cg_k%d" % k]) < TOL, k
    x10 = cg_solve(hvp, grad, x_0=grad.copy(), cg_iters=10)
    opol = O.GaussianPolicy(g["n"], g["m"], g["hidden"], g["params"], **(g["transforms"] or {}))
    tp = O.TruthPolicy(opol)
    xs = O.truth_cg(tp, obs, tp.vpg(obs, act, g["b1_adv_white"]))
    assert rel_l2(x10, xs) <= max(TOL, 1.05 * rel_l2(g["cg_k10"], xs))
    # early exit: a huge tolerance stops after the first update, like the reference (:19-20)
    np.testing.assert_array_equal(cg_solve(hvp, grad, cg_iters=10, residual_tol=1e30),
                                  cg_solve(hvp, grad, cg_iters=1))


@pytest.mark.parametrize("case", ["c3_hopper", "c2_swimmer"])
def test_consecutive_updates_refresh_the_jacobian_cache(case, monkeypatch):
    """Two train_from_paths calls on ONE agent over a caller-resident batch (same obs buffer, parameters
    changed by the first update): the second solve must run on Jacobian rows of the CURRENT policy.  The
    device parameter tensor of every update is a fresh allocation that the caching allocator places at
    the address of the previous one, which once made the stale rows look current."""
    from conftest import load_golden
    from mjrl_b200.algos import engine
    from mjrl_b200.algos.npg_cg import NPG
    from oracle import mjrl_oracle as O
    g = load_golden(case)
    iters = 3            # k <= 3 iterates of the reference agree with any summation order to 3e-7
    finals = {}
    for jac_min_n, tc_min_n in ((0, 1 << 30), (1 << 30, 1 << 30), (1 << 30, 0)):
        monkeypatch.setattr(engine, "JAC_MIN_N", jac_min_n)
        monkeypatch.setattr(engine, "TC_MIN_N", tc_min_n)
        p1 = prepared_paths(g)
        pol = make_policy(g)
        ag = agent(NPG, g, pol, normalized_step_size=0.01, FIM_invert_args={'iters': iters, 'damping': 1e-4})
        ag.train_from_paths(p1)
        mid = pol.get_param_values().copy()
        ag.train_from_paths(p1)
        finals[(jac_min_n, tc_min_n)] = (mid, pol.get_param_values().copy())
    opol = O.GaussianPolicy(g["n"], g["m"], g["hidden"], g["params"], **(g["transforms"] or {}))
    p1 = prepared_paths(g)
    o1 = O.npg_update(opol, p1, cg_iters=iters)["new_params"].copy()
    o2 = O.npg_update(opol, p1, cg_iters=iters)["new_params"].copy()
    for mid, fin in finals.values():
        assert rel_l2(mid - g["params"], o1 - g["params"]) < TOL
        assert rel_l2(fin.astype(np.float64) - mid, o2.astype(np.float64) - o1) < 10 * TOL


@pytest.mark.parametrize("n,m,N,mid", [(11, 3, 1000, False), (11, 3, 777, True), (12, 4, 96, True), (8, 2, 97, False),
                                       (3, 1, 64, False), (15, 8, 4097, True), (14, 6, 2049, True), (11, 3, 40000, True)])
def test_fvp_tensor_memory_variant(n, m, N, mid):
    """cache_mode 4 (csrc/fvp_tc.cu, hidden 64 x 64): the product on tcgen05 tensor cores with accumulators and A
    operands in tensor memory, reading h1|h2 from a tile-ordered copy of the cache (one relayout per solve).
    Same product as the rows-in-registers kernels to fp32 rounding, deterministic, exact against the fp64
    Gauss-Newton truth, several vectors per fill, in/out transforms, ragged N (partial last tile, several tiles
    per CTA at N = 40000), act_dim up to 8."""
    import torch
    from mjrl_b200 import _lib
    from mjrl_b200.policies.gaussian_mlp import MLP
    from mjrl_b200.utils.gym_env import EnvSpec
    from mjrl_b200.algos.engine import UpdateEngine
    from oracle import mjrl_oracle as O
    rng = np.random.default_rng(n * 100 + N)
    pol = MLP(EnvSpec(n, m, 100, 1), hidden_sizes=(64, 64), seed=5)
    flat = pol.get_param_values()
    tr = None
    if mid:
        flat, tr = syn.mid_training_variant(flat, n, m, 77)
        pol.model.set_transformations(**tr); pol.old_model.set_transformations(**tr)
    else:
        flat = flat + 0.3 * rng.standard_normal(flat.shape).astype(np.float32)
    pol.set_param_values(flat.copy(), True, True)
    eng = UpdateEngine(pol)
    dev = eng.device
    assert _lib.load().mjrl_policy_fvp_tc_ok(eng.ctx.h, eng.dv.desc_ref()) == 1
    obs = torch.from_numpy((3 * rng.standard_normal((N, n))).astype(np.float32)).to(dev)
    params = eng.dv.params("new")
    tp = O.TruthPolicy(O.GaussianPolicy(n, m, (64, 64), flat.copy(), **(tr or {})))

    def product(vec, mode):
        out = torch.zeros(flat.shape[0], dtype=torch.float64, device=dev)
        eng.cg_flags.zero_()
        eng.fvp(params, obs, vec, out, cache_mode=mode)
        return out.cpu().numpy()

    vecs = [torch.from_numpy(rng.standard_normal(flat.shape[0]).astype(np.float32) * s).to(dev) for s in (0.1, 3.0, 1e-3)]
    ref0 = product(vecs[0], 1)                    # fills h1|h2
    assert eng._tc_prepare(params, obs)           # relayout (once: the second call is a no-op)
    assert eng._tc_prepare(params, obs)
    for k, vec in enumerate(vecs):
        ref = ref0 if k == 0 else product(vec, 2)
        a, b = product(vec, 4), product(vec, 4)
        np.testing.assert_array_equal(a, b)
        assert rel_l2(a, ref) < 3e-6, (k, rel_l2(a, ref))
        if N <= 5000:
            truth = N * tp.fvp(obs.cpu().numpy(), vec.cpu().numpy().astype(np.float64))
            assert rel_l2(a, truth) < 1e-5, (k, rel_l2(a, truth))
        # per block of the flat vector (a wrong bias or head block hides in the norm of W1)
        F = [0, 64 * n, 64 * n + 64, 64 * n + 64 + 4096, 64 * n + 128 + 4096, 64 * n + 128 + 4096 + 64 * m,
             64 * n + 128 + 4096 + 64 * m + m]
        for lo, hi in zip(F[:-1], F[1:]):
            assert rel_l2(a[lo:hi], ref[lo:hi]) < 1e-5, (k, lo, hi, rel_l2(a[lo:hi], ref[lo:hi]))
        np.testing.assert_array_equal(a[-m:], ref[-m:])


@pytest.mark.parametrize("n,m,N,mid", [(111, 8, 1000, True), (45, 24, 777, True), (24, 3, 96, False), (63, 16, 4097, True),
                                       (64, 9, 2049, False), (127, 24, 500, True), (31, 1, 97, False), (111, 8, 40000, True),
                                       (45, 24, 30000, False)])
def test_fvp_tensor_memory_wide_variant(n, m, N, mid):
    """cache_mode 4 for the shapes the fused kernel does not hold (csrc/fvp_tc_wide.cu: obs_dim + 1 <= 128, act_dim <=
    24; Ant 111 / 8, Adroit 45 / 24): first layer as two tcgen05 passes of its own around the main kernel, fed by the
    gradient pass (planes + x~ image).  Same product as the shared-memory tile kernel to fp32 rounding, deterministic,
    exact against the fp64 Gauss-Newton truth, per block of the flat vector, ragged N, several tiles per CTA."""
    import torch
    from mjrl_b200 import _lib
    from mjrl_b200.policies.gaussian_mlp import MLP
    from mjrl_b200.utils.gym_env import EnvSpec
    from mjrl_b200.algos import engine as E
    from oracle import mjrl_oracle as O
    rng = np.random.default_rng(n * 100 + N)
    pol = MLP(EnvSpec(n, m, 100, 1), hidden_sizes=(64, 64), seed=5)
    flat = pol.get_param_values()
    tr = None
    if mid:
        flat, tr = syn.mid_training_variant(flat, n, m, 77)
        pol.model.set_transformations(**tr); pol.old_model.set_transformations(**tr)
    else:
        flat = flat + 0.1 * rng.standard_normal(flat.shape).astype(np.float32)
    pol.set_param_values(flat.copy(), True, True)
    eng = E.UpdateEngine(pol)
    dev = eng.device
    assert _lib.load().mjrl_policy_fvp_tc_ok(eng.ctx.h, eng.dv.desc_ref()) == 2
    obs = torch.from_numpy((2 * rng.standard_normal((N, n))).astype(np.float32)).to(dev)
    act = torch.from_numpy(rng.standard_normal((N, m)).astype(np.float32)).to(dev)
    w = torch.from_numpy(rng.standard_normal(N).astype(np.float32)).to(dev)
    params = eng.dv.params("new")
    tp = O.TruthPolicy(O.GaussianPolicy(n, m, (64, 64), flat.copy(), **(tr or {})))
    keep, E.TC_MIN_N = E.TC_MIN_N, 1 << 60
    try:
        g_plain = eng.gradient(params, obs, act, w).clone()
        E.TC_MIN_N = 0
        g_tc = eng.gradient(params, obs, act, w).clone()          # also leaves the planes and the x~ image
    finally:
        E.TC_MIN_N = keep
    assert rel_l2(g_tc.cpu().numpy()[:-1], g_plain.cpu().numpy()[:-1]) < 2e-5     # tcgen05 pass (csrc/pol_tc.cu) vs mma.sync tile kernel
    assert eng._tc_key is not None

    def product(vec, mode):
        out = torch.zeros(flat.shape[0], dtype=torch.float64, device=dev)
        eng.cg_flags.zero_()
        eng.fvp(params, obs, vec, out, cache_mode=mode)
        return out.cpu().numpy()

    F = [0, 64 * n, 64 * n + 64, 64 * n + 64 + 4096, 64 * n + 128 + 4096, 64 * n + 128 + 4096 + 64 * m,
         64 * n + 128 + 4096 + 64 * m + m]
    for k, s in enumerate((0.1, 3.0, 1e-3)):
        vec = torch.from_numpy(rng.standard_normal(flat.shape[0]).astype(np.float32) * s).to(dev)
        a, b = product(vec, 4), product(vec, 4)
        np.testing.assert_array_equal(a, b)
        truth = None
        if N <= 5000:
            truth = N * tp.fvp(obs.cpu().numpy(), vec.cpu().numpy().astype(np.float64))
            assert rel_l2(a, truth) < 1e-5, (k, rel_l2(a, truth))
        if n + m > 140:
            ref = truth       # the tile kernel's product does not fit shared memory at this shape (its gradient pass does)
        else:
            ref = product(vec, 0)
            assert rel_l2(a, ref) < 3e-6, (k, rel_l2(a, ref))
        for lo, hi in zip(F[:-1], F[1:]):
            assert rel_l2(a[lo:hi], ref[lo:hi]) < 1e-5, (k, lo, hi, rel_l2(a[lo:hi], ref[lo:hi]))
        if ref is not truth:
            np.testing.assert_array_equal(a[-m:], ref[-m:])


@pytest.mark.parametrize("n,m,N,mid", [(11, 3, 1000, True), (11, 3, 97, False), (15, 8, 4097, True), (3, 1, 96, False), (12, 4, 40000, True),
                                       (111, 8, 1000, True), (45, 24, 777, True), (24, 3, 96, False), (63, 16, 4097, False),
                                       (17, 6, 30000, True)])
def test_gradient_and_evaluation_on_tensor_memory(n, m, N, mid):
    """csrc/pol_tc.cu: flat_vpg and the surrogate / KL evaluation on tcgen05 (narrow shapes fused, wide shapes around
    the first-layer passes of fvp_tc_wide.cu) against the mma.sync kernels on the same inputs: gradient per block of
    the flat vector, surrogate sum, per-timestep LL and mean, KL sum; the planes it leaves feed the product."""
    import torch
    from mjrl_b200.policies.gaussian_mlp import MLP
    from mjrl_b200.utils.gym_env import EnvSpec
    from mjrl_b200.algos import engine as E
    from oracle import mjrl_oracle as O
    rng = np.random.default_rng(n * 1000 + N)
    pol = MLP(EnvSpec(n, m, 100, 1), hidden_sizes=(64, 64), seed=5)
    flat = pol.get_param_values()
    tr = None
    if mid:
        flat, tr = syn.mid_training_variant(flat, n, m, 77)
        pol.model.set_transformations(**tr); pol.old_model.set_transformations(**tr)
    else:
        flat = flat + 0.1 * rng.standard_normal(flat.shape).astype(np.float32)
    pol.set_param_values(flat.copy(), True, True)
    eng = E.UpdateEngine(pol)
    dev = eng.device
    assert eng._tc_kind() in (1, 2)
    obs = torch.from_numpy((2 * rng.standard_normal((N, n))).astype(np.float32)).to(dev)
    w = torch.from_numpy(rng.standard_normal(N).astype(np.float32)).to(dev)
    params = eng.dv.params("new")
    d = flat.shape[0]
    # actions drawn from the policy itself (|LL| of order act_dim: fp32 log-likelihoods stay well conditioned)
    _, mean_now = eng.old_cache(params, obs, torch.zeros(N, m, dtype=torch.float32, device=dev))
    act = (mean_now + torch.exp(params[-m:]) * torch.from_numpy(rng.standard_normal((N, m)).astype(np.float32)).to(dev)).contiguous()
    old_ll, old_mean = eng.old_cache(params, obs, act)
    old_ll = old_ll + torch.from_numpy((0.05 * rng.standard_normal(N)).astype(np.float32)).to(dev)   # a ratio that is not 1
    keep = E.TC_MIN_N
    out = {}
    try:
        for name, thr in (("plain", 1 << 60), ("tc", 0)):
            E.TC_MIN_N = thr
            ll = torch.zeros(N, dtype=torch.float32, device=dev)
            mean = torch.zeros(N, m, dtype=torch.float32, device=dev)
            g = eng.gradient(params, obs, act, w, old_ll=old_ll, cache=(ll, mean)).clone().cpu().numpy()
            pnew = params + 0.01 * torch.from_numpy(rng.standard_normal(d).astype(np.float32)).to(dev) if name == "plain" else pnew
            ols = params[-m:].clone()
            ev = eng.evaluate(pnew, obs, act, w, old_ll, old_mean, ols, N, raw=True).clone().cpu().numpy()
            out[name] = (g, ll.cpu().numpy(), mean.cpu().numpy(), ev)
        # the planes the tcgen05 gradient pass left are the ones the product reads
        vec = torch.from_numpy(rng.standard_normal(d).astype(np.float32) * 0.1).to(dev)
        res = torch.zeros(d, dtype=torch.float64, device=dev)
        eng.cg_flags.zero_()
        E.TC_MIN_N = 0
        eng.gradient(params, obs, act, w, old_ll=old_ll)
        eng.fvp(params, obs, vec, res, cache_mode=4)
        g_unit = eng.gradient(params, obs, act, w).clone().cpu().numpy()      # ratio 1: the analytic fp64 gradient applies
    finally:
        E.TC_MIN_N = keep
    (g0, ll0, mean0, ev0), (g1, ll1, mean1, ev1) = out["plain"], out["tc"]
    F = [0, 64 * n, 64 * n + 64, 64 * n + 64 + 4096, 64 * n + 128 + 4096, 64 * n + 128 + 4096 + 64 * m,
         64 * n + 128 + 4096 + 64 * m + m, d]
    for lo, hi in zip(F[:-1], F[1:]):
        assert rel_l2(g1[lo:hi], g0[lo:hi]) < 2e-5, (lo, hi, rel_l2(g1[lo:hi], g0[lo:hi]))
    assert abs(g1[d] - g0[d]) < 1e-5 * N ** 0.5 + 1e-5 * abs(g0[d])      # surrogate sum: N mixed-sign terms of order 1
    assert rel_l2(g1[:d], g0[:d]) < 2e-5
    np.testing.assert_allclose(ll1, ll0, rtol=2e-5, atol=2e-5)
    np.testing.assert_allclose(mean1, mean0, rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(ev1, ev0, rtol=2e-5, atol=1e-4 * N ** 0.5)
    if N <= 5000:
        tp = O.TruthPolicy(O.GaussianPolicy(n, m, (64, 64), flat.copy(), **(tr or {})))
        truth = N * tp.fvp(obs.cpu().numpy(), vec.cpu().numpy().astype(np.float64))
        assert rel_l2(res.cpu().numpy(), truth) < 1e-5
        g_truth = N * tp.vpg(obs.cpu().numpy(), act.cpu().numpy(), w.cpu().numpy())
        assert rel_l2(g_unit[:d], g_truth) < 1e-5, rel_l2(g_unit[:d], g_truth)


def test_fvp_tensor_memory_variant_not_offered_for_other_shapes():
    from mjrl_b200 import _lib
    from mjrl_b200.policies.gaussian_mlp import MLP
    from mjrl_b200.utils.gym_env import EnvSpec
    from mjrl_b200.algos.engine import UpdateEngine
    for n, m, hidden in ((11, 3, (32, 32)), (11, 3, (64, 32)), (128, 3, (64, 64)), (45, 25, (64, 64)), (111, 8, (32, 32))):
        eng = UpdateEngine(MLP(EnvSpec(n, m, 100, 1), hidden_sizes=hidden, seed=5))
        assert _lib.load().mjrl_policy_fvp_tc_ok(eng.ctx.h, eng.dv.desc_ref()) == 0, (n, m, hidden)
    # obs_dim 17 (HalfCheetah / Walker shapes): too wide for the fused kernel and for the rows-in-registers plan -> three-pass form
    eng = UpdateEngine(MLP(EnvSpec(17, 6, 100, 1), hidden_sizes=(64, 64), seed=5))
    assert _lib.load().mjrl_policy_fvp_tc_ok(eng.ctx.h, eng.dv.desc_ref()) == 2


def test_replaced_advantages_and_returns_are_honoured():
    """The reference re-reads path['advantages'] / path['returns'] from the dicts on every call (npg_cg.py:87-91,
    linear_baseline.py:22-23).  A caller that replaces them between compute_advantages and train_from_paths /
    baseline.fit must see the replacement used, not the device copy of the earlier computation."""
    from conftest import load_golden
    from mjrl_b200.algos.npg_cg import NPG
    from mjrl_b200.baselines.linear_baseline import LinearBaseline
    from mjrl_b200.utils import process_samples as ps
    from mjrl_b200.utils.gym_env import EnvSpec
    from oracle import mjrl_oracle as O
    g = load_golden("c3_hopper")
    paths = syn.flat_to_paths(batch_of(g, "b1_"))
    spec = EnvSpec(g["n"], g["m"], g["horizon"], 1)
    base = LinearBaseline(spec)
    base._coeffs = g["fit0_coeffs"].copy()
    ps.compute_returns(paths, g["gamma"])
    ps.compute_advantages(paths, base, g["gamma"], g["gae_lambda"])      # scatters device results into the dicts
    rng = np.random.default_rng(5)
    for p in paths:                                                      # user post-processing: NEW arrays
        p["advantages"] = p["advantages"] * rng.uniform(0.5, 2.0) + 0.3
        p["returns"] = p["returns"] - 1.5
    o_paths = [dict(p) for p in paths]
    pol = make_policy(g)
    ag = agent(NPG, g, pol, normalized_step_size=0.01, FIM_invert_args={'iters': 3, 'damping': 1e-4})
    ag.train_from_paths(paths)
    opol = O.GaussianPolicy(g["n"], g["m"], g["hidden"], g["params"], **(g["transforms"] or {}))
    out = O.npg_update(opol, o_paths, cg_iters=3)
    assert rel_l2(pol.get_param_values() - g["params"], out["new_params"] - g["params"]) < TOL
    base.fit(paths)
    ob = O.LinearBaseline(); ob.fit(o_paths)
    X = np.concatenate([O.linear_features(p) for p in o_paths])
    assert rel_l2(X @ np.asarray(base._coeffs), X @ ob._coeffs) < 1e-6
