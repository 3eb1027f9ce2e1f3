"""Generate golden fixtures by running the LIVE reference (bennevans/mjrl).

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden.py
It imports the reference's own modules (with empty `gym` / `matplotlib` stubs,
because mjrl.algos imports its samplers and logger at module level), feeds them
synthetic batches from `mjrl_b200.utils.synthetic`, and stores inputs and the
reference's outputs in `tests/golden/*.npz`.  Nothing here runs on the GPU box.

Versions recorded in each fixture: numpy / torch of this container.
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = os.environ.get("MJRL_REFERENCE", "/root/reference")


def import_reference():
    for name in ("gym", "matplotlib", "matplotlib.pyplot"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["matplotlib"].use = lambda *a, **k: None
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.path.insert(0, REF)
    saved = {k: os.environ.get(k) for k in ("CUDA_VISIBLE_DEVICES", "CUDA_DEVICE_ORDER")}
    import mjrl.utils.process_samples as ps
    import mjrl.utils.cg_solve as cg
    import mjrl.baselines.linear_baseline as lb
    import mjrl.baselines.mlp_baseline as mb   # sets CUDA_VISIBLE_DEVICES at import (F10)
    import mjrl.policies.gaussian_mlp as gm
    import mjrl.policies.gaussian_linear as gl
    import mjrl.algos.npg_cg as npg
    import mjrl.algos.trpo as trpo
    import mjrl.algos.dapg as dapg
    import mjrl.algos.batch_reinforce as br
    from mjrl.utils.gym_env import EnvSpec
    for k, v in saved.items():
        if v is None:
            os.environ.pop(k, None)
        else:
            os.environ[k] = v
    return types.SimpleNamespace(ps=ps, cg=cg, lb=lb, mb=mb, gm=gm, gl=gl, npg=npg,
                                 trpo=trpo, dapg=dapg, br=br, EnvSpec=EnvSpec)


def build_ref_policy(R, n, m, hidden, seed, flat=None, transforms=None, min_log_std=-3):
    spec = R.EnvSpec(n, m, 1000, 1)
    if hidden:
        pol = R.gm.MLP(spec, hidden_sizes=tuple(hidden), seed=seed, min_log_std=min_log_std)
    else:
        pol = R.gl.LinearPolicy(spec, seed=seed, min_log_std=min_log_std)
    init = pol.get_param_values()
    if transforms is not None:
        pol.model.set_transformations(**transforms)
        pol.old_model.set_transformations(**transforms)
    if flat is not None:
        pol.set_param_values(flat.copy(), set_new=True, set_old=True)
    return pol, init


def run_case(R, name, n, m, hidden, horizon, N, seed, mid, algo="npg"):
    import torch
    from mjrl_b200.utils import synthetic as syn
    spec = R.EnvSpec(n, m, horizon, 1)
    pol, init = build_ref_policy(R, n, m, hidden, seed=123)
    flat, tr = init, None
    if mid:
        flat, tr = syn.mid_training_variant(init, n, m, seed)
        pol, _ = build_ref_policy(R, n, m, hidden, seed=123, flat=flat, transforms=tr)
    gamma, lam = 0.995, 0.97
    out = dict(n=n, m=m, hidden=np.asarray(hidden, dtype=np.int64), horizon=horizon, N=N,
               seed=seed, gamma=gamma, gae_lambda=lam, init_params=init, params=flat,
               numpy_version=np.__version__, torch_version=torch.__version__)
    if tr is not None:
        out.update({"tr_" + k: v for k, v in tr.items()})

    # two batches: batch 0 warms the baseline (F8), batch 1 is the measured update
    b0 = syn.make_flat_batch(N, n, m, horizon, seed, flat, hidden, tr)
    b1 = syn.make_flat_batch(N, n, m, horizon, seed + 1, flat, hidden, tr)
    for tag, b in (("b0_", b0), ("b1_", b1)):
        out.update({tag + k: v for k, v in b.items()})
    p0, p1 = syn.flat_to_paths(b0), syn.flat_to_paths(b1)

    base = R.lb.LinearBaseline(spec)
    R.ps.compute_returns(p0, gamma)
    out["b0_returns"] = np.concatenate([p["returns"] for p in p0])
    # iteration 0: predict() returns zeros (F8)
    R.ps.compute_advantages(p0, base, gamma, lam)
    out["b0_adv_zero_baseline"] = np.concatenate([p["advantages"] for p in p0])
    eb, ea = base.fit(p0, return_errors=True)
    out["fit0_coeffs"] = base._coeffs.copy()
    out["fit0_errors"] = np.asarray([eb, ea])

    R.ps.compute_returns(p1, gamma)
    R.ps.compute_advantages(p1, base, gamma, lam)
    out["b1_returns"] = np.concatenate([p["returns"] for p in p1])
    out["b1_baseline"] = np.concatenate([p["baseline"] for p in p1])
    out["b1_adv"] = np.concatenate([p["advantages"] for p in p1])
    # plain (non-GAE) mode and the normalize flag
    import copy
    pp = copy.deepcopy(p1)
    R.ps.compute_advantages(pp, base, gamma, None, normalize=True)
    out["b1_adv_plain_norm"] = np.concatenate([p["advantages"] for p in pp])

    obs = np.concatenate([p["observations"] for p in p1])
    act = np.concatenate([p["actions"] for p in p1])
    adv = np.concatenate([p["advantages"] for p in p1])
    adv = (adv - np.mean(adv)) / (np.std(adv) + 1e-6)
    out["b1_adv_white"] = adv

    agent = R.npg.NPG(None, pol, base, normalized_step_size=0.01,
                      FIM_invert_args={"iters": 10, "damping": 1e-4})
    out["LL"] = pol.log_likelihood(obs, act)
    out["mean"] = pol.mean_LL(obs, act)[0].data.numpy()
    out["surr_before"] = agent.CPI_surrogate(obs, act, adv).data.numpy().ravel()[0]
    g = agent.flat_vpg(obs, act, adv)
    out["vpg_grad"] = g
    out["hvp_of_grad"] = agent.HVP(obs, act, g, 1e-4)
    out["hvp_of_grad_nodamp"] = agent.HVP(obs, act, g, 0.0)
    hv = agent.build_Hvp_eval([obs, act], regu_coef=1e-4)
    for k in (1, 2, 3, 5, 10):
        out["cg_k%d" % k] = R.cg.cg_solve(hv, g, x_0=g.copy(), cg_iters=k)

    # perturbed "new" parameters: exercises LR != 1 and KL > 0 evaluation
    rng = np.random.default_rng(seed + 17)
    pert = flat + np.float32(0.02) * rng.standard_normal(flat.shape).astype(np.float32)
    pol.set_param_values(pert.copy(), set_new=True, set_old=False)
    out["pert_params"] = pert
    out["pert_surr"] = agent.CPI_surrogate(obs, act, adv).data.numpy().ravel()[0]
    out["pert_kl"] = agent.kl_old_new(obs, act).data.numpy().ravel()[0]
    out["pert_vpg"] = agent.flat_vpg(obs, act, adv)
    out["pert_LL"] = pol.log_likelihood(obs, act)
    pol.set_param_values(flat.copy(), set_new=True, set_old=True)

    # full updates
    agent.train_from_paths(p1)
    out["npg_new_params"] = pol.get_param_values()
    eb, ea = base.fit(p1, return_errors=True)
    out["fit1_coeffs"] = base._coeffs.copy()
    out["fit1_errors"] = np.asarray([eb, ea])

    # TRPO from the same starting point (captures per-trial decisions via stdout-free replay)
    pol, _ = build_ref_policy(R, n, m, hidden, seed=123, flat=flat, transforms=tr)
    t = R.trpo.TRPO(None, pol, base, kl_dist=0.01, FIM_invert_args={"iters": 10, "damping": 1e-4})
    import io, contextlib
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        t.train_from_paths(p1)
    out["trpo_new_params"] = pol.get_param_values()
    out["trpo_backtracks"] = buf.getvalue().count("Backtracking")

    # DAPG with 5 demo paths
    pol, _ = build_ref_policy(R, n, m, hidden, seed=123, flat=flat, transforms=tr)
    db = syn.make_flat_batch(5 * min(horizon, 50), n, m, min(horizon, 50), seed + 2, flat, hidden, tr)
    out.update({"demo_" + k: v for k, v in db.items()})
    dp = syn.flat_to_paths(db)
    dg = R.dapg.DAPG(None, pol, base, demo_paths=dp, kl_dist=0.01,
                     FIM_invert_args={"iters": 10, "damping": 1e-4}, lam_0=1.0, lam_1=0.95)
    dg.train_from_paths(p1)
    out["dapg_new_params"] = pol.get_param_values()

    # batch REINFORCE
    pol, _ = build_ref_policy(R, n, m, hidden, seed=123, flat=flat, transforms=tr)
    br = R.br.BatchREINFORCE(None, pol, base, learn_rate=0.01)
    br.train_from_paths(p1)
    out["reinforce_new_params"] = pol.get_param_values()

    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **out)
    print("wrote", path, "%.1f KB" % (os.path.getsize(path) / 1024))


def run_mlpbase_case(R, name, n, horizon, N, seed, epochs, reg_coef):
    """MLPBaseline (mlp_baseline.py): seed-for-seed init, predictions, one full fit() with the
    minibatch permutations it drew from numpy's global stream, errors, final weights."""
    import torch
    from mjrl_b200.utils import synthetic as syn
    spec = R.EnvSpec(n, 3, horizon, 1)
    zero = np.zeros(n * 4 + 4 + 4 * 4 + 4 + 4 * 3 + 3 + 3, dtype=np.float32)
    b = syn.make_flat_batch(N, n, 3, horizon, seed, zero, (4, 4))
    paths = syn.flat_to_paths(b)
    R.ps.compute_returns(paths, 0.995)
    torch.manual_seed(seed)
    base = R.mb.MLPBaseline(spec, reg_coef=reg_coef, batch_size=64, epochs=epochs)
    flat = lambda: np.concatenate([p.data.numpy().ravel() for p in base.model.parameters()]).astype(np.float32)
    out = dict(n=n, horizon=horizon, N=N, seed=seed, epochs=epochs, reg_coef=reg_coef, batch_size=64,
               numpy_version=np.__version__, torch_version=torch.__version__,
               observations=b["observations"], rewards=b["rewards"], path_offsets=b["path_offsets"],
               terminated=b["terminated"], returns=np.concatenate([p["returns"] for p in paths]),
               init_params=flat(), features=base._features(paths).astype(np.float32))
    out["pred_before"] = np.concatenate([base.predict(p) for p in paths])
    np.random.seed(seed + 1)
    eb, ea = base.fit(paths, return_errors=True)
    np.random.seed(seed + 1)
    out["perms"] = np.stack([np.random.permutation(N) for _ in range(epochs)])
    out["errors"] = np.asarray([eb, ea], dtype=np.float64)
    out["fit_params"] = flat()
    out["pred_after"] = np.concatenate([base.predict(p) for p in paths])
    # the optimiser state persists across fit() calls: a second fit continues the Adam chain
    base.fit(paths)
    out["fit2_params"] = flat()
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **out)
    print("wrote", path, "%.1f KB" % (os.path.getsize(path) / 1024))


def run_quadbase_case(R, name, n, horizon, N, seed):
    """QuadraticBaseline (quadratic_baseline.py): features, first fit (zero predictions before),
    second fit on a second batch (non-trivial error_before), predictions."""
    from mjrl_b200.utils import synthetic as syn
    import mjrl.baselines.quadratic_baseline as qb
    spec = R.EnvSpec(n, 3, horizon, 1)
    zero = np.zeros(n * 4 + 4 + 4 * 4 + 4 + 4 * 3 + 3 + 3, dtype=np.float32)
    out = dict(n=n, horizon=horizon, N=N, seed=seed, numpy_version=np.__version__)
    base = qb.QuadraticBaseline(spec)
    for k in (0, 1):
        b = syn.make_flat_batch(N, n, 3, horizon, seed + k, zero, (4, 4))
        paths = syn.flat_to_paths(b)
        R.ps.compute_returns(paths, 0.995)
        pre = "b%d_" % k
        out[pre + "observations"] = b["observations"]; out[pre + "rewards"] = b["rewards"]
        out[pre + "path_offsets"] = b["path_offsets"]; out[pre + "terminated"] = b["terminated"]
        out[pre + "returns"] = np.concatenate([p["returns"] for p in paths])
        if k == 0:
            out["features_head"] = base._features(paths[:2])
        out[pre + "pred_before"] = np.concatenate([base.predict(p) for p in paths])
        out[pre + "errors"] = np.asarray(base.fit(paths, return_errors=True), dtype=np.float64)
        out[pre + "coeffs"] = np.asarray(base._coeffs, dtype=np.float64)
        out[pre + "pred_after"] = np.concatenate([base.predict(p) for p in paths])
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **out)
    print("wrote", path, "%.1f KB" % (os.path.getsize(path) / 1024))


def run_ppo_case(R, name, n, m, hidden, horizon, N, seed, mid, epochs):
    """PPO (ppo_clip.py): full train_from_paths of the live reference, with the minibatch rows it
    drew from numpy's global stream (np.random.choice, :89) recorded."""
    import torch
    from mjrl_b200.utils import synthetic as syn
    import mjrl.algos.ppo_clip as ppo
    spec = R.EnvSpec(n, m, horizon, 1)
    pol, init = build_ref_policy(R, n, m, hidden, seed=123)
    flat, tr = init, None
    if mid:
        flat, tr = syn.mid_training_variant(init, n, m, seed)
        pol, _ = build_ref_policy(R, n, m, hidden, seed=123, flat=flat, transforms=tr)
    b = syn.make_flat_batch(N, n, m, horizon, seed, flat, hidden, tr)
    paths = syn.flat_to_paths(b)
    rng = np.random.default_rng(seed + 5)
    for p in paths:   # synthetic advantages: PPO only reads paths[...]["advantages"]
        p["advantages"] = rng.standard_normal(len(p["rewards"]))
    out = dict(n=n, m=m, hidden=np.asarray(hidden, dtype=np.int64), horizon=horizon, N=N, seed=seed,
               epochs=epochs, params=flat, init_params=init, numpy_version=np.__version__,
               torch_version=torch.__version__, advantages=np.concatenate([p["advantages"] for p in paths]))
    out.update(b)
    if tr is not None:
        out.update({"tr_" + k: v for k, v in tr.items()})
    agent = ppo.PPO(None, pol, R.lb.LinearBaseline(spec), clip_coef=0.2, epochs=epochs, mb_size=64,
                    learn_rate=3e-4)
    steps = epochs * int(N / 64)
    np.random.seed(seed + 9)
    out["minibatches"] = np.stack([np.random.choice(N, size=64) for _ in range(steps)])
    np.random.seed(seed + 9)
    stats = agent.train_from_paths(paths)
    out["stats"] = np.asarray(stats)
    out["ppo_new_params"] = pol.get_param_values()
    # second iteration on the same paths: Adam state persists in agent.optimizer
    agent.train_from_paths(paths)
    out["ppo_new_params_2"] = pol.get_param_values()
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **out)
    print("wrote", path, "%.1f KB" % (os.path.getsize(path) / 1024))


def run_bc_case(R, name, n, m, hidden, horizon, N, seed, mode, epochs):
    """Behaviour cloning: the live reference's BC class (behavior_cloning.py = "mle",
    behavior_cloning_2.py = "mse") trained twice on synthetic expert paths, minibatch rows from
    numpy's global stream (np.random.choice, :56 / :66) recorded."""
    import torch
    from mjrl_b200.utils import synthetic as syn
    import importlib
    mod = importlib.import_module("mjrl.algos.behavior_cloning" + ("" if mode == "mle" else "_2"))
    pol, init = build_ref_policy(R, n, m, hidden, seed=123)
    # expert data: actions of a DIFFERENT (perturbed) policy so the loss has something to fit
    expert_flat, _ = syn.mid_training_variant(init, n, m, seed)
    b = syn.make_flat_batch(N, n, m, horizon, seed, expert_flat, hidden, None)
    paths = syn.flat_to_paths(b)
    out = dict(n=n, m=m, hidden=np.asarray(hidden, dtype=np.int64), horizon=horizon, N=N, seed=seed,
               epochs=epochs, mode=mode, init_params=init, numpy_version=np.__version__,
               torch_version=torch.__version__)
    out.update(b)
    steps = epochs * int(N / 64)
    np.random.seed(seed + 9)
    out["minibatches"] = np.stack([np.random.choice(N, size=64) for _ in range(2 * steps)])
    agent = mod.BC(paths, pol, epochs=epochs, batch_size=64, lr=1e-3)
    out["params_after_init"] = pol.get_param_values()
    for k in ("in_shift", "in_scale", "out_shift", "out_scale"):
        out["tr_" + k] = getattr(pol.model, k).data.numpy().copy()
    np.random.seed(seed + 9)
    agent.train()
    out["bc_new_params"] = pol.get_param_values()
    out["bc_losses"] = np.asarray(agent.logger.log["loss"], dtype=np.float64)
    agent.train()            # Adam state persists in agent.optimizer
    out["bc_new_params_2"] = pol.get_param_values()
    out["bc_losses_2"] = np.asarray(agent.logger.log["loss"], dtype=np.float64)[len(out["bc_losses"]):]
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **out)
    print("wrote", path, "%.1f KB" % (os.path.getsize(path) / 1024), "losses", out["bc_losses"])


def run_checkpoint_case(R):
    """Checkpoints exactly as train_agent.py:61-66 writes them: pickle.dump of the live reference's
    policy / baseline objects, plus the values a loader must reproduce from them."""
    import pickle
    import torch
    from mjrl_b200.utils import synthetic as syn
    out = dict(numpy_version=np.__version__, torch_version=torch.__version__)
    rng = np.random.default_rng(77)
    # MLP policy, mid-training parameters set through set_param_values (float32 -> new/old alias, F14)
    n, m, hidden = 11, 3, (64, 64)
    pol, init = build_ref_policy(R, n, m, hidden, seed=123)
    flat, tr = syn.mid_training_variant(init, n, m, 4242)
    pol, _ = build_ref_policy(R, n, m, hidden, seed=123, flat=flat, transforms=tr)
    pickle.dump(pol, open(os.path.join(HERE, "stock_policy_mlp.pickle"), "wb"))
    obs = rng.standard_normal((40, n)) * 3
    act = rng.standard_normal((40, m))
    out.update(mlp_params=pol.get_param_values(), mlp_obs=obs, mlp_act=act,
               mlp_ll=pol.log_likelihood(obs, act), mlp_mean=pol.new_dist_info(obs, act)[1].data.numpy())
    # LinearPolicy straight from its constructor (new/old NOT aliased)
    lp, linit = build_ref_policy(R, 6, 2, (), seed=5)
    pickle.dump(lp, open(os.path.join(HERE, "stock_policy_linear.pickle"), "wb"))
    out.update(lin_params=linit)
    # LinearBaseline after one fit
    spec = R.EnvSpec(n, m, 100, 1)
    b = syn.make_flat_batch(1500, n, m, 100, 4243, flat, hidden, tr)
    paths = syn.flat_to_paths(b)
    R.ps.compute_returns(paths, 0.995)
    lb = R.lb.LinearBaseline(spec)
    lb.fit(paths)
    pickle.dump(lb, open(os.path.join(HERE, "stock_baseline_linear.pickle"), "wb"))
    out.update({"b_" + k: v for k, v in b.items()})
    out.update(returns=np.concatenate([p["returns"] for p in paths]), lb_coeffs=lb._coeffs,
               lb_pred=np.concatenate([lb.predict(p) for p in paths]))
    # MLPBaseline after one fit (Adam state inside .optimizer), and what a SECOND fit then gives
    torch.manual_seed(11)
    mb = R.mb.MLPBaseline(spec, reg_coef=1e-3, batch_size=64, epochs=1)
    np.random.seed(12)
    mb.fit(paths)
    pickle.dump(mb, open(os.path.join(HERE, "stock_baseline_mlp.pickle"), "wb"))
    flatp = lambda: np.concatenate([p.data.numpy().ravel() for p in mb.model.parameters()]).astype(np.float32)
    out.update(mb_params=flatp(), mb_pred=np.concatenate([mb.predict(p) for p in paths]))
    np.random.seed(13)
    mb.fit(paths)
    out.update(mb_params_2=flatp())
    path = os.path.join(HERE, "stock_checkpoints.npz")
    np.savez_compressed(path, **out)
    for f in ("stock_policy_mlp", "stock_policy_linear", "stock_baseline_linear", "stock_baseline_mlp"):
        print(f, "%.1f KB" % (os.path.getsize(os.path.join(HERE, f + ".pickle")) / 1024))


def run_rollout_case(R, name, hidden, N, pegasus_seed):
    """The reference's own rollout loops (mjrl/samplers/base_sampler.py:10-88,
    evaluation_sampler.py:11-85) on tests/stub_env.StubEnv with a reference policy."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import stub_env
    import mjrl.samplers.base_sampler as bs
    import mjrl.samplers.evaluation_sampler as es
    env = stub_env.make_stub()
    pol, init = build_ref_policy(R, env.n, env.m, hidden, seed=77)
    rng = np.random.default_rng(pegasus_seed)
    flat = init + 0.2 * rng.standard_normal(init.shape)
    flat[-env.m:] = -0.7
    pol.set_param_values(flat.copy(), set_new=True, set_old=True)
    out = dict(hidden=np.asarray(hidden, dtype=np.int64), flat=pol.get_param_values(), N=N,
               pegasus_seed=pegasus_seed, horizon=env.horizon)
    for tag, paths in (("s", bs.do_rollout(N, pol, T=1e6, env=stub_env.make_stub(), pegasus_seed=pegasus_seed)),
                       ("e", es.do_evaluation_rollout(N, pol, T=1e6, env=stub_env.make_stub(),
                                                      pegasus_seed=pegasus_seed))):
        out[tag + "_lens"] = np.asarray([len(q["rewards"]) for q in paths])
        out[tag + "_terminated"] = np.asarray([bool(q["terminated"]) for q in paths])
        for key in ("observations", "actions", "rewards"):
            out[tag + "_" + key] = np.concatenate([q[key] for q in paths])
        out[tag + "_mean"] = np.concatenate([q["agent_infos"]["mean"] for q in paths])
        out[tag + "_info_t"] = np.concatenate([q["env_infos"]["t"] for q in paths])
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **out)
    print("wrote", path, "%.1f KB" % (os.path.getsize(path) / 1024), "lens", out["s_lens"])


if __name__ == "__main__":
    R = import_reference()
    if "--bc-only" in sys.argv:
        run_bc_case(R, "bc_mle_hopper", 11, 3, (64, 64), 100, 640, 8100, "mle", 2)
        run_bc_case(R, "bc_mse_adroit", 45, 24, (32, 32), 50, 512, 8101, "mse", 2)
        run_bc_case(R, "bc_mle_linear", 6, 2, (), 25, 320, 8102, "mle", 3)
        sys.exit(0)
    if "--checkpoint-only" in sys.argv:
        run_checkpoint_case(R)
        sys.exit(0)
    if "--ppo-only" in sys.argv:
        run_ppo_case(R, "ppo_hopper_mid", 11, 3, (64, 64), 100, 640, 8000, True, 2)
        run_ppo_case(R, "ppo_linear", 6, 2, (), 25, 320, 8001, False, 3)
        sys.exit(0)
    if "--rollout-only" in sys.argv:
        run_rollout_case(R, "rollout_stub_mlp", (32, 32), 12, 500)
        run_rollout_case(R, "rollout_stub_linear", (), 6, 600)
        sys.exit(0)
    if "--quadbase-only" in sys.argv:
        run_quadbase_case(R, "quadbase_pointmass", 6, 25, 1000, 9000)
        run_quadbase_case(R, "quadbase_hopper", 11, 200, 3000, 9001)
        sys.exit(0)
    if "--c4c5-only" in sys.argv:
        # BASELINE.json configs[3] / configs[4] shapes at the class-default (64, 64): Ant obs 111 / act 8 with
        # mid-training transforms (TRPO backtrack count included), Adroit obs 45 / act 24
        run_case(R, "c4_ant_mid", 111, 8, (64, 64), 100, 1200, 4000, True)
        run_case(R, "c5_adroit_64", 45, 24, (64, 64), 100, 1500, 5001, False)
        sys.exit(0)
    if "--mlpbase-only" in sys.argv:
        run_mlpbase_case(R, "mlpbase_1step", 11, 60, 150, 7000, 1, 0.0)
        run_mlpbase_case(R, "mlpbase_hopper", 11, 200, 1500, 7001, 2, 1e-3)
        sys.exit(0)
    #            name            n   m  hidden    hor   N     seed  mid
    run_case(R, "c1_linear",     6,  2, (),        25, 1000, 1000, False)
    run_case(R, "c2_swimmer",   12,  4, (32, 32), 500, 4000, 2000, False)
    run_case(R, "c3_hopper",    11,  3, (64, 64), 200, 3000, 3000, False)
    run_case(R, "c3_hopper_mid", 11, 3, (64, 64), 200, 3000, 3001, True)
    run_case(R, "c5_adroit_mid", 45, 24, (32, 32), 100, 2000, 5000, True)
    run_case(R, "c1_linear_mid", 6,  2, (),        25, 1000, 1001, True)
    run_case(R, "c4_ant_mid", 111, 8, (64, 64), 100, 1200, 4000, True)
    run_case(R, "c5_adroit_64", 45, 24, (64, 64), 100, 1500, 5001, False)
    run_mlpbase_case(R, "mlpbase_1step", 11, 60, 150, 7000, 1, 0.0)
    run_mlpbase_case(R, "mlpbase_hopper", 11, 200, 1500, 7001, 2, 1e-3)
    run_ppo_case(R, "ppo_hopper_mid", 11, 3, (64, 64), 100, 640, 8000, True, 2)
    run_ppo_case(R, "ppo_linear", 6, 2, (), 25, 320, 8001, False, 3)
    run_quadbase_case(R, "quadbase_pointmass", 6, 25, 1000, 9000)
    run_quadbase_case(R, "quadbase_hopper", 11, 200, 3000, 9001)
    run_rollout_case(R, "rollout_stub_mlp", (32, 32), 12, 500)
    run_rollout_case(R, "rollout_stub_linear", (), 6, 600)
