"""Prints the parity figures of the policy passes against the live-reference goldens and the fp64
truth (oracle TruthPolicy), for the tensor-core and the fp32-FMA kernel variants.
Usage on the GPU box: [MJRL_B200_POLICY_PATH=fma] python profiles/probes/parity_report.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from conftest import load_golden, batch_of, rel_l2
from test_gpu_policy import make_policy, agent, prepared_paths
from mjrl_b200.algos.npg_cg import NPG
from oracle import mjrl_oracle as O

print("kernel variant:", os.environ.get("MJRL_B200_POLICY_PATH", "mma (3xTF32 tensor cores)"))
for name in ["c2_swimmer", "c3_hopper", "c3_hopper_mid", "c5_adroit_mid", "c4_ant_mid", "c5_adroit_64"]:
    g = load_golden(name)
    b1 = batch_of(g, "b1_")
    obs, act, adv = b1["observations"], b1["actions"], g["b1_adv_white"]
    pol = make_policy(g)
    ag = agent(NPG, g, pol)
    tr = g["transforms"] or {}
    tp = O.TruthPolicy(O.GaussianPolicy(g["n"], g["m"], g["hidden"], g["params"], **tr))
    grad = ag.flat_vpg(obs, act, adv)
    hv = ag.HVP(obs, act, g["vpg_grad"], 0.0)
    t_grad = tp.vpg(obs, act, adv)
    t_hv = tp.fvp(obs, g["vpg_grad"].astype(np.float64), 0.0)
    print("%-14s grad: vs ref %.2e | vs fp64 truth %.2e (ref itself %.2e)   FVP: vs ref %.2e | vs truth %.2e (ref %.2e)   LL %.2e"
          % (name, rel_l2(grad, g["vpg_grad"]), rel_l2(grad, t_grad), rel_l2(g["vpg_grad"], t_grad),
             rel_l2(hv, g["hvp_of_grad_nodamp"]), rel_l2(hv, t_hv), rel_l2(g["hvp_of_grad_nodamp"], t_hv),
             rel_l2(pol.log_likelihood(obs, act), g["LL"])))


# TRPO line search (trpo.py:105-124) fed the reference's own direction: per-trial KL, distance from the 0.01
# threshold and whether the backtrack count equals the reference's.  tests/test_gpu_policy.py only asserts the
# count when every trial is more than 0.1 % away from the threshold; this table makes a skipped
# assertion visible.
from mjrl_b200.algos.trpo import TRPO
print()
print("TRPO decisions on the reference's direction (kl_dist = 0.01):")
for name in ["c1_linear", "c1_linear_mid", "c2_swimmer", "c3_hopper", "c3_hopper_mid", "c5_adroit_mid", "c4_ant_mid", "c5_adroit_64"]:
    g = load_golden(name)
    pol = make_policy(g)
    ag = agent(TRPO, g, pol, kl_dist=0.01)
    ag.train_from_paths(prepared_paths(g), direction_override=g["cg_k10"], verbose=False)
    trials = ag.last_update["trials"]
    margin = min(abs(t[0] - 0.01) / 0.01 for t in trials)
    back = sum(1 for t in trials if not t[2])
    step_ref = g["trpo_new_params"].astype(np.float64) - g["params"]
    step_new = pol.get_param_values().astype(np.float64) - g["params"]
    print("%-14s trials %d  backtracks %d (reference %d)  kl per trial %s  min margin to threshold %.3f %%  %s  step vs ref %.2e"
          % (name, len(trials), back, int(g["trpo_backtracks"]), ["%.6f" % t[0] for t in trials], 100 * margin,
             "ASSERTED" if margin > 1e-3 else "not asserted (within 0.1 % of the threshold)", rel_l2(step_new, step_ref)))
