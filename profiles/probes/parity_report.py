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
for name in ["c2_swimmer", "c3_hopper", "c3_hopper_mid", "c5_adroit_mid"]:
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
