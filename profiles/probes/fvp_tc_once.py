"""A few launches of the tensor-memory FVP at the headline shape (for ncu).  Usage: fvp_tc_once.py [N]"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from mjrl_b200.policies.gaussian_mlp import MLP
from mjrl_b200.utils.gym_env import EnvSpec
from mjrl_b200.algos.engine import UpdateEngine
N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
pol = MLP(EnvSpec(11, 3, 1000, 1), hidden_sizes=(64, 64), seed=123)
eng = UpdateEngine(pol); dev = eng.device
obs = torch.from_numpy((3 * np.random.default_rng(0).standard_normal((N, 11))).astype(np.float32)).to(dev)
params = eng.dv.params("new"); d = params.numel()
vec = torch.randn(d, dtype=torch.float32, device=dev) * 0.01
out = torch.empty(d, dtype=torch.float64, device=dev)
eng.cg_flags.zero_()
eng.fvp(params, obs, vec, out, cache_mode=1)
assert eng._tc_prepare(params, obs)
for _ in range(4):
    eng.fvp(params, obs, vec, out, cache_mode=4)
torch.cuda.synchronize()
print("ok", float(out.abs().sum()))
