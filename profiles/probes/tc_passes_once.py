"""One gradient pass, two products and one evaluation through the tcgen05 kernels at a given shape (for ncu).
Usage: tc_passes_once.py N n m"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from mjrl_b200.policies.gaussian_mlp import MLP
from mjrl_b200.utils.gym_env import EnvSpec
from mjrl_b200.algos import engine as E
N, n, m = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
pol = MLP(EnvSpec(n, m, 1000, 1), hidden_sizes=(64, 64), seed=123)
eng = E.UpdateEngine(pol); dev = eng.device
g = torch.Generator(device=dev); g.manual_seed(0)
obs = torch.randn(N, n, dtype=torch.float32, device=dev, generator=g) * 2
act = torch.randn(N, m, dtype=torch.float32, device=dev, generator=g)
w = torch.randn(N, dtype=torch.float32, device=dev, generator=g)
params = eng.dv.params("new"); d = params.numel()
vec = torch.randn(d, dtype=torch.float32, device=dev) * 0.01
out = torch.empty(d, dtype=torch.float64, device=dev)
eng.cg_flags.zero_()
E.TC_MIN_N = 0
ll = torch.zeros(N, dtype=torch.float32, device=dev); mean = torch.zeros(N, m, dtype=torch.float32, device=dev)
eng.gradient(params, obs, act, w, cache=(ll, mean))
for _ in range(2):
    eng.fvp(params, obs, vec, out, cache_mode=4)
eng.evaluate(params + 0.01 * vec, obs, act, w, ll, mean, params[-m:].clone(), N, raw=True)
torch.cuda.synchronize()
print("ok", float(out.abs().sum()))
