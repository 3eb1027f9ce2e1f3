"""Time of one wide tcgen05 product (three kernels + reduce, csrc/fvp_tc_wide.cu) against the shared-memory tile
kernel at the Ant / Adroit shapes.  CUDA events, median of 10.  Usage: python profiles/probes/fvp_wide_time.py N n m"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from mjrl_b200.policies.gaussian_mlp import MLP
from mjrl_b200.utils.gym_env import EnvSpec
from mjrl_b200.algos import engine as E

N = int(sys.argv[1]); n = int(sys.argv[2]); m = int(sys.argv[3])
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

def t(fn, reps=10):
    for _ in range(2): fn()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    torch.cuda.synchronize()
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return sorted(a.elapsed_time(b) for a, b in ev)[reps // 2]

keep = E.TC_MIN_N
E.TC_MIN_N = 1 << 60
tg0 = t(lambda: eng.gradient(params, obs, act, w))
t0 = t(lambda: eng.fvp(params, obs, vec, out, cache_mode=0), 5)
r0 = out.cpu().numpy().copy()
E.TC_MIN_N = 0
tg1 = t(lambda: eng.gradient(params, obs, act, w))
t4 = t(lambda: eng.fvp(params, obs, vec, out, cache_mode=4))
r4 = out.cpu().numpy()
print("N=%d n=%d m=%d d=%d: gradient %.3f ms (plain) %.3f ms (+planes, x~ image); product tile kernel %.3f ms, tcgen05 wide %.3f ms; rel %.2e"
      % (N, n, m, d, tg0, tg1, t0, t4, float(np.linalg.norm(r4 - r0) / np.linalg.norm(r0))))
