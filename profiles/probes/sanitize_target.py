"""Small workload for compute-sanitizer (memcheck / racecheck / synccheck): one launch each of the kernels with the
race-prone constructs -- the tensor-memory FVP (mbarrier hand-offs between row, loader and issuer warps, tcgen05
commit), the Jacobian-cached rows FVP (named-barrier half-CTAs), the gradient pass in both cache layouts, and a short
MLPBaseline Adam chain (8-CTA cluster, DSMEM exchanges).  The cross-GPU peer exchange is left out: its kernels spin
on flags written by other launches and must not be serialised by a tool.
Usage: compute-sanitizer --tool memcheck python profiles/probes/sanitize_target.py"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from mjrl_b200.policies.gaussian_mlp import MLP
from mjrl_b200.utils.gym_env import EnvSpec
from mjrl_b200.algos.engine import UpdateEngine
from mjrl_b200.algos import engine

N, n, m = 300, 11, 3
pol = MLP(EnvSpec(n, m, 100, 1), hidden_sizes=(64, 64), seed=5)
eng = UpdateEngine(pol); dev = eng.device
rng = np.random.default_rng(0)
obs = torch.from_numpy(rng.standard_normal((N, n)).astype(np.float32)).to(dev)
act = torch.from_numpy(rng.standard_normal((N, m)).astype(np.float32)).to(dev)
w = torch.from_numpy(rng.standard_normal(N).astype(np.float32)).to(dev)
params = eng.dv.params("new"); d = params.numel()
vec = torch.randn(d, dtype=torch.float32, device=dev) * 0.1
out = torch.zeros(d, dtype=torch.float64, device=dev)
eng.cg_flags.zero_()
res = {}
for mode in (1, 2):
    eng.fvp(params, obs, vec, out, cache_mode=mode); res[mode] = out.cpu().numpy().copy()
assert eng._jac_prepare(params, obs)
eng.fvp(params, obs, vec, out, cache_mode=3); res[3] = out.cpu().numpy().copy()
assert eng._tc_prepare(params, obs)
eng.fvp(params, obs, vec, out, cache_mode=4); res[4] = out.cpu().numpy().copy()
engine.TC_MIN_N = 0
eng.gradient(params, obs, act, w)                       # cache_mode 5: tile-ordered h1|h2
eng.fvp(params, obs, vec, out, cache_mode=4); res[5] = out.cpu().numpy().copy()
for k in (2, 3, 4, 5):
    r = np.linalg.norm(res[k] - res[1]) / np.linalg.norm(res[1])
    assert r < 1e-5, (k, r)
# MLPBaseline: a short Adam chain on the cluster kernel
from mjrl_b200.baselines.mlp_baseline import MLPBaseline
from mjrl_b200.utils import synthetic as syn
flat = pol.get_param_values()
hb = syn.make_flat_batch(640, n, m, 100, 3, flat, (64, 64))
paths = syn.flat_to_paths(hb)
from mjrl_b200.utils import process_samples as ps
ps.compute_returns(paths, 0.995)
torch.manual_seed(0); np.random.seed(0)
mb = MLPBaseline(EnvSpec(n, m, 100, 1), epochs=1)
mb.fit(paths)
torch.cuda.synchronize()
print("SANITIZE_TARGET_OK")
