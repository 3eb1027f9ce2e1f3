"""Per-phase cycle counts of the MLP kernels (debug instrumentation, MJRL_B200_PHASE_CYCLES=1).
Usage on the GPU box: MJRL_B200_PHASE_CYCLES=1 [MJRL_B200_POLICY_PATH=fma] python profiles/probes/phase_cycles.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from mjrl_b200.utils import synthetic as syn
from mjrl_b200.utils.gym_env import EnvSpec
from mjrl_b200.policies.gaussian_mlp import MLP
from mjrl_b200.baselines.linear_baseline import LinearBaseline
from mjrl_b200.algos.npg_cg import NPG
from mjrl_b200.batch import PathBatch

shape = sys.argv[1] if len(sys.argv) > 1 else "C3-hopper"
n, m, hidden, horizon, N = syn.SHAPES[shape]
N = min(N, 1_000_000)
spec = EnvSpec(n, m, horizon, 1)
pol = MLP(spec, hidden_sizes=hidden, seed=123)
hb = syn.make_flat_batch(N, n, m, horizon, 1, pol.get_param_values(), hidden)
b = PathBatch(hb["observations"], hb["actions"], hb["rewards"], hb["path_offsets"], hb["terminated"])
ag = NPG(None, pol, LinearBaseline(spec))
eng = ag._engine()
p = eng.dv.params("new")
w = torch.randn(N, device=eng.device)
print("== grad"); eng.gradient(p, b.obs, b.act, w)
ll, mean = eng.old_cache(p, b.obs, b.act)
print("== eval"); eng.evaluate(p, b.obs, b.act, w, ll, mean, p[-m:], float(N))
print("== fvp"); eng.cg_flags.zero_(); eng.fvp(p, b.obs, eng.cg_p32, eng.cg_z)
torch.cuda.synchronize()
