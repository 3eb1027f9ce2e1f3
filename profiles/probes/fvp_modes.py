"""Per-launch time of every Fisher-vector-product variant at the headline shape (CUDA events, median of 20,
L2-sized inputs): cache_mode 1 (fill), 2 (activation cache), 3 (Jacobian rows), 4 (tcgen05 / tensor memory),
plus the fill passes.  Usage: python profiles/probes/fvp_modes.py [N] [n] [m]"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from mjrl_b200.policies.gaussian_mlp import MLP
from mjrl_b200.utils.gym_env import EnvSpec
from mjrl_b200.algos.engine import UpdateEngine

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
n = int(sys.argv[2]) if len(sys.argv) > 2 else 11
m = int(sys.argv[3]) if len(sys.argv) > 3 else 3
pol = MLP(EnvSpec(n, m, 1000, 1), hidden_sizes=(64, 64), seed=123)
eng = UpdateEngine(pol); dev = eng.device
rng = np.random.default_rng(0)
obs = torch.from_numpy((3 * rng.standard_normal((N, n))).astype(np.float32)).to(dev)
params = eng.dv.params("new"); d = params.numel()
vec = torch.randn(d, dtype=torch.float32, device=dev) * 0.01
out = torch.empty(d, dtype=torch.float64, device=dev)
eng.cg_flags.zero_()

def t(fn, reps=20):
    for _ in range(3): fn()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    torch.cuda.synchronize()
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return sorted(a.elapsed_time(b) for a, b in ev)[reps // 2]

res = {}
res["mode1_fill"] = t(lambda: eng.fvp(params, obs, vec, out, cache_mode=1))
res["mode2_act_cache"] = t(lambda: eng.fvp(params, obs, vec, out, cache_mode=2))
r2 = out.cpu().numpy().copy()
if eng._jac_prepare(params, obs):
    def jf(): eng._jac_key = None; eng._jac_prepare(params, obs)
    res["jac_fill"] = t(jf, 5)
    res["mode3_jacobian"] = t(lambda: eng.fvp(params, obs, vec, out, cache_mode=3))
if eng._tc_prepare(params, obs):
    def tf(): eng._tc_key = None; eng._tc_prepare(params, obs)
    res["tc_fill"] = t(tf, 5)
    res["mode4_tcgen05"] = t(lambda: eng.fvp(params, obs, vec, out, cache_mode=4))
    r4 = out.cpu().numpy()
    res["rel_mode4_vs_mode2"] = float(np.linalg.norm(r4 - r2) / np.linalg.norm(r2))
print("N=%d n=%d m=%d d=%d" % (N, n, m, d))
for k, v in res.items():
    print("  %-22s %.4g" % (k, v) + (" ms" if not k.startswith("rel") else ""))
