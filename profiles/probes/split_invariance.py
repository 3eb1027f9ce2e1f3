"""Single-GPU emulation of the 2-rank sharding: gradient / FVP sums over the whole batch vs the
sum of two halves (what the all-reduce adds), then CG iterates of both.  Shows how much of the
multi-GPU difference is summation order and how CG amplifies it."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from conftest import load_golden, batch_of, rel_l2
from test_gpu_policy import make_policy, agent
from mjrl_b200.algos.npg_cg import NPG
from mjrl_b200.algos.batch_reinforce import _dev32

name = sys.argv[1] if len(sys.argv) > 1 else "c3_hopper_mid"
g = load_golden(name)
b1 = batch_of(g, "b1_")
pol = make_policy(g)
ag = agent(NPG, g, pol)
eng = ag._engine()
dev = eng.device
obs, act = _dev32(b1["observations"], dev), _dev32(b1["actions"], dev)
w = _dev32(g["b1_adv_white"], dev)
N = obs.shape[0]; h = (N // 2 // 64) * 64 + 17
p = eng.dv.params("new")
d = eng.dv.d
def grad(o, a, ww):
    eng.gradient(p, o, a, ww); return eng.grad[:d].clone()
g_all = grad(obs, act, w); g_split = grad(obs[:h], act[:h], w[:h]) + grad(obs[h:].contiguous(), act[h:].contiguous(), w[h:].contiguous())
print("gradient sum: whole vs halves rel %.2e" % rel_l2(g_split.cpu().numpy(), g_all.cpu().numpy()))
def fvp(o, v):
    out = torch.empty(d, dtype=torch.float64, device=dev); eng.cg_flags.zero_(); eng.fvp(p, o, v, out); return out
v = (g_all / N).float()
f_all = fvp(obs, v); f_split = fvp(obs[:h], v) + fvp(obs[h:].contiguous(), v)
print("FVP sum:      whole vs halves rel %.2e" % rel_l2(f_split.cpu().numpy(), f_all.cpu().numpy()))
# CG with each operator (fp64 vectors like the engine), damping 1e-4
def cg(op, b, iters):
    x = torch.zeros_like(b); r = b.clone(); pp = b.clone(); rr = r.dot(r); xs = []
    for _ in range(iters):
        z = op(pp.float()) / N + 1e-4 * pp
        a_ = rr / pp.dot(z); x = x + a_ * pp; r = r - a_ * z; rn = r.dot(r); pp = r + (rn / rr) * pp; rr = rn
        xs.append(x.clone())
    return xs
b = (g_all / N)
xa = cg(lambda q: fvp(obs, q), b, 10)
xs = cg(lambda q: fvp(obs[:h], q) + fvp(obs[h:].contiguous(), q), b, 10)
for k in (1, 2, 3, 5, 10):
    print("CG k=%2d: whole vs halves rel %.2e" % (k, rel_l2(xs[k - 1].cpu().numpy(), xa[k - 1].cpu().numpy())))
