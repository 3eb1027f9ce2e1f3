"""Per-block error of the wide tcgen05 product (csrc/fvp_tc_wide.cu) against the tile kernel and the fp64 truth."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "tests"))
import torch
from mjrl_b200.utils import synthetic as syn
from mjrl_b200.policies.gaussian_mlp import MLP
from mjrl_b200.utils.gym_env import EnvSpec
from mjrl_b200.algos import engine as E
from oracle import mjrl_oracle as O


def rel(a, b):
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def run(n, m, N, mid):
    rng = np.random.default_rng(n * 100 + N)
    pol = MLP(EnvSpec(n, m, 100, 1), hidden_sizes=(64, 64), seed=5)
    flat = pol.get_param_values()
    tr = None
    if mid:
        flat, tr = syn.mid_training_variant(flat, n, m, 77)
        pol.model.set_transformations(**tr); pol.old_model.set_transformations(**tr)
    else:
        flat = flat + 0.1 * rng.standard_normal(flat.shape).astype(np.float32)
    pol.set_param_values(flat.copy(), True, True)
    eng = E.UpdateEngine(pol)
    dev = eng.device
    obs = torch.from_numpy((2 * rng.standard_normal((N, n))).astype(np.float32)).to(dev)
    act = torch.from_numpy(rng.standard_normal((N, m)).astype(np.float32)).to(dev)
    w = torch.from_numpy(rng.standard_normal(N).astype(np.float32)).to(dev)
    params = eng.dv.params("new")
    E.TC_MIN_N = 0
    eng.gradient(params, obs, act, w)
    vec = torch.from_numpy(rng.standard_normal(flat.shape[0]).astype(np.float32) * 0.1).to(dev)

    def product(mode):
        out = torch.zeros(flat.shape[0], dtype=torch.float64, device=dev)
        eng.cg_flags.zero_()
        eng.fvp(params, obs, vec, out, cache_mode=mode)
        return out.cpu().numpy()
    ref, a = product(0), product(4)
    F = [0, 64 * n, 64 * n + 64, 64 * n + 64 + 4096, 64 * n + 128 + 4096, 64 * n + 128 + 4096 + 64 * m,
         64 * n + 128 + 4096 + 64 * m + m]
    names = ["W0", "b0", "W1", "b1", "W2", "b2"]
    print((n, m, N, mid), "total %.2e" % rel(a, ref), " ".join("%s %.1e" % (nm, rel(a[lo:hi], ref[lo:hi])) for nm, lo, hi in zip(names, F[:-1], F[1:])))
    if rel(a[:64 * n], ref[:64 * n]) > 1e-4:
        A, R = a[:64 * n].reshape(64, n), ref[:64 * n].reshape(64, n)
        colerr = np.abs(A - R).max(axis=0) / (np.abs(R).max() + 1e-30)
        print("   W0 column errors:", np.array2string(colerr, precision=1, max_line_width=250))


if __name__ == "__main__":
    for c in [(24, 3, 96, False), (24, 3, 96, True), (24, 3, 1000, False), (31, 1, 97, False), (40, 3, 96, False), (63, 16, 4097, True),
              (64, 9, 2049, False), (127, 24, 500, True), (111, 8, 1000, False)]:
        run(*c)
