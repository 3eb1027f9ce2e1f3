"""Launch each HBM-bound kernel of the path a few times at N timesteps (for an ncu capture).
Usage: python profiles/probes/hbm_kernels_once.py [N] [horizon]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np, torch
from mjrl_b200.batch import PathBatch
from mjrl_b200.utils import process_samples as ps
from mjrl_b200.utils.gym_env import EnvSpec
from mjrl_b200.baselines.linear_baseline import LinearBaseline

N = int(sys.argv[1]) if len(sys.argv) > 1 else 16_000_000
horizon = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
n, m = 11, 3
P = N // horizon
rng = np.random.default_rng(0)
obs = torch.randn(N, n, dtype=torch.float32).numpy(); rew = rng.standard_normal(N)
off = np.arange(P + 1, dtype=np.int64) * horizon
b = PathBatch(obs, None, rew, off, np.zeros(P, dtype=np.uint8))
base = LinearBaseline(EnvSpec(n, m, horizon, 1))
for _ in range(2):
    ps.returns_on_device(b, 0.995)
    base.fit_batch(b)
    bv = base.predict_batch(b)
    ps.advantages_on_device(b, bv, 0.995, 0.97)
torch.cuda.synchronize()
print("done")
