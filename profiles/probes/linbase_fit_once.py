"""LinearBaseline fit (Gram + solve) on a synthetic batch of a given observation width (for ncu).  Usage: linbase_fit_once.py N n"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from mjrl_b200.utils import synthetic as syn
from mjrl_b200.utils.gym_env import EnvSpec
from mjrl_b200.baselines.linear_baseline import LinearBaseline
from mjrl_b200.batch import PathBatch
from mjrl_b200.utils import process_samples as ps
N, n = int(sys.argv[1]), int(sys.argv[2])
m, horizon = 3, 1000
zero = np.zeros(n * 4 + 4 + 4 * 4 + 4 + 4 * m + m + m, dtype=np.float32)
hb = syn.make_flat_batch(N, n, m, horizon, 1, zero, (4, 4))
b = PathBatch(hb["observations"], hb["actions"], hb["rewards"], hb["path_offsets"], hb["terminated"])
ps.returns_on_device(b, 0.995)
base = LinearBaseline(EnvSpec(n, m, horizon, 1))
ev = lambda: torch.cuda.Event(enable_timing=True)
for it in range(3):
    e0, e1 = ev(), ev()
    e0.record(); base.fit_batch(b, sync=False); e1.record()
    torch.cuda.synchronize()
    print("fit %d: %.3f ms" % (it, e0.elapsed_time(e1)))
