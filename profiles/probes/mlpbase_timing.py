"""Times MLPBaseline.predict / fit (csrc/mlpbase.cu) on a Hopper-shaped batch.
Usage on the GPU box: python profiles/probes/mlpbase_timing.py [N]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from mjrl_b200.utils import synthetic as syn
from mjrl_b200.utils.gym_env import EnvSpec
from mjrl_b200.baselines.mlp_baseline import MLPBaseline
from mjrl_b200.batch import PathBatch
from mjrl_b200.utils import process_samples as ps

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
n, m, hidden, horizon, _ = syn.SHAPES["C3-hopper"]
zero = np.zeros(n * 4 + 4 + 4 * 4 + 4 + 4 * m + m + m, dtype=np.float32)
hb = syn.make_flat_batch(N, n, m, horizon, 1, zero, (4, 4))
b = PathBatch(hb["observations"], hb["actions"], hb["rewards"], hb["path_offsets"], hb["terminated"])
ps.returns_on_device(b, 0.995)
torch.manual_seed(0); np.random.seed(0)
base = MLPBaseline(EnvSpec(n, m, horizon, 1), epochs=1)
ev = lambda: torch.cuda.Event(enable_timing=True)
for it in range(3):
    e0, e1, e2 = ev(), ev(), ev()
    e0.record(); base.predict_batch(b); e1.record()
    errs = base.fit_batch(b, return_errors=(it == 2), sync=False); e2.record()
    torch.cuda.synchronize()
    print("iter %d: predict %.3f ms (%.1f TFLOP/s fp32)  fit %d steps %.1f ms (%.2f us/step)  errors %s"
          % (it, e0.elapsed_time(e1), 36.9e3 * N / e0.elapsed_time(e1) / 1e9, base.last_fit_steps,
             e1.elapsed_time(e2), 1e3 * e1.elapsed_time(e2) / max(base.last_fit_steps, 1), errs))
