"""HBM-bound kernels at sizes where launch latency no longer dominates: returns / GAE scans,
LinearBaseline predict, Gram, whitening, on N timesteps of Hopper-shaped paths (horizon 1000).
Usage on the GPU box: python profiles/probes/scan_bandwidth.py [N ...]"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np, torch
from mjrl_b200.batch import PathBatch
from mjrl_b200.utils import process_samples as ps
from mjrl_b200.utils.gym_env import EnvSpec
from mjrl_b200.baselines.linear_baseline import LinearBaseline
from mjrl_b200.algos.engine import UpdateEngine

peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
n, m, horizon = 11, 3, 1000
def timed(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps
for N in [int(x) for x in sys.argv[1:]] or [1_000_000, 4_000_000, 16_000_000, 64_000_000]:
    P = N // horizon
    rng = np.random.default_rng(0)
    obs = torch.randn(N, n, dtype=torch.float32).numpy(); rew = rng.standard_normal(N)
    off = np.arange(P + 1, dtype=np.int64) * horizon
    b = PathBatch(obs, None, rew, off, np.zeros(P, dtype=np.uint8))
    base = LinearBaseline(EnvSpec(n, m, horizon, 1))
    ps.returns_on_device(b, 0.995)
    base.fit_batch(b)
    rows = [("returns scan", timed(lambda: ps.returns_on_device(b, 0.995)), 16),
            ("linbase predict", timed(lambda: base.predict_batch(b)), 4 * n + 8)]
    bv = base.predict_batch(b)
    rows.append(("GAE scan", timed(lambda: ps.advantages_on_device(b, bv, 0.995, 0.97)), 24))
    mom = torch.empty(4, dtype=torch.float64, device=b.device)
    from mjrl_b200._lib import c_i64, ptr, stream_ptr
    rows.append(("moments (2 passes)", timed(lambda: b.ctx.call("mjrl_moments", ptr(b.adv), c_i64(N), ptr(mom), stream_ptr())), 16))
    rows.append(("linbase Gram+solve", timed(lambda: base.fit_batch(b, sync=False)), 4 * n + 8))
    print("N = %d timesteps (%d paths x %d), working set > L2: %s" % (N, P, horizon, N * 16 > 126e6))
    for name, ms, bpt in rows:
        gbs = N * bpt / (ms * 1e-3) / 1e9
        print("   %-20s %8.3f ms  %7.1f GB/s algorithmic = %4.1f %% of measured HBM peak (%.0f GB/s)" % (name, ms, gbs, 100 * gbs / peak, peak))
    del b, obs, rew
    torch.cuda.empty_cache()
