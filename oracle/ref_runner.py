"""TEST / BENCH INFRASTRUCTURE -- drives the reference's OWN update code (staged in oracle/_ref by
oracle/build_ref.py) on synthetic batches: the CPU baseline of record of BASELINE.md section 3.

What runs is exactly what `BatchREINFORCE.train_step` does after sampling (mjrl/algos/batch_reinforce.py:86-101)
with the reference's objects: `process_samples.compute_returns`, `compute_advantages`, `NPG.train_from_paths`
(npg_cg.py:84-165) and `baseline.fit`.  `gym` and `matplotlib` are absent here and on the GPU box; the
reference's algos import them at module level (samplers, logger), so empty stub modules stand in
(SURVEY 8(c)); no stubbed function is ever called with `env=None, save_logs=False`."""
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


def available():
    return os.path.isdir(os.path.join(REF_DIR, "mjrl", "algos"))


_R = None


def load():
    """Import the staged reference (once).  Keeps CUDA_VISIBLE_DEVICES as it was: importing
    mjrl.baselines.mlp_baseline overwrites it (mlp_baseline.py:1-5)."""
    global _R
    if _R is not None:
        return _R
    if not available():
        raise RuntimeError("oracle/_ref is not staged: run python oracle/build_ref.py where /root/reference exists")
    for name in ("gym", "matplotlib", "matplotlib.pyplot"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["matplotlib"].use = lambda *a, **k: None
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    saved = {k: os.environ.get(k) for k in ("CUDA_VISIBLE_DEVICES", "CUDA_DEVICE_ORDER")}
    sys.path.insert(0, REF_DIR)
    try:
        import mjrl.utils.process_samples as ps
        import mjrl.baselines.linear_baseline as lb
        import mjrl.policies.gaussian_mlp as gm
        import mjrl.policies.gaussian_linear as gl
        import mjrl.algos.npg_cg as npg
        import mjrl.algos.trpo as trpo
        from mjrl.utils.gym_env import EnvSpec
    finally:
        sys.path.remove(REF_DIR)
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    _R = types.SimpleNamespace(ps=ps, lb=lb, gm=gm, gl=gl, npg=npg, trpo=trpo, EnvSpec=EnvSpec)
    return _R


def make_agent(n, m, hidden, horizon, flat, step=0.01, cg_iters=10, damping=1e-4, algo="npg"):
    """Reference policy + LinearBaseline + NPG/TRPO agent with the given flat parameters."""
    R = load()
    spec = R.EnvSpec(n, m, horizon, 1)
    pol = R.gm.MLP(spec, hidden_sizes=tuple(hidden), seed=123) if hidden else R.gl.LinearPolicy(spec, seed=123)
    pol.set_param_values(flat.copy(), set_new=True, set_old=True)
    base = R.lb.LinearBaseline(spec)
    args = dict(FIM_invert_args={"iters": cg_iters, "damping": damping}, save_logs=False)
    if algo == "trpo":
        agent = R.trpo.TRPO(None, pol, base, kl_dist=0.5 * step, **args)
    else:
        agent = R.npg.NPG(None, pol, base, normalized_step_size=step, **args)
    return pol, base, agent


def full_iteration(agent, paths, gamma, gae_lambda):
    """batch_reinforce.py:86-101 after sampling; `paths` is the reference's list of dicts (mutated)."""
    R = load()
    R.ps.compute_returns(paths, gamma)
    R.ps.compute_advantages(paths, agent.baseline, gamma, gae_lambda)
    stats = agent.train_from_paths(paths)
    agent.baseline.fit(paths)
    return stats
