"""Run under torchrun (one rank per GPU): sharded NPG update must equal the single-GPU update.
Used by tests/test_gpu_multi.py; rank 0 prints 'MULTI_OK <rel>'."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))

from conftest import load_golden, batch_of, rel_l2  # noqa: E402
from mjrl_b200.sharding import init_distributed, shard_paths  # noqa: E402
from mjrl_b200.utils import synthetic as syn  # noqa: E402
from mjrl_b200.utils.gym_env import EnvSpec  # noqa: E402
from mjrl_b200.policies.gaussian_mlp import MLP  # noqa: E402
from mjrl_b200.baselines.linear_baseline import LinearBaseline  # noqa: E402
from mjrl_b200.algos.npg_cg import NPG  # noqa: E402
from mjrl_b200.algos.engine import Dist  # noqa: E402
from mjrl_b200.batch import drop_cache  # noqa: E402


def run(world_mode):
    g = load_golden("c3_hopper_mid")
    spec = EnvSpec(g["n"], g["m"], g["horizon"], 1)
    pol = MLP(spec, hidden_sizes=g["hidden"], seed=123)
    pol.model.set_transformations(**g["transforms"]); pol.old_model.set_transformations(**g["transforms"])
    pol.set_param_values(g["params"].copy())
    base = LinearBaseline(spec)
    base._coeffs = g["fit0_coeffs"].copy()
    paths = syn.flat_to_paths(batch_of(g, "b1_"))
    # the full 10 CG iterations: shards own whole paths and DIFFERENT timestep counts, every mean
    # divides by the all-reduced global count, and only the fp32 summation order differs
    ag = NPG(None, pol, base, normalized_step_size=0.01, FIM_invert_args={"iters": 10, "damping": 1e-4})
    if world_mode:
        ag.dist = Dist()
        paths = shard_paths(paths, ag.dist.rank, ag.dist.world)
    else:
        ag.dist = Dist.__new__(Dist); ag.dist.td = None; ag.dist.world = 1; ag.dist.rank = 0; ag.dist.calls = 0
        ag.dist.group = None
    drop_cache()
    ag.train_step(len(paths), gamma=g["gamma"], gae_lambda=g["gae_lambda"], paths=paths)
    global PEER_NOTE
    if world_mode:
        PEER_NOTE = "%s, %d exchanges" % (ag.dist.peer_note, ag.dist.peer.exchanges if ag.dist.peer else 0)
    return pol.get_param_values().astype(np.float64) - g["params"], np.asarray(base._coeffs), g


PEER_NOTE = "-"


def run_trpo_dapg(world_mode):
    """TRPO (line search: one exchange per trial) and DAPG (demo term on top of the sharded gradient) under
    the same sharding: updated parameters and the backtrack count must not depend on the rank count."""
    from mjrl_b200.algos.trpo import TRPO
    from mjrl_b200.algos.dapg import DAPG
    g = load_golden("c3_hopper_mid")
    spec = EnvSpec(g["n"], g["m"], g["horizon"], 1)
    out = {}
    for name in ("trpo", "dapg"):
        pol = MLP(spec, hidden_sizes=g["hidden"], seed=123)
        pol.model.set_transformations(**g["transforms"]); pol.old_model.set_transformations(**g["transforms"])
        pol.set_param_values(g["params"].copy())
        base = LinearBaseline(spec)
        base._coeffs = g["fit0_coeffs"].copy()
        paths = syn.flat_to_paths(batch_of(g, "b1_"))
        if name == "trpo":
            ag = TRPO(None, pol, base, kl_dist=0.002, FIM_invert_args={"iters": 10, "damping": 1e-4})
        else:
            demos = syn.flat_to_paths(batch_of(g, "b1_"))[:3]      # replicated on every rank
            ag = DAPG(None, pol, base, demo_paths=demos, kl_dist=0.01, lam_0=1.0, lam_1=0.95,
                      FIM_invert_args={"iters": 10, "damping": 1e-4})
        if world_mode:
            ag.dist = Dist()
            paths = shard_paths(paths, ag.dist.rank, ag.dist.world)
        else:
            ag.dist = Dist.__new__(Dist); ag.dist.td = None; ag.dist.world = 1; ag.dist.rank = 0; ag.dist.calls = 0
            ag.dist.group = None
        drop_cache()
        ag.train_step(len(paths), gamma=g["gamma"], gae_lambda=g["gae_lambda"], paths=paths)
        out[name] = (pol.get_param_values().astype(np.float64) - g["params"],
                     len(ag.last_update.get("trials", [])) if name == "trpo" else 0)
    return out


def run_ppo(world_mode, rank, world):
    """PPO: the minibatch chain is replicas-only (every rank gathers all rows in rank order, rank 0 draws the
    minibatches), the full-batch passes are sharded.  Must equal a single-GPU run on the shards concatenated
    in rank order with the same numpy seed, up to the summation order of the whitening moments."""
    from mjrl_b200.algos.ppo_clip import PPO
    g = load_golden("c3_hopper_mid")
    spec = EnvSpec(g["n"], g["m"], g["horizon"], 1)
    pol = MLP(spec, hidden_sizes=g["hidden"], seed=123)
    pol.model.set_transformations(**g["transforms"]); pol.old_model.set_transformations(**g["transforms"])
    pol.set_param_values(g["params"].copy())
    base = LinearBaseline(spec)
    base._coeffs = g["fit0_coeffs"].copy()
    paths = syn.flat_to_paths(batch_of(g, "b1_"))
    shards = [shard_paths(paths, r, world) for r in range(world)]
    ag = PPO(None, pol, base, clip_coef=0.2, epochs=2, mb_size=64, learn_rate=3e-4)
    np.random.seed(77)
    if world_mode:
        ag.dist = Dist()
        mine = shards[rank]
    else:
        ag.dist = Dist.__new__(Dist); ag.dist.td = None; ag.dist.world = 1; ag.dist.rank = 0; ag.dist.calls = 0
        ag.dist.group = None
        mine = [p for sh in shards for p in sh]
    drop_cache()
    ag.train_step(len(mine), gamma=g["gamma"], gae_lambda=g["gae_lambda"], paths=mine)
    return pol.get_param_values().astype(np.float64), ag.last_update


def run_mlpbase(world_mode, rank, world):
    """MLPBaseline.fit is replicas-only: every rank gathers all rows and runs the same chain, so the
    result must equal (bit for bit) a single-GPU fit on the shards concatenated in rank order."""
    from mjrl_b200.baselines.mlp_baseline import MLPBaseline
    from mjrl_b200.utils import process_samples as ps
    from mjrl_b200.batch import batch_for
    g = load_golden("c3_hopper_mid")
    spec = EnvSpec(g["n"], g["m"], g["horizon"], 1)
    paths = syn.flat_to_paths(batch_of(g, "b1_"))
    shards = [shard_paths(paths, r, world) for r in range(world)]
    torch.manual_seed(5); np.random.seed(11)
    mb = MLPBaseline(spec, epochs=1)
    drop_cache()
    if world_mode:
        b = batch_for(shards[rank])
        ps.returns_on_device(b, g["gamma"])
        errs = mb.fit_batch(b, return_errors=True, dist=Dist())
    else:
        b = batch_for([p for sh in shards for p in sh])
        ps.returns_on_device(b, g["gamma"])
        errs = mb.fit_batch(b, return_errors=True)
    return mb.get_param_values(), errs, mb.last_fit_steps


if __name__ == "__main__":
    rank, world, local = init_distributed("nccl")
    torch.cuda.set_device(local)
    step_multi, coef_multi, g = run(True)
    mlp_multi, errs_multi, steps_multi = run_mlpbase(True, rank, world)
    td_multi = run_trpo_dapg(True)
    ppo_multi, ppo_lu = run_ppo(True, rank, world)
    # replicas: every rank must hold the same parameters bit for bit
    gathered = [None] * world
    torch.distributed.all_gather_object(gathered, ppo_multi.tobytes())
    ppo_replicas_same = all(b == gathered[0] for b in gathered)
    if rank == 0:
        ppo_one, ppo_lu1 = run_ppo(False, rank, world)
        d_ppo = float(np.abs(ppo_multi - ppo_one).max())
        ppo_ok = ppo_replicas_same and d_ppo < 2e-5 and ppo_lu["steps"] == ppo_lu1["steps"] and \
            abs(ppo_lu["kl_dist"] - ppo_lu1["kl_dist"]) < 1e-5
        print("PPO_%s max|dp|=%.3e steps=%d replicas_bitexact=%s" % ("OK" if ppo_ok else "FAIL", d_ppo,
                                                                     ppo_lu["steps"], ppo_replicas_same))
        td_one = run_trpo_dapg(False)
        r_trpo = rel_l2(td_multi["trpo"][0], td_one["trpo"][0])
        r_dapg = rel_l2(td_multi["dapg"][0], td_one["dapg"][0])
        td_ok = r_trpo < 1e-4 and r_dapg < 1e-4 and td_multi["trpo"][1] == td_one["trpo"][1]
        print("TRPO_DAPG_%s trpo_rel=%.3e (%d trials) dapg_rel=%.3e" % ("OK" if td_ok else "FAIL", r_trpo,
                                                                         td_multi["trpo"][1], r_dapg))
        step_one, coef_one, _ = run(False)
        mlp_one, errs_one, steps_one = run_mlpbase(False, rank, world)
        r_step, r_coef = rel_l2(step_multi, step_one), rel_l2(coef_multi, coef_one)
        mlp_same = steps_multi == steps_one and np.array_equal(mlp_multi, mlp_one)
        errs_ok = all(abs(a - b) < 1e-9 * max(1.0, abs(b)) for a, b in zip(errs_multi, errs_one))
        ok = r_step < 1e-4 and r_coef < 1e-6 and mlp_same and errs_ok and td_ok and ppo_ok
        print("MULTI_%s world=%d step_rel=%.3e coef_rel=%.3e mlpbase_bitexact=%s (%d steps) errs_ok=%s peer=[%s]"
              % ("OK" if ok else "FAIL", world, r_step, r_coef, mlp_same, steps_multi, errs_ok, PEER_NOTE))
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()
