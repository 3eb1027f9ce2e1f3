"""world_size-2 gloo test (CPU) of the N>1 host logic: whole-path partitioning, and that the
sums each rank would hand to the all-reduce (Gram, advantage moments, gradient, FVP)
add up to the single-process result.  The per-rank partial sums are produced with the
oracle's fp64 maths -- the CUDA kernels produce the same partials on a GPU box."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import torch.distributed as td
    from conftest import load_golden, batch_of
    from mjrl_b200.sharding import init_distributed, shard_paths
    from mjrl_b200.algos.engine import Dist
    from mjrl_b200.utils import synthetic as syn
    from oracle import mjrl_oracle as O
    init_distributed("gloo")
    dist = Dist()
    g = load_golden("c3_hopper_mid")
    paths = syn.flat_to_paths(batch_of(g, "b1_"))
    off = g["b1_path_offsets"]
    for i, p in enumerate(paths):
        p["advantages"] = g["b1_adv"][off[i]:off[i + 1]]
        p["returns"] = g["b1_returns"][off[i]:off[i + 1]]
    mine = shard_paths(paths, rank, world)
    # (1) advantage moments: (N, sum) then centred squares, as engine.whiten does
    adv = np.concatenate([p["advantages"] for p in mine])
    part = torch.tensor([float(adv.size), adv.sum()], dtype=torch.float64)
    dist.allreduce(part)
    mean = part[1] / part[0]
    ss = torch.tensor([((adv - mean.item()) ** 2).sum()], dtype=torch.float64)
    dist.allreduce(ss)
    std = float(torch.sqrt(ss / part[0]))
    # (2) Gram of the LinearBaseline
    X = np.concatenate([O.linear_features(p) for p in mine])
    y = np.concatenate([p["returns"] for p in mine])
    gram = torch.from_numpy(np.concatenate([(X.T @ np.c_[X, y]).ravel(), [y @ y]]))
    dist.allreduce(gram)
    # (3) gradient and FVP sums with globally whitened advantages
    pol = O.GaussianPolicy(g["n"], g["m"], g["hidden"], g["params"], **g["transforms"])
    tp = O.TruthPolicy(pol)
    obs = np.concatenate([p["observations"] for p in mine])
    act = np.concatenate([p["actions"] for p in mine])
    w = (adv - mean.item()) / (std + 1e-6)
    gsum = torch.from_numpy(tp.vpg(obs, act, w) * obs.shape[0])
    dist.allreduce(gsum)
    fsum = torch.from_numpy(tp.fvp(obs, g["vpg_grad"]) * obs.shape[0])
    dist.allreduce(fsum)
    # (4) replicas-only chains (PPO minibatch epochs): all rows in rank order on every rank, rank 0's schedule
    rows, sizes = dist.allgather_rows(torch.from_numpy(np.c_[obs, act].astype(np.float32)))
    sched = torch.arange(6, dtype=torch.int32) * (7 if rank == 0 else 1)
    dist.broadcast0(sched)
    if rank == 0:
        q.put(dict(N=float(part[0]), mean=float(mean), std=std, gram=gram.numpy(), gsum=gsum.numpy(),
                   fsum=fsum.numpy(), calls=dist.calls, n_mine=len(mine), rows=rows.numpy(), sizes=sizes,
                   sched=sched.numpy()))
    else:
        q.put(dict(rank=rank, rows_sum=float(rows.double().sum()), sched=sched.numpy()))
    td.barrier()
    td.destroy_process_group()


def test_partition_balances_whole_paths():
    from mjrl_b200.sharding import partition_paths
    rng = np.random.default_rng(0)
    lens = np.r_[np.full(900, 1000), rng.integers(2, 1000, 100)]
    for world in (1, 2, 4, 8):
        parts = partition_paths(lens, world)
        allidx = np.sort(np.concatenate(parts))
        np.testing.assert_array_equal(allidx, np.arange(len(lens)))      # every path exactly once
        loads = np.array([lens[p].sum() for p in parts])
        assert loads.max() - loads.min() <= 1000                          # within one path
        for p in parts:
            assert np.all(np.diff(p) > 0)


@pytest.mark.timeout(300)
def test_two_rank_sums_match_single_process():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from conftest import load_golden, batch_of, rel_l2
    from mjrl_b200.utils import synthetic as syn
    from oracle import mjrl_oracle as O
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    got = [q.get(timeout=240), q.get(timeout=240)]
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    res = [r for r in got if "N" in r][0]
    other = [r for r in got if "N" not in r][0]
    g = load_golden("c3_hopper_mid")
    paths = syn.flat_to_paths(batch_of(g, "b1_"))
    off = g["b1_path_offsets"]
    adv = g["b1_adv"]
    assert res["N"] == adv.size and 0 < res["n_mine"] < len(paths)
    assert abs(res["mean"] - adv.mean()) < 1e-12 and abs(res["std"] - adv.std()) < 1e-12
    for i, p in enumerate(paths):
        p["returns"] = g["b1_returns"][off[i]:off[i + 1]]
    X = np.concatenate([O.linear_features(p) for p in paths])
    y = g["b1_returns"]
    full = np.concatenate([(X.T @ np.c_[X, y]).ravel(), [y @ y]])
    assert rel_l2(res["gram"], full) < 1e-12
    # the all-reduced gradient sum / N equals the reference's gradient (whitening eps 1e-6)
    assert rel_l2(res["gsum"] / adv.size, g["vpg_grad"]) < 1e-5
    assert rel_l2(res["fsum"] / adv.size, g["hvp_of_grad_nodamp"]) < 1e-5
    assert res["calls"] == 5
    # replicas-only chains: both ranks hold all rows, in rank order, and rank 0's schedule
    from mjrl_b200.sharding import shard_paths
    want = np.concatenate([np.c_[np.concatenate([p["observations"] for p in sh]),
                                 np.concatenate([p["actions"] for p in sh])]
                           for sh in (shard_paths(paths, r, 2) for r in range(2))]).astype(np.float32)
    np.testing.assert_array_equal(res["rows"], want)
    assert sum(res["sizes"]) == adv.size and abs(other["rows_sum"] - float(want.astype(np.float64).sum())) < 1e-6
    np.testing.assert_array_equal(res["sched"], np.arange(6) * 7)
    np.testing.assert_array_equal(other["sched"], np.arange(6) * 7)
