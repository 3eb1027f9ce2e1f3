"""Multi-GPU sharding of the update: every rank owns WHOLE paths.

Every per-timestep quantity depends only on its own path and every global quantity is a sum
over timesteps, so the data path needs no collective; only the small fp64 sums
(advantage moments, Gram, gradient+surrogate, each FVP, surrogate/KL) are all-reduced
(algos/engine.py: Dist).  One process per GPU, launched by torchrun.
"""
import os

import numpy as np


def partition_paths(lengths, world):
    """Greedy longest-first assignment of whole paths to `world` ranks balancing timesteps.
    Returns a list of index arrays (ascending inside each rank, so path order is stable)."""
    lengths = np.asarray(lengths, dtype=np.int64)
    order = np.argsort(-lengths, kind="stable")
    load = np.zeros(world, dtype=np.int64)
    owner = np.empty(len(lengths), dtype=np.int64)
    for i in order:
        r = int(np.argmin(load))
        owner[i] = r
        load[r] += lengths[i]
    return [np.flatnonzero(owner == r) for r in range(world)]


def shard_paths(paths, rank, world):
    if world == 1:
        return paths
    idx = partition_paths([len(p["rewards"]) for p in paths], world)[rank]
    return [paths[i] for i in idx]


def init_distributed(backend=None):
    """torch.distributed from the torchrun environment; returns (rank, world, local_rank)."""
    import torch
    import torch.distributed as td
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not td.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        backend = backend or ("nccl" if torch.cuda.is_available() else "gloo")
        if backend == "nccl":
            torch.cuda.set_device(local)
            td.init_process_group(backend, device_id=torch.device("cuda", local))
        else:
            td.init_process_group(backend)
    return rank, world, local
