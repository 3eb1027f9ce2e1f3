"""Host side of csrc/peer.cu: one symmetric device block per rank, mapped by every peer through
CUDA IPC, over which the small fp64 all-reduces of the sharded update run as single
kernels instead of NCCL launches, and the CG update runs fused with its all-reduce.

One process per GPU (torchrun).  `torch.distributed` carries only the 64-byte IPC handles at start-up
and the agreement that every rank mapped every peer; after that the data path never touches NCCL.
"""
import ctypes
import os

import torch

from . import _lib
from ._lib import c_double, c_i64, c_int, c_void_p, ptr, stream_ptr


class PeerGroup:
    MIN_COUNT = 16384      # covers d+1 of every policy shape of BASELINE.json and Ant's Gram (13,341)

    def __init__(self, ctx, rank, world, max_count):
        self.ctx, self.rank, self.world = ctx, int(rank), int(world)
        self.max_count = int(max(max_count, self.MIN_COUNT))
        self.lib = ctx.lib
        self.lib.mjrl_peer_slot.restype = c_void_p
        self.lib.mjrl_peer_base.restype = c_void_p
        self.h = c_void_p()
        nb = int(self.lib.mjrl_peer_handle_bytes())
        self.handle = (ctypes.c_char * nb)()
        _lib.check(self.lib.mjrl_peer_create(ctx.h, c_int(self.rank), c_int(self.world), c_i64(self.max_count),
                                             ctypes.byref(self.h), self.handle), "mjrl_peer_create")
        self.epoch = 0
        self.exchanges = 0

    # ---- wiring ------------------------------------------------------------------------
    def connect_ipc(self, handles):
        """handles: list of `world` bytes objects in rank order (all_gather of `bytes(self.handle)`)."""
        blob = b"".join(bytes(h) for h in handles)
        buf = ctypes.create_string_buffer(blob, len(blob))
        _lib.check(self.lib.mjrl_peer_connect(self.h, buf), "mjrl_peer_connect")

    def connect_local(self, groups):
        """Same-process ranks (tests): `groups` = the PeerGroup of every rank, in rank order."""
        arr = (c_void_p * self.world)(*[c_void_p(self.lib.mjrl_peer_base(g.h)) for g in groups])
        _lib.check(self.lib.mjrl_peer_connect_ptrs(self.h, arr), "mjrl_peer_connect_ptrs")

    def set_timeout(self, seconds):
        _lib.check(self.lib.mjrl_peer_set_timeout(self.h, c_double(seconds)), "mjrl_peer_set_timeout")

    def close(self):
        if self.h:
            self.lib.mjrl_peer_destroy(self.h)
            self.h = c_void_p()

    # ---- exchanges (every rank calls the same sequence) ------------------------------------
    def next_slot(self):
        """Advance the epoch and return (epoch, device pointer of the local contribution slot): the
        producer of the next exchange may write its vector there instead of having it copied in."""
        self.epoch += 1
        return self.epoch, c_void_p(self.lib.mjrl_peer_slot(self.h, c_i64(self.epoch)))

    def takes(self, t):
        return (t.is_cuda and t.dtype == torch.float64 and t.is_contiguous() and 0 < t.numel() <= self.max_count)

    def allreduce(self, t):
        """In-place sum over ranks of a contiguous fp64 device tensor."""
        self.epoch += 1
        self.ctx.call("mjrl_peer_allreduce_f64", self.h, ptr(t), ptr(t), c_i64(t.numel()), c_i64(self.epoch),
                      stream_ptr())
        self.exchanges += 1
        return t

    def cg_update(self, cg_state, d, inv_count, damping, tol, epoch):
        """Fused all-reduce + CG update; the local product sum is in the slot `next_slot` returned."""
        self.ctx.call("mjrl_cg_update_peer", self.h, ctypes.byref(cg_state), c_i64(d), c_double(inv_count),
                      c_double(damping), c_double(tol), c_i64(epoch), stream_ptr())
        self.exchanges += 1

    def check(self):
        """Synchronous: raise if a wait on a peer timed out since creation."""
        err = ctypes.c_int32(0)
        _lib.check(self.lib.mjrl_peer_status(self.h, ctypes.byref(err)), "mjrl_peer_status")
        if err.value:
            raise _lib.MjrlError("peer exchange: a rank did not arrive within the timeout; results are invalid")


def connect_process_group(ctx, td, group, max_count):
    """Create this rank's block, exchange the IPC handles over `td` and map every peer.  Returns a
    PeerGroup, or None when ANY rank could not map its peers (all ranks agree on the outcome)."""
    rank, world = td.get_rank(group), td.get_world_size(group)
    dev = torch.device("cuda", ctx.device)
    pg, ok, why = None, 1, ""
    try:
        pg = PeerGroup(ctx, rank, world, max_count)
        handles = [None] * world
        td.all_gather_object(handles, bytes(pg.handle), group=group)
        pg.connect_ipc(handles)
    except Exception as e:       # noqa: BLE001 - any failure means "no peer path on this box"
        ok, why = 0, str(e)
        if pg is None:           # keep the collective sequence identical on every rank
            td.all_gather_object([None] * world, b"", group=group)
    flag = torch.tensor([ok], dtype=torch.int32, device=dev)
    td.all_reduce(flag, op=td.ReduceOp.MIN, group=group)
    if int(flag.item()) == 0:
        if pg is not None:
            pg.close()
        return None, why or "a peer rank could not map the blocks"
    td.barrier(group=group)      # every block is zeroed and mapped before the first flag is raised
    # ranks run their host code independently between exchanges (data staging, Python jitter, 8 processes
    # sharing the host cores): give a late peer a minute before declaring the update invalid
    pg.set_timeout(float(os.environ.get("MJRL_B200_PEER_TIMEOUT", "60")))
    return pg, ""


def mode():
    """MJRL_B200_PEER = 0 (NCCL only) | 1 (required: raise when unavailable) | auto (default)."""
    return os.environ.get("MJRL_B200_PEER", "auto").strip().lower()
