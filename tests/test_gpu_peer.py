"""csrc/peer.cu on ONE GPU: W ranks living in this process (blocks wired with mjrl_peer_connect_ptrs,
one stream per rank) run the same exchange kernels the torchrun ranks run over NVLink.  Checks the
protocol (signal / wait / slot reuse over many epochs), rank-order determinism, the fused CG update
against all-reduce + mjrl_cg_update, and the timeout path.  The real multi-process run is
tests/test_gpu_multi.py."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def make_groups(world, max_count=0):
    from mjrl_b200 import _lib
    from mjrl_b200.peer import PeerGroup
    ctx = _lib.context()
    groups = [PeerGroup(ctx, r, world, max_count) for r in range(world)]
    for g in groups:
        g.connect_local(groups)
    return ctx, groups, [torch.cuda.Stream() for _ in range(world)]


@pytest.mark.parametrize("world", [2, 4, 8])
def test_allreduce_many_epochs_bit_identical_across_ranks(world):
    ctx, groups, streams = make_groups(world)
    dev = torch.device("cuda", ctx.device)
    rng = np.random.default_rng(world)
    try:
        for it, count in enumerate([1, 3, 5127, 13341, 16384, 2, 5127, 5127, 100, 16384]):
            host = [rng.standard_normal(count) * 10.0 ** rng.integers(-3, 4) for _ in range(world)]
            ts = [torch.from_numpy(h.copy()).to(dev) for h in host]
            torch.cuda.synchronize()
            for r in range(world):
                with torch.cuda.stream(streams[r]):
                    groups[r].allreduce(ts[r])
            torch.cuda.synchronize()
            want = np.zeros(count)
            for h in host:                      # rank order, like the kernel
                want = want + h
            for r in range(world):
                np.testing.assert_array_equal(ts[r].cpu().numpy(), want)
        for g in groups:
            g.check()
            assert g.epoch == 10
    finally:
        for g in groups:
            g.close()


def test_oversize_and_dtype_are_not_taken():
    ctx, groups, _ = make_groups(2)
    dev = torch.device("cuda", ctx.device)
    try:
        g = groups[0]
        assert g.takes(torch.zeros(16384, dtype=torch.float64, device=dev))
        assert not g.takes(torch.zeros(16385, dtype=torch.float64, device=dev))
        assert not g.takes(torch.zeros(8, dtype=torch.float32, device=dev))
        assert not g.takes(torch.zeros(8, dtype=torch.float64))
        assert not g.takes(torch.zeros(8, 2, dtype=torch.float64, device=dev)[:, 0])
    finally:
        for g in groups:
            g.close()


@pytest.mark.parametrize("d", [16, 5126, 8192, 11856])
def test_fused_cg_update_equals_allreduce_then_update(d):
    from mjrl_b200._lib import CGState, c_double, c_i64, c_void_p, ptr, stream_ptr
    world = 2
    ctx, groups, streams = make_groups(world, d + 1)
    dev = torch.device("cuda", ctx.device)
    rng = np.random.default_rng(d)

    def state():
        f64 = lambda: torch.zeros(d, dtype=torch.float64, device=dev)
        t = dict(x=f64(), r=f64(), p=f64(), z=f64(), scal=torch.zeros(8, dtype=torch.float64, device=dev),
                 p32=torch.zeros(d, dtype=torch.float32, device=dev),
                 flags=torch.zeros(2, dtype=torch.int32, device=dev))
        cg = CGState(*[c_void_p(t[k].data_ptr()) for k in ("x", "r", "p", "z", "scal", "p32", "flags")])
        return t, cg

    try:
        b = torch.from_numpy(rng.standard_normal(d)).to(dev)
        ref_t, ref_cg = state()
        rank_states = [state() for _ in range(world)]
        for t, cg in [(ref_t, ref_cg)] + rank_states:
            ctx.call("mjrl_cg_init", ctypes.byref(cg), c_i64(d), ptr(b), c_double(0.5), stream_ptr())
        torch.cuda.synchronize()
        for it in range(4):
            # an SPD operator z = A p split into two "shards": z_r = A_r p, A = A_0 + A_1 (diagonal here)
            diag = [torch.from_numpy(rng.uniform(0.5, 2.0, d)).to(dev) for _ in range(world)]
            parts = [diag[r] * ref_t["p"] for r in range(world)]
            ref_t["z"].copy_(parts[0] + parts[1])          # rank-order sum
            ctx.call("mjrl_cg_update", ctypes.byref(ref_cg), c_i64(d), c_double(1.0 / 7.0), c_double(1e-4),
                     c_double(1e-10), stream_ptr())
            torch.cuda.synchronize()
            for r in range(world):
                epoch, slot = groups[r].next_slot()
                t, cg = rank_states[r]
                with torch.cuda.stream(streams[r]):
                    # the producer (mjrl_policy_fvp's partial reduce in the engine) writes into the slot
                    src = parts[r].contiguous()
                    _memcpy_d2d(slot.value, src.data_ptr(), 8 * d)
                    groups[r].cg_update(cg, d, 1.0 / 7.0, 1e-4, 1e-10, epoch)
            torch.cuda.synchronize()
            for r in range(world):
                t, _ = rank_states[r]
                for k in ("x", "r", "p", "z", "p32", "scal"):
                    np.testing.assert_array_equal(t[k].cpu().numpy(), ref_t[k].cpu().numpy(), err_msg="%s it %d" % (k, it))
                assert int(t["flags"][1].item()) == it + 1
        for g in groups:
            g.check()
    finally:
        for g in groups:
            g.close()


def _memcpy_d2d(dst, src, nbytes):
    rt = ctypes.CDLL("libcudart.so.12")
    rc = rt.cudaMemcpyAsync(ctypes.c_void_p(dst), ctypes.c_void_p(src), ctypes.c_size_t(nbytes), ctypes.c_int(3),
                            ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, rc


def test_missing_peer_times_out_instead_of_hanging():
    from mjrl_b200 import _lib
    ctx, groups, streams = make_groups(2)
    dev = torch.device("cuda", ctx.device)
    try:
        groups[0].set_timeout(0.2)
        t = torch.ones(8, dtype=torch.float64, device=dev)
        groups[0].allreduce(t)            # rank 1 never arrives
        torch.cuda.synchronize()          # returns after ~0.2 s
        with pytest.raises(_lib.MjrlError):
            groups[0].check()
        groups[1].check()
    finally:
        for g in groups:
            g.close()
