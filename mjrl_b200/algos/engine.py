"""On-device update engine shared by BatchREINFORCE / NPG / TRPO / DAPG.

It strings the C-ABI calls of one update together on the current CUDA stream without any
host synchronisation until the final scalars are read:

    whiten -> gradient (+surr_before, caches old LL/mean) -> [all-reduce]
           -> K x ( FVP -> [all-reduce] -> CG update ) -> step -> eval (+surr_after, KL)

Multi-GPU: every rank owns whole paths; `dist` (torch.distributed, NCCL) all-reduces only the
small fp64 vectors (gradient+surrogate, each FVP, two scalars) -- SURVEY.md section 8(e).
"""
import ctypes
import os

import numpy as np
import torch

from .. import _lib
from .._lib import CGState, c_double, c_float, c_i64, c_int, c_void_p, ptr, stream_ptr


# Below this many local timesteps a CG solve is launch-latency sized and the extra fill pass of the
# Jacobian-cached product does not pay (C2 swimmer, 10k timesteps: 1.26 ms with, 1.28 without).
JAC_MIN_N = int(os.environ.get("MJRL_B200_FVP_JAC_MIN_N", "32768"))
# Same threshold for the tensor-memory product (csrc/fvp_tc.cu): it needs a relayout pass per solve.
TC_MIN_N = int(os.environ.get("MJRL_B200_FVP_TC_MIN_N", "32768"))
# surrogate / KL evaluation on the tcgen05 pass (csrc/pol_tc.cu) for the shapes and batch sizes the product takes
EVAL_TC = os.environ.get("MJRL_B200_EVAL_TC", "1") != "0"


class Dist:
    """Sum all-reduce of small device tensors across the ranks of a torch.distributed group; identity
    when the process group is absent (single GPU).

    Data path (world > 1, CUDA): the peer-memory exchange kernels of csrc/peer.cu (`enable_peer`),
    i.e. one single-CTA kernel per all-reduce over NVLink-mapped blocks, and the CG update fused with
    its all-reduce; NCCL carries the tensors the peer block does not take (not fp64, or longer than the
    block) and everything when MJRL_B200_PEER=0 or the box cannot map peer memory."""

    peer = None
    peer_note = "not tried"

    def __init__(self, group=None):
        import torch.distributed as td
        self.td = td if (td.is_available() and td.is_initialized()) else None
        self.group = group
        self.world = self.td.get_world_size(group) if self.td else 1
        self.rank = self.td.get_rank(group) if self.td else 0
        self.calls = 0
        self.peer = None
        self.peer_note = "not tried"

    def enable_peer(self, ctx, max_count):
        """Set up the peer-memory path once (collective: every rank of the group calls it)."""
        from .. import peer as _peer
        if self.peer is not None or self.td is None or self.world == 1 or self.peer_note != "not tried":
            return self.peer
        mode = _peer.mode()
        if mode in ("0", "off", "nccl") or self.td.get_backend(self.group) != "nccl":
            self.peer_note = "disabled"
            return None
        self.peer, why = _peer.connect_process_group(ctx, self.td, self.group, max_count)
        self.peer_note = "active" if self.peer is not None else "unavailable: " + why
        if self.peer is None and mode in ("1", "on", "required"):
            raise _lib.MjrlError("MJRL_B200_PEER=1 but the peer-memory path is " + self.peer_note)
        if self.peer is None and self.rank == 0:
            import warnings
            warnings.warn("mjrl_b200: peer-memory exchange " + self.peer_note + "; using NCCL all-reduces")
        return self.peer

    def allreduce(self, t):
        if self.td is not None and self.world > 1:
            if self.peer is not None and self.peer.takes(t):
                self.peer.allreduce(t)
            else:
                self.td.all_reduce(t, op=self.td.ReduceOp.SUM, group=self.group)
            self.calls += 1
        return t

    def allgather_rows(self, t):
        """Rows of a [N_r, C] device tensor of every rank, concatenated in rank order on every rank
        (ranks own different row counts).  Used by the replicas-only serial chains (PPO minibatch epochs),
        never by the sharded passes.  Returns (tensor [sum N_r, C], [N_r per rank])."""
        if self.td is None or self.world == 1:
            return t, [int(t.shape[0])]
        n_local = torch.tensor([t.shape[0]], dtype=torch.int64, device=t.device)
        sizes = [torch.zeros_like(n_local) for _ in range(self.world)]
        self.td.all_gather(sizes, n_local, group=self.group)
        sizes = [int(v.item()) for v in sizes]
        padded = torch.zeros((max(sizes),) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        padded[:t.shape[0]] = t
        parts = [torch.empty_like(padded) for _ in range(self.world)]
        self.td.all_gather(parts, padded, group=self.group)
        return torch.cat([p[:k] for p, k in zip(parts, sizes)]), sizes

    def broadcast0(self, t):
        """Rank 0's tensor on every rank (in place)."""
        if self.td is not None and self.world > 1:
            src = self.td.get_global_rank(self.group, 0) if self.group is not None else 0
            self.td.broadcast(t, src=src, group=self.group)
        return t

    def check(self):
        """After a host synchronisation: raise if a peer exchange timed out."""
        if self.peer is not None:
            self.peer.check()

    def global_count(self, n_local, device=None):
        """Sum of per-rank sample counts as a host float.  Shards are whole paths, so ranks
        generally own DIFFERENT numbers of timesteps: every mean over the batch divides by this,
        never by n_local * world.  One tiny all-reduce + host read; callers cache it per batch."""
        if self.td is None or self.world == 1:
            return float(n_local)
        t = torch.tensor([float(n_local)], dtype=torch.float64, device=device)
        self.allreduce(t)
        return float(t.item())


class UpdateEngine:
    def __init__(self, policy, dist=None):
        self.policy = policy
        self.dv = policy.device_view()
        self.ctx = self.dv.ctx
        self.device = self.dv.device
        self.dist = dist or Dist()
        d = self.dv.d
        if self.dist.world > 1:
            self.dist.enable_peer(self.ctx, d + 1)
        dev = self.device
        f64 = lambda k: torch.zeros(k, dtype=torch.float64, device=dev)
        # [gradient sum | surrogate sum] packed for one all-reduce, followed by the update's scalar results
        # (surrogate / KL sums, step scalars): one contiguous device->host read at the end of an update
        self.gbuf = f64(d + 1 + 6)
        self.grad = self.gbuf[:d + 1]
        self.cg_x, self.cg_r, self.cg_p, self.cg_z = f64(d), f64(d), f64(d), f64(d)
        self.cg_scal = f64(8)
        self.cg_p32 = torch.zeros(d, dtype=torch.float32, device=dev)
        self.cg_flags = torch.zeros(2, dtype=torch.int32, device=dev)
        self.cg = CGState(*[c_void_p(t.data_ptr()) for t in
                            (self.cg_x, self.cg_r, self.cg_p, self.cg_z, self.cg_scal, self.cg_p32,
                             self.cg_flags)])
        self.sums2 = self.gbuf[d + 1:d + 3]
        self.scal4 = self.gbuf[d + 3:d + 7]
        self.mom = f64(4)
        self.new_params = torch.zeros(d, dtype=torch.float32, device=dev)
        self.launches = 0  # kernels launched by this engine (bench's gpu_launches)
        self._cache = None  # forward activations shared by the FVPs of one CG solve
        self._cache_key = None
        self._jac_key = None  # (params, obs) the Jacobian rows G1 of the cache were filled for
        self._jac_ok = None
        self._tc_key = None   # same for the tile-ordered copy the tensor-memory product reads
        self._tc_ok = None
        self._ximg_key = None  # obs the [x~ | 1] image of the wide tcgen05 passes was built from
        self._held = None     # tensors of the current cache keys (see _hold)

    # ------------------------------------------------------------------ building blocks
    def whiten(self, adv64, eps=1e-6, scale=1.0, N_global=None):
        """npg_cg.py:91 on device: fp32 (adv-mean)/(std+eps).  Multi-GPU: global moments."""
        N = adv64.numel()
        w = torch.empty(N, dtype=torch.float32, device=self.device)
        if self.dist.world == 1:
            self.ctx.call("mjrl_moments", ptr(adv64), c_i64(N), ptr(self.mom), stream_ptr())
            self.launches += 4
        else:
            # two-pass global moments: all-reduce {N, sum}, then the centred sums of squares
            self.ctx.call("mjrl_moments_stage", ptr(adv64), c_i64(N), ptr(self.mom), c_int(0), stream_ptr())
            self.dist.allreduce(self.mom[0:2])
            self.ctx.call("mjrl_moments_stage", ptr(adv64), c_i64(N), ptr(self.mom), c_int(1), stream_ptr())
            self.dist.allreduce(self.mom[2:3])
            self.launches += 5
        self.ctx.call("mjrl_whiten_f32", ptr(adv64), c_i64(N), ptr(self.mom), c_double(eps),
                      c_double(scale), ptr(w), stream_ptr())
        self.launches += 1
        return w

    def old_cache(self, params_old, obs, act):
        """Old-policy LL [N] and mean [N,m] (what old_dist_info returns, gaussian_mlp.py:116-118)."""
        mean, ll = self.dv.forward(params_old, obs, act)
        self.launches += 1
        return ll, mean

    def gradient(self, params, obs, act, w, old_ll=None, cache=None, accumulate=False):
        """flat_vpg sums (batch_reinforce.py:51-55) into self.grad[:d], surrogate sum into
        self.grad[d].  `cache=(ll, mean)` tensors receive LL/mean of `params`."""
        d = self.dv.d
        N = obs.shape[0]
        tgt = self.grad if not accumulate else torch.empty_like(self.grad)
        ll_out, mean_out = cache if cache is not None else (None, None)
        # the pass computes h1, h2 of (params, obs) anyway: leave them for the FVPs of the CG solve
        # (not for an accumulated second term -- DAPG's demonstration batch, dapg.py:93-98: the CG solve runs on the
        # sampled batch, whose planes the first call left in the cache)
        fcache = self._fvp_cache(N) if not accumulate else None
        direct_tc = fcache is not None and self._tc_kind() > 0 and N >= TC_MIN_N
        # hidden 64 x 64 at a batch the tensor-memory product takes: h1, h2 go straight into its tile order
        self.ctx.call("mjrl_policy_grad_cached_tc" if direct_tc else "mjrl_policy_grad_cached",
                      self.dv.desc_ref(), ptr(params), ptr(obs), ptr(act), ptr(w),
                      ptr(old_ll), c_i64(N), ptr(tgt), c_void_p(tgt.data_ptr() + 8 * d),
                      ptr(ll_out), ptr(mean_out), ptr(fcache), stream_ptr())
        # the pass refilled h1|h2 for these tensors; Jacobian rows of any earlier policy / batch are stale
        if not accumulate:
            key = self._hold(params, obs) if fcache is not None else None
            self._cache_key = None if direct_tc else key     # fragment-ordered h1|h2 (modes 2, 3)
            self._tc_key = key if direct_tc else None        # tile-ordered copy (mode 4)
            self._ximg_key = (self._key(obs, obs), self.dv._tr_key) if direct_tc and self._tc_kind() == 2 else None   # [x~ | 1] image (wide shapes)
            self._jac_key = None
        self.launches += 3
        if accumulate:
            self.grad[:d] += tgt[:d]   # demo term: only the gradient, not the surrogate (dapg.py:93,98)
        return self.grad

    def fvp(self, params, obs, vec32, out64, cache_mode=0):
        """Fisher-vector product sum over obs.  cache_mode 1 / 2: fill / reuse the forward
        activations of `params` on `obs` (same tensors in both calls -- one CG solve)."""
        cache = self._fvp_cache(obs.shape[0]) if cache_mode else None
        if cache is None:
            cache_mode = 0
        self.ctx.call("mjrl_policy_fvp_cached", self.dv.desc_ref(), ptr(params), ptr(obs),
                      c_i64(obs.shape[0]), ptr(vec32), ptr(out64), ptr(self.cg_flags), ptr(cache),
                      c_int(cache_mode), stream_ptr())
        self.launches += 2

    def _jac_prepare(self, params, obs):
        """With the h1|h2 cache valid for (params, obs): make the Jacobian rows valid too (one fill pass
        per CG solve).  False when this policy shape has no Jacobian-cached product."""
        if self._jac_ok is None:
            self._jac_ok = bool(_lib.load().mjrl_policy_fvp_jac_ok(self.ctx.h, self.dv.desc_ref()))
        if not self._jac_ok or self._cache is None:
            return False
        key = self._key(params, obs)
        if self._jac_key != key:
            self.ctx.call("mjrl_policy_fvp_jac_fill", self.dv.desc_ref(), ptr(params), c_i64(obs.shape[0]),
                          ptr(self._cache), stream_ptr())
            self._jac_key = key
        return True

    def _tc_kind(self):
        """0: no tcgen05 product for this policy shape; 1: the fused kernel (csrc/fvp_tc.cu; its planes come from the
        gradient pass or from a relayout of the fragment-ordered cache); 2: the three-pass form for wide observation /
        action vectors (csrc/fvp_tc_wide.cu; planes and x~ image from the gradient pass only)."""
        if self._tc_ok is None:
            self._tc_ok = int(_lib.load().mjrl_policy_fvp_tc_ok(self.ctx.h, self.dv.desc_ref()))
        return self._tc_ok

    def _tc_prepare(self, params, obs):
        """With the h1|h2 cache valid for (params, obs): make its tile-ordered copy for the tensor-memory product
        (csrc/fvp_tc.cu) valid too -- one relayout pass per CG solve.  False when the shape has no such product."""
        if self._tc_kind() != 1 or self._cache is None:
            return False
        key = self._key(params, obs)
        if self._tc_key != key:
            self.ctx.call("mjrl_policy_fvp_tc_fill", self.dv.desc_ref(), c_i64(obs.shape[0]), ptr(self._cache),
                          stream_ptr())
            self._tc_key = key
        return True

    def _cached_mode(self, params, obs):
        """Product variant for a solve whose h1|h2 cache is valid: 4 = tensor-memory product (tcgen05), 3 =
        Jacobian-cached rows kernel, 2 = activation-cached rows kernel."""
        N = obs.shape[0]
        if N >= TC_MIN_N and self._tc_prepare(params, obs):
            return 4
        if N >= JAC_MIN_N and self._jac_prepare(params, obs):
            return 3
        return 2

    @staticmethod
    def _key(params, obs):
        """Identity + in-place-modification counters of the tensors a cache was filled from."""
        return (params.data_ptr(), params._version, obs.data_ptr(), obs._version, obs.shape[0])

    def _hold(self, params, obs):
        """Key of a cache fill.  The engine keeps the two tensors alive while the key is: `dv.params()` builds a
        fresh device tensor per update and the caching allocator would hand a freed one's address (and version
        0) to the next, making a stale cache look current.  Every update's gradient pass refills h1|h2 and drops
        the Jacobian key, so rows written through raw pointers into a reused obs store are covered as well."""
        self._held = (params, obs)
        return self._key(params, obs)

    def _fvp_cache(self, N):
        """Device buffer for the cached-activation FVP (None when the kernel variant that uses it
        does not take this policy shape, or MJRL_B200_FVP_CACHE=0)."""
        if os.environ.get("MJRL_B200_FVP_CACHE", "1") == "0":
            return None
        lib = _lib.load()
        cnt = int(lib.mjrl_policy_fvp_cache_count(self.ctx.h, self.dv.desc_ref(), c_i64(N)))
        if cnt <= 0:
            return None
        if self._cache is None or self._cache.numel() < cnt:
            self._cache = None   # release before growing
            self._cache = torch.empty(cnt, dtype=torch.float32, device=self.device)
        return self._cache

    def cg_solve(self, params, obs, g_scale, N_global, iters, damping, tol=1e-10, trace=None,
                 subsample=None):
        """cg_solve.py:3-22 entirely on device; b = self.grad[:d]*g_scale.  Returns self.cg_x."""
        d = self.dv.d
        peer = self.dist.peer if self.dist.world > 1 else None
        if subsample is None and trace is None and (self.dist.world == 1 or peer is not None) \
                and os.environ.get("MJRL_B200_CG_LOOP", "c") == "c":
            return self._cg_solve_native(params, obs, g_scale, N_global, iters, damping, tol, peer)
        self.ctx.call("mjrl_cg_init", ctypes.byref(self.cg), c_i64(d), ptr(self.grad), c_double(g_scale),
                      stream_ptr())
        self.launches += 1
        for k in range(iters):
            o = obs if subsample is None else subsample()
            cnt = N_global if subsample is None else self.dist.global_count(o.shape[0], self.device)
            # params and obs are fixed across the iterations of one solve (unless the operator
            # subsamples): the first product fills the activation cache, the others reuse it
            if subsample is not None:
                mode = 0
            elif self._tc_key is not None and self._tc_key == self._key(params, o):
                mode = 4      # the gradient pass left h1|h2 in the tile order of the tensor-memory product
            elif self._cache_key is not None and self._cache_key == self._key(params, o):
                # h1|h2 filled by the gradient pass (or the first product) at these tensors; small action
                # dimensions also cache the output-Jacobian rows and run the cheaper product (fvp_jac.cu)
                mode = self._cached_mode(params, o)
            else:
                mode = 1
            peer = self.dist.peer if self.dist.world > 1 else None
            if peer is not None:
                # the product's partial reduce writes straight into this rank's exchange slot; the
                # all-reduce over NVLink peer memory happens inside the CG update kernel
                epoch, slot = peer.next_slot()
                self.fvp(params, o, self.cg_p32, slot, cache_mode=mode)
            else:
                self.fvp(params, o, self.cg_p32, self.cg_z, cache_mode=mode)
            if mode == 1 and self._cache is not None:
                self._cache_key = self._hold(params, o)
                self._jac_key = None
                self._tc_key = None
            if peer is not None:
                peer.cg_update(self.cg, d, 1.0 / cnt, damping, tol, epoch)
                self.dist.calls += 1
            else:
                self.dist.allreduce(self.cg_z)
                self.ctx.call("mjrl_cg_update", ctypes.byref(self.cg), c_i64(d), c_double(1.0 / cnt),
                              c_double(damping), c_double(tol), stream_ptr())
            self.launches += 1
            if trace is not None:
                trace.append(self.cg_x.cpu().numpy().copy())
        return self.cg_x

    def _cg_solve_native(self, params, obs, g_scale, N_global, iters, damping, tol, peer):
        """The same solve through ONE C-ABI call (`mjrl_cg_solve`): the host loop over product / update
        launches runs in C, which matters when the batch is small enough for the update to be bound by the
        rate at which Python can enqueue kernels (small configs, strong scaling)."""
        N = obs.shape[0]
        key = self._key(params, obs)
        cache = self._fvp_cache(N)
        fill_jac = 0
        if cache is None:
            first = rest = 0
        elif self._tc_key is not None and self._tc_key == key:
            first = rest = 4      # the gradient pass left h1|h2 in the tile order of the tensor-memory product
        elif self._cache_key is not None and self._cache_key == key:
            first = rest = self._cached_mode(params, obs)
        else:
            first, rest = 1, 2
            key = self._hold(params, obs)
            self._jac_key = None
            self._tc_key = None
            if iters > 1:
                if self._jac_ok is None:
                    self._jac_ok = bool(_lib.load().mjrl_policy_fvp_jac_ok(self.ctx.h, self.dv.desc_ref()))
                if self._tc_kind() == 1 and N >= TC_MIN_N:
                    rest, fill_jac = 4, 2        # relayout pass after the first (cache-filling) product
                    self._tc_key = key
                elif self._jac_ok and N >= JAC_MIN_N:
                    rest, fill_jac = 3, 1
                    self._jac_key = key
            self._cache_key = key
        epoch0 = 0
        if peer is not None:
            epoch0 = peer.epoch
            peer.epoch += iters
            peer.exchanges += iters
            self.dist.calls += iters
        self.ctx.call("mjrl_cg_solve", self.dv.desc_ref(), ptr(params), ptr(obs), c_i64(N), ctypes.byref(self.cg),
                      ptr(self.grad), c_double(g_scale), c_int(iters), c_double(1.0 / N_global),
                      c_double(damping), c_double(tol), ptr(cache), c_int(first), c_int(rest), c_int(fill_jac),
                      peer.h if peer is not None else c_void_p(0), c_i64(epoch0), stream_ptr())
        self.launches += 1 + 3 * iters + (1 if fill_jac else 0)
        self.ctx.launches += 3 * iters + (1 if fill_jac else 0)     # ctx.call counted one
        return self.cg_x

    def step(self, curr, g_scale, delta, const_alpha=None, alpha_scale=1.0):
        """npg_cg.py:130-141: alpha, new_params = curr + alpha*alpha_scale*x, log_std clamped."""
        self.ctx.call("mjrl_npg_step", c_i64(self.dv.d), c_int(self.dv.m), ptr(self.grad),
                      c_double(g_scale), ptr(self.cg_x), ptr(curr), c_double(delta),
                      c_double(-1.0 if const_alpha is None else const_alpha), c_double(alpha_scale),
                      c_float(self.policy.min_log_std), ptr(self.new_params), ptr(self.scal4),
                      stream_ptr())
        self.launches += 1
        return self.new_params

    def evaluate(self, params_new, obs, act, w, old_ll, old_mean, old_log_std, N_global, raw=False):
        """(surrogate, KL) of params_new against the cached old policy -> device fp64 [2] means (`raw`: the sums,
        left in `self.sums2` for `tail()`: the caller divides on the host)."""
        N = obs.shape[0]
        kind = self._tc_kind() if N >= TC_MIN_N and EVAL_TC else 0
        self.dv.desc_ref()      # refreshes the transform key the image key below is compared with
        if kind == 2 and not (self._cache is not None and self._ximg_key is not None and self._ximg_key == (self._key(obs, obs), self.dv._tr_key)):
            kind = 0      # wide shapes need the [x~ | 1] image the gradient pass of THIS batch left in the cache
        if kind:
            self.ctx.call("mjrl_policy_eval_tc", self.dv.desc_ref(), ptr(params_new), ptr(obs), ptr(act),
                          ptr(w), ptr(old_ll), ptr(old_mean), ptr(old_log_std), c_i64(N),
                          ptr(self.sums2), ptr(self._cache), stream_ptr())
            self.launches += 1 if kind == 2 else 0
        else:
            self.ctx.call("mjrl_policy_eval", self.dv.desc_ref(), ptr(params_new), ptr(obs), ptr(act),
                          ptr(w), ptr(old_ll), ptr(old_mean), ptr(old_log_std), c_i64(N),
                          ptr(self.sums2), stream_ptr())
        self.launches += 2
        self.dist.allreduce(self.sums2)
        return self.sums2 if raw else self.sums2 / N_global

    def tail(self):
        """Device view [surrogate-before sum, surrogate-after sum, KL sum, alpha, g.x, n_step_size, alpha*scale]:
        what an update reads back, contiguous behind the gradient."""
        d = self.dv.d
        return self.gbuf[d:d + 7]
