"""Lock-step batched rollouts: the B200-first replacement of the reference's sampling loop.

The reference runs `policy.get_action(o)` once per environment step inside every worker process
(base_sampler.py:64-66: a batch-1 torch forward plus `np.random.randn`), pickles the policy out
to the workers and the finished `paths` back (trajectory_sampler.py:71-78), and every consumer
then concatenates the list again.  Here E environment slots advance together:

    per simulator step:  obs of all slots --one H2D--> mjrl_rollout_act (forward + noise + filing the
                         rows into the device-resident timestep store) --one D2H--> actions
                         -> env.step per slot (in-process, or in worker processes through a
                            page-locked shared-memory block, `_envs.WorkerEnvs`)

Episode `ep` always uses seed `pegasus_seed + ep` (base_sampler.py:44-50) for its environment AND for
its exploration noise, so the batch is the same whatever E, the worker count or the slot assignment.
`noise="reference"` draws the noise on the host from `RandomState(seed).randn` -- the reference's own
stream -- which makes the paths comparable value for value with `base_sampler.do_rollout`;
`noise="philox"` (default) generates it inside the kernel.

When the rollout ends the observations and actions are already on the GPU in the flat path-major
layout the update kernels take; the returned `paths` (host views, reference format) carry that
`PathBatch`, so `train_step` uploads only the rewards.
"""
import numpy as np
import torch

from .._lib import c_i64, c_int, ptr, stream_ptr
from ..utils import tensor_utils
from ..utils.get_environment import factory_for
from ._envs import LocalEnvs, StepBlock, WorkerEnvs


class RolloutPaths(list):
    """`paths` in the reference's format plus the device-resident batch they were filed into."""
    batch = None


class DeviceActor:
    """obs [E,n] -> actions, through `mjrl_rollout_act`; owns the device mirrors of the step block and
    the padded [episode][T] stores."""

    def __init__(self, policy, block, pinned_host, rows):
        self.dv = policy.device_view()
        self.ctx, self.device = self.dv.ctx, self.dv.device
        self.blk = block
        self.host = pinned_host                       # uint8 torch tensor over the block's bytes
        self.up = block.span(StepBlock.UPLOAD)
        self.down = block.span(StepBlock.DOWNLOAD)
        self.dev = torch.empty(block.nbytes, dtype=torch.uint8, device=self.device)
        self.obs_store = torch.empty((rows, block.n), dtype=torch.float32, device=self.device)
        self.act_store = torch.empty((rows, block.m), dtype=torch.float32, device=self.device)
        self.params = self.dv.params("new")

    def _p(self, name):
        return ptr(int(self.dev.data_ptr() + self.blk.off[name][0]))

    def __call__(self, sample, use_noise):
        lo, nb = self.up
        self.dev[lo:lo + nb].copy_(self.host[lo:lo + nb], non_blocking=True)
        null = ptr(None)
        self.ctx.call("mjrl_rollout_act", self.dv.desc_ref(), ptr(self.params), self._p("obs32"),
                      c_int(self.blk.E), null if use_noise else self._p("seeds"),
                      null if use_noise else self._p("tsteps"), self._p("noise") if use_noise else null,
                      c_int(1 if sample else 0), self._p("dst_rows"), ptr(self.obs_store),
                      ptr(self.act_store), self._p("act"), self._p("mean"), stream_ptr())
        lo, nb = self.down
        self.host[lo:lo + nb].copy_(self.dev[lo:lo + nb], non_blocking=True)
        torch.cuda.current_stream().synchronize()

    def finish(self, src_rows, N):
        """Flat [N,n] / [N,m] device arrays: the stores themselves when every path ran to T."""
        if src_rows is None:
            return self.obs_store[:N], self.act_store[:N]
        idx = torch.from_numpy(src_rows).to(self.device)
        obs = torch.empty((N, self.blk.n), dtype=torch.float32, device=self.device)
        act = torch.empty((N, self.blk.m), dtype=torch.float32, device=self.device)
        for src, dst, w in ((self.obs_store, obs, self.blk.n), (self.act_store, act, self.blk.m)):
            self.ctx.call("mjrl_rows_gather_f32", ptr(src), ptr(idx), c_i64(N), c_int(w), ptr(dst), stream_ptr())
        return obs, act


class LockstepSampler:
    """E environment slots stepped together.

    env / env_name : a zero-argument factory (picklable when num_workers > 0), a list of ready
                     environments (in-process only), or a name known to utils.get_environment.
    num_slots      : E, environments alive at once (default 64, never more than the episodes asked for).
    num_workers    : 0 steps the environments in this process; W > 0 keeps them in W worker processes.
    actor_factory  : seam for the host-logic tests; the default (and only shipped) actor is the GPU one.
    """

    def __init__(self, env=None, env_name=None, num_slots=64, num_workers=0, collect_env_infos=True,
                 actor_factory=None, worker_timeout=300.0):
        if env is None and env_name is None:
            raise ValueError("one of env / env_name must be given")
        self.envs_given = list(env) if isinstance(env, (list, tuple)) else None
        self.factory = None if self.envs_given is not None else (env if callable(env) else factory_for(env_name))
        if self.envs_given is not None and num_workers:
            raise ValueError("ready-made environments cannot be moved into worker processes; pass a factory")
        self.num_slots, self.num_workers = int(num_slots), int(num_workers)
        self.collect = collect_env_infos
        self.actor_factory = actor_factory
        self.worker_timeout = worker_timeout
        self._backend = None
        self._shape = None
        self._pinned = None

    # ---- set-up (kept across calls: worker start-up is the expensive part) -------------------------
    def _ensure(self, E, n, m):
        if self._backend is not None and self._shape == (E, n, m):
            return
        self.close()
        _, nbytes = StepBlock.layout(E, n, m)
        if self.num_workers > 0:
            be = WorkerEnvs(self.factory, E, n, m, self.num_workers, self.collect, self.worker_timeout)
            blk = be.blk
            host = torch.frombuffer(be.shm.buf, dtype=torch.uint8, count=nbytes)
            if torch.cuda.is_available():       # page-lock the shared block: it is the DMA end of every step
                rc = torch.cuda.cudart().cudaHostRegister(host.data_ptr(), nbytes, 0)
                self._registered = host.data_ptr() if int(rc) == 0 else None
        else:
            host = torch.empty(nbytes, dtype=torch.uint8, pin_memory=torch.cuda.is_available())
            blk = StepBlock(memoryview(host.numpy()), E, n, m)
            envs = self.envs_given[:E] if self.envs_given is not None else [self.factory() for _ in range(E)]
            be = LocalEnvs(envs, blk, self.collect)
        self._backend, self._blk, self._pinned, self._shape = be, blk, host, (E, n, m)

    def close(self):
        if self._backend is not None:
            reg = getattr(self, "_registered", None)
            if reg:
                torch.cuda.cudart().cudaHostUnregister(reg)
                self._registered = None
            self._pinned = None
            self._blk = None
            self._backend.close()
            self._backend = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- the rollout ----------------------------------------------------------------------------
    def sample(self, N, policy, T=1e6, pegasus_seed=None, mode="sample", noise="philox", episode_offset=0):
        """N trajectories (base_sampler.py:10-88 / evaluation_sampler.py:11-85 for mode='evaluation').
        `episode_offset` shifts the episode numbers (a rank's share of a sharded rollout)."""
        if mode not in ("sample", "evaluation"):
            raise ValueError("mode has to be 'sample' or 'evaluation'")
        if noise not in ("philox", "reference"):
            raise ValueError("noise has to be 'philox' or 'reference'")
        N = int(N)
        n, m = int(policy.n), int(policy.m)
        if self.envs_given is not None:
            E = min(N, len(self.envs_given))
        else:
            E = min(N, self.num_slots)
        self._ensure(E, n, m)
        be, blk = self._backend, self._blk
        T = int(min(T, be.horizon))
        sample = mode == "sample"
        use_noise = sample and noise == "reference"
        if pegasus_seed is None:        # base_sampler.py:51-52: fresh entropy per episode
            ep_seeds = np.random.SeedSequence().generate_state(N, dtype=np.uint64) >> np.uint64(1)
        else:
            ep_seeds = (int(pegasus_seed) + int(episode_offset) + np.arange(N)).astype(np.uint64)

        rows = N * T
        actor = (self.actor_factory or DeviceActor)(policy, blk, self._pinned, rows)
        obs_h = np.empty((rows, n), dtype=np.float64)
        act_h = np.empty((rows, m), dtype=np.float32)
        mean_h = np.empty((rows, m), dtype=np.float32)
        rew_h = np.empty(rows, dtype=np.float64)
        lens = np.zeros(N, dtype=np.int64)
        term = np.zeros(N, dtype=bool)
        infos = [[] for _ in range(N)] if self.collect else None

        slot_ep = np.arange(E, dtype=np.int64)          # episode each slot is running; -1 = retired
        slot_t = np.zeros(E, dtype=np.int64)
        slot_noise = [None] * E

        def begin(slots):
            be.reset(slots, [ep_seeds[slot_ep[s]] if pegasus_seed is not None or use_noise else None
                             for s in slots])
            for s in slots:
                slot_t[s] = 0
                blk.seeds[s] = ep_seeds[slot_ep[s]]
                if use_noise:       # the reference's stream: np.random.seed(seed); randn(m) per step
                    slot_noise[s] = np.random.RandomState(int(ep_seeds[slot_ep[s]]) % (2 ** 32)).randn(T, m)

        begin(list(range(E)))
        blk.noise[:] = 0
        while True:
            live = np.nonzero(slot_ep >= 0)[0]
            if live.size == 0:
                break
            dst = slot_ep[live] * T + slot_t[live]
            blk.dst_rows[:] = -1
            blk.dst_rows[live] = dst
            blk.tsteps[:] = slot_t
            blk.obs32[live] = blk.obs64[live]
            obs_h[dst] = blk.obs64[live]
            if use_noise:
                for s in live:
                    blk.noise[s] = slot_noise[s][slot_t[s]]
            actor(sample, use_noise)
            act_h[dst] = blk.act[live]
            mean_h[dst] = blk.mean[live]
            step_infos = be.step(live)
            rew_h[dst] = blk.rew[live]
            slot_t[live] += 1
            if self.collect:
                for s, info in zip(live, step_infos):
                    infos[slot_ep[s]].append(info)
            ended = live[(blk.done[live] != 0) | (slot_t[live] >= T)]
            restart = []
            for s in ended:
                ep = slot_ep[s]
                lens[ep] = slot_t[s]
                term[ep] = bool(blk.done[s])
                nxt = ep + E
                slot_ep[s] = nxt if nxt < N else -1
                if nxt < N:
                    restart.append(int(s))
            if restart:
                begin(restart)

        return self._assemble(policy, actor, N, T, lens, term, obs_h, act_h, mean_h, rew_h, infos)

    def _assemble(self, policy, actor, N, T, lens, term, obs_h, act_h, mean_h, rew_h, infos):
        off = np.zeros(N + 1, dtype=np.int64)
        np.cumsum(lens, out=off[1:])
        total = int(off[-1])
        ragged = total != N * T
        src_rows = None
        if ragged:      # flat row r of path p sits at p*T + (r - off[p]) in the padded stores
            src_rows = np.repeat(np.arange(N, dtype=np.int64) * T - off[:-1], lens) + np.arange(total, dtype=np.int64)
        log_std = np.asarray(policy.log_std_val, dtype=np.float64).reshape(1, -1)
        paths = RolloutPaths()
        for ep in range(N):
            a, b = ep * T, ep * T + int(lens[ep])
            paths.append(dict(
                observations=obs_h[a:b], actions=act_h[a:b], rewards=rew_h[a:b],
                agent_infos=dict(mean=mean_h[a:b], log_std=np.broadcast_to(log_std, (b - a, log_std.shape[1])),
                                 evaluation=mean_h[a:b]),
                env_infos=tensor_utils.stack_tensor_dict_list(infos[ep]) if infos is not None else {},
                terminated=bool(term[ep])))
        finish = getattr(actor, "finish", None)
        if finish is not None:
            from ..batch import PathBatch
            obs_d, act_d = finish(src_rows, total)
            rew_flat = rew_h[src_rows] if ragged else rew_h[:total]
            paths.batch = PathBatch.from_device(obs_d, act_d, rew_flat, off, term, paths=paths)
        return paths
