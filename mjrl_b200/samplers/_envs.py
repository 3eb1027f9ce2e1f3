"""Environment back-ends of the lock-step sampler: where `env.reset()` / `env.step(a)` run.

`LocalEnvs` steps E environments in the calling process.  `WorkerEnvs` keeps them in W host worker
processes (the reference's `mp.Pool` of trajectory_sampler.py:71-74) that exchange observations,
actions, rewards and done flags with the parent through ONE shared-memory block, which the parent
page-locks so the same bytes are the DMA source/target of the per-step copies: nothing is pickled
per step except the env_info dicts (and those only on request).

This module must stay importable without torch: the worker processes import it.
"""
import multiprocessing as mp
from multiprocessing import shared_memory

import numpy as np

from ..utils.get_environment import seed_environment


def _align(x, a=256):
    return (x + a - 1) // a * a


class StepBlock:
    """Typed views into one byte block: the per-step exchange area, [E] slots.

    upload part (parent -> device):  obs32 [E,n] f32 | dst_rows [E] i64 | seeds [E] u64 |
                                     tsteps [E] i32 | noise [E,m] f32
    download part (device -> parent): act [E,m] f32 | mean [E,m] f32
    host-only part:                   obs64 [E,n] f64 | rew [E] f64 | done [E] u8
    """
    FIELDS = (("obs32", "f4", "n"), ("dst_rows", "i8", 1), ("seeds", "u8", 1), ("tsteps", "i4", 1),
              ("noise", "f4", "m"), ("act", "f4", "m"), ("mean", "f4", "m"),
              ("obs64", "f8", "n"), ("rew", "f8", 1), ("done", "u1", 1))
    UPLOAD = ("obs32", "dst_rows", "seeds", "tsteps", "noise")
    DOWNLOAD = ("act", "mean")

    @classmethod
    def layout(cls, E, n, m):
        off, pos = {}, 0
        for name, dt, w in cls.FIELDS:
            width = {"n": n, "m": m}.get(w, w)
            nbytes = E * width * np.dtype(dt).itemsize
            off[name] = (pos, nbytes, dt, width)
            pos = _align(pos + nbytes)
        return off, pos

    def __init__(self, buf, E, n, m):
        self.E, self.n, self.m = E, n, m
        self.off, self.nbytes = self.layout(E, n, m)
        self.buf = buf
        for name, (pos, nbytes, dt, width) in self.off.items():
            a = np.frombuffer(buf, dtype=dt, count=E * width, offset=pos)
            setattr(self, name, a.reshape(E, width) if width > 1 or name in ("obs32", "obs64", "noise", "act", "mean")
                    else a)

    def span(self, names):
        """(byte offset, byte length) covering the named consecutive fields."""
        lo = self.off[names[0]][0]
        last = self.off[names[-1]]
        return lo, _align(last[0] + last[1]) - lo


class LocalEnvs:
    """E environments stepped in the calling process."""

    def __init__(self, envs, block, collect_env_infos=True):
        self.envs = list(envs)
        self.blk = block
        self.collect = collect_env_infos

    @property
    def horizon(self):
        return self.envs[0].horizon

    def reset(self, slots, seeds):
        for s, seed in zip(slots, seeds):
            env = self.envs[s]
            if seed is not None:
                seed_environment(env, int(seed))
                np.random.seed(int(seed) % (2 ** 32))   # base_sampler.py:50 (for resets drawing from the global stream)
            self.blk.obs64[s] = np.asarray(env.reset(), dtype=np.float64).reshape(-1)

    def step(self, slots):
        infos = []
        blk = self.blk
        for s in slots:
            o, r, d, info = self.envs[s].step(np.asarray(blk.act[s], dtype=np.float64))
            blk.obs64[s] = np.asarray(o, dtype=np.float64).reshape(-1)
            blk.rew[s] = r
            blk.done[s] = 1 if d == True else 0    # noqa: E712
            if self.collect:
                infos.append(info)
        return infos if self.collect else None

    def close(self):
        self.envs = []


def _worker_main(conn, factory, shm_name, E, n, m, my_slots, collect):
    shm = shared_memory.SharedMemory(name=shm_name)
    try:
        blk = StepBlock(shm.buf, E, n, m)
        envs = {s: factory() for s in my_slots}
        local = LocalEnvs([None] * E, blk, collect)
        for s, e in envs.items():
            local.envs[s] = e
        conn.send(("ready", int(next(iter(envs.values())).horizon) if envs else 0))
        while True:
            cmd, a, b = conn.recv()
            if cmd == "reset":
                local.reset(a, b)
                conn.send(("ok", None))
            elif cmd == "step":
                conn.send(("ok", local.step(a)))
            else:
                break
    except BaseException as e:          # report instead of dying silently: the parent is blocked on recv()
        try:
            conn.send(("error", repr(e)))
        except Exception:
            pass
    finally:
        del blk, local
        shm.close()


class WorkerEnvs:
    """E environments spread over W worker processes; slot s lives in worker s % W."""

    def __init__(self, factory, E, n, m, num_workers, collect_env_infos=True, timeout=300.0):
        self.E, self.W = E, max(1, min(int(num_workers), E))
        self.collect = collect_env_infos
        self.timeout = timeout
        _, nbytes = StepBlock.layout(E, n, m)
        self.shm = shared_memory.SharedMemory(create=True, size=nbytes)
        self.blk = StepBlock(self.shm.buf, E, n, m)
        ctx = mp.get_context("spawn")       # the parent holds a CUDA context: never fork it
        self.conns, self.procs = [], []
        for w in range(self.W):
            mine = list(range(w, E, self.W))
            parent, child = ctx.Pipe()
            p = ctx.Process(target=_worker_main, daemon=True,
                            args=(child, factory, self.shm.name, E, n, m, mine, collect_env_infos))
            p.start()
            child.close()
            self.conns.append(parent); self.procs.append(p)
        self._horizon = None
        for c in self.conns:
            tag, val = self._recv(c)
            self._horizon = val if self._horizon is None else min(self._horizon, val)

    @property
    def horizon(self):
        return self._horizon

    def _recv(self, conn):
        if not conn.poll(self.timeout):     # trajectory_sampler.py:80-88: a stuck worker is an error, not a hang
            self.close()
            raise TimeoutError("rollout worker did not answer within %.0f s" % self.timeout)
        tag, val = conn.recv()
        if tag == "error":
            self.close()
            raise RuntimeError("rollout worker failed: " + str(val))
        return tag, val

    def _fanout(self, cmd, slots, extra=None):
        per = [[] for _ in range(self.W)]
        per_x = [[] for _ in range(self.W)]
        for i, s in enumerate(slots):
            per[int(s) % self.W].append(int(s))
            if extra is not None:
                per_x[int(s) % self.W].append(extra[i])
        busy = [w for w in range(self.W) if per[w]]
        for w in busy:
            self.conns[w].send((cmd, per[w], per_x[w] if extra is not None else None))
        return busy, per, [self._recv(self.conns[w])[1] for w in busy]

    def reset(self, slots, seeds):
        self._fanout("reset", slots, [None if s is None else int(s) for s in seeds])

    def step(self, slots):
        busy, per, answers = self._fanout("step", slots)
        if not self.collect:
            return None
        by_slot = {}
        for w, infos in zip(busy, answers):
            for s, info in zip(per[w], infos):
                by_slot[s] = info
        return [by_slot[int(s)] for s in slots]

    def close(self):
        for c in self.conns:
            try:
                c.send(("close", None, None))
            except Exception:
                pass
        for p in self.procs:
            p.join(timeout=5)
            if p.is_alive():
                p.terminate()
        self.conns, self.procs = [], []
        if self.shm is not None:
            blk, self.blk = self.blk, None
            del blk
            try:
                self.shm.close()
                self.shm.unlink()
            except Exception:
                pass
            self.shm = None
