#!/usr/bin/env python
"""Benchmark of the policy-update hot path: NPG update timesteps/s (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A "step" is ONE full NPG update of a synthetic Hopper-shaped batch (obs 11, act 3,
MLP(64,64) policy, 1M timesteps per GPU): returns scan + baseline predict + GAE scan +
advantage whitening + policy gradient + 10 x (Fisher-vector product + CG update) + step +
surrogate/KL evaluation + LinearBaseline fit (Gram + Cholesky).  The policy is reset to the
same parameters before every step so each step does identical work.

Prints ONE JSON line (rank 0).  `value` = device-resident throughput (CUDA events, L2 flushed
between steps, max over ranks); `e2e` = the same update through the public API from pinned
HOST fp64 buffers, H2D and the D2H of the results inside the timed region; `roofline`
describes the dominant kernel (the FVP pass); `cpu_baseline` is the oracle port timed on this
box's host cores on a bounded 100k-timestep sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOAD = "C3-hopper"            # BASELINE.json configs[2] shapes; LinearBaseline (north_star (b))
GAMMA, LAM, DAMPING, CG_ITERS, STEP = 0.995, 0.97, 1e-4, 10, 0.01
CPU_SAMPLE_N = 100_000


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=WORKLOAD)
    ap.add_argument("--timesteps", type=int, default=0, help="override timesteps per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--baseline", default="linear", choices=["linear", "mlp"],
                    help="value baseline inside the timed update: LinearBaseline (north_star (b), default) or "
                         "MLPBaseline (BASELINE.json configs[2]; its fit is a serial chain of Adam steps)")
    ap.add_argument("--phases", action="store_true", help="also print a per-phase breakdown to stderr")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


# ------------------------------------------------------------------------------------------
# CPU baseline / reference arm: the oracle port (the reference is pure Python and cannot travel
# to the GPU box; oracle/mjrl_oracle.py is pinned bit-exact to it by tests/golden)
# ------------------------------------------------------------------------------------------
def cpu_update_setup(shape, N, seed):
    import torch
    from mjrl_b200.utils import synthetic as syn
    from oracle import mjrl_oracle as O
    n, m, hidden, horizon, _ = syn.SHAPES[shape]
    torch.set_num_threads(os.cpu_count())
    flat = O.init_policy_params(n, m, hidden, seed=123)
    b0 = syn.make_flat_batch(N, n, m, horizon, seed, flat, hidden)
    b1 = syn.make_flat_batch(N, n, m, horizon, seed + 1, flat, hidden)
    base = O.LinearBaseline()
    p0 = syn.flat_to_paths(b0)
    O.compute_returns(p0, GAMMA)
    base.fit(p0)                                   # steady state: predict() is non-trivial (F8)
    return O, syn, (n, m, hidden), flat, b1, base


def cpu_update_once(O, syn, dims, flat, b1, base):
    import copy
    n, m, hidden = dims
    pol = O.GaussianPolicy(n, m, hidden, flat)
    paths = syn.flat_to_paths(b1)
    bl = copy.deepcopy(base)
    t = time.perf_counter()
    O.full_npg_iteration(pol, bl, paths, GAMMA, LAM, normalized_step_size=STEP,
                         cg_iters=CG_ITERS, damping=DAMPING)
    return time.perf_counter() - t


def cpu_baseline(shape, reps=5):
    import torch
    ctx = cpu_update_setup(shape, CPU_SAMPLE_N, 900)
    cpu_update_once(*ctx)                          # warm-up (first torch call is slow)
    ts = sorted(cpu_update_once(*ctx) for _ in range(reps))
    med = ts[len(ts) // 2]
    return {"value": CPU_SAMPLE_N / med, "unit": "timesteps/s", "cores": torch.get_num_threads(),
            "kind": "port",
            "sample": "oracle port of the reference update (numpy + CPU torch autograd, python-loop "
                      "scans are single-threaded), %d timesteps of %s, median of %d after 1 warm-up"
                      % (CPU_SAMPLE_N, shape, reps)}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    ctx = cpu_update_setup(args.workload, CPU_SAMPLE_N, 900)
    for _ in range(max(1, args.warmup)):
        cpu_update_once(*ctx)
    ts = [cpu_update_once(*ctx) for _ in range(args.steps)]
    tot = sum(ts)
    val = CPU_SAMPLE_N * args.steps / tot
    cb = {"value": val, "unit": "timesteps/s", "cores": torch.get_num_threads(), "kind": "port",
          "sample": "%d-timestep sample of %s per step" % (CPU_SAMPLE_N, args.workload)}
    print(json.dumps({
        "impl": "reference", "metric": "NPG update timesteps/sec", "value": val,
        "unit": "timesteps/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * tot / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "%s NPG + MLP(64,64) + LinearBaseline, CPU oracle port on a bounded "
                               "sample of %d timesteps/step" % (args.workload, CPU_SAMPLE_N)},
        "cpu_baseline": cb,
        "e2e": {"value": val, "unit": "timesteps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


# ------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "20",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            pass
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx = float(r[2])
            except Exception:
                continue
            for k, nm in enumerate(names):
                if len(r) > 5 + k and r[5 + k].lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        busy = [v for v in sm if mx and v > 0.5 * mx] or sm
        return {"sm_mhz": busy[len(busy) // 2] if busy else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}


def run_b200(args):
    import torch
    import torch.distributed as td
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        td.init_process_group("nccl", device_id=dev)

    from mjrl_b200 import _lib
    from mjrl_b200.utils import synthetic as syn
    from mjrl_b200.utils.gym_env import EnvSpec
    from mjrl_b200.policies.gaussian_mlp import MLP
    from mjrl_b200.policies.gaussian_linear import LinearPolicy
    from mjrl_b200.baselines.linear_baseline import LinearBaseline
    from mjrl_b200.algos.npg_cg import NPG
    from mjrl_b200.batch import PathBatch

    n, m, hidden, horizon, N = syn.SHAPES[args.workload]
    if args.timesteps:
        N = args.timesteps
    spec = EnvSpec(n, m, horizon, 1)
    pol = MLP(spec, hidden_sizes=hidden, seed=123) if hidden else LinearPolicy(spec, seed=123)
    flat0 = pol.get_param_values()
    use_mlp = args.baseline == "mlp"
    if use_mlp:
        from mjrl_b200.baselines.mlp_baseline import MLPBaseline
        torch.manual_seed(7)
        np.random.seed(7 + rank)
        base = MLPBaseline(spec, epochs=1)
    else:
        base = LinearBaseline(spec)
    agent = NPG(None, pol, base, normalized_step_size=STEP,
                FIM_invert_args={"iters": CG_ITERS, "damping": DAMPING})

    # synthetic trajectories (seed depends on the rank: every GPU owns its own whole paths)
    hb = syn.make_flat_batch(N, n, m, horizon, 3000 + rank, flat0, hidden)
    pin = lambda a: torch.from_numpy(a).pin_memory()
    h_obs, h_act, h_rew = pin(hb["observations"]), pin(hb["actions"]), pin(hb["rewards"])
    off, term = hb["path_offsets"], hb["terminated"]
    ctx = _lib.context(local)

    def barrier():
        if world > 1:
            td.barrier()
        torch.cuda.synchronize()

    # steady state (F8): fit the baseline once so predict() is non-trivial, remember its coeffs
    warm = PathBatch(h_obs, h_act, h_rew, off, term, device=local)
    from mjrl_b200.utils import process_samples as ps
    ps.returns_on_device(warm, GAMMA)
    base.fit_batch(warm, dist=agent._engine().dist if world > 1 else None)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2
    if use_mlp:
        st0 = {k: v.clone() for k, v in base._state(dev).items()}
        step0 = base.adam_step

        def reset():
            pol.set_param_values(flat0.copy(), True, True)
            for k, v in base._state(dev).items():
                v.copy_(st0[k])
            base.adam_step = step0
    else:
        coeffs0 = np.asarray(base._coeffs).copy()

        def reset():
            pol.set_param_values(flat0.copy(), True, True)
            base._coeffs = coeffs0.copy()

    resident = PathBatch(h_obs, h_act, h_rew, off, term, device=local)

    def step_resident():
        resident.returns = resident.baseline = resident.adv = None
        return agent.train_step(N, gamma=GAMMA, gae_lambda=LAM, paths=resident)

    def step_e2e():
        b = PathBatch(h_obs, h_act, h_rew, off, term, device=local)   # H2D from pinned host memory
        out = agent.train_step(N, gamma=GAMMA, gae_lambda=LAM, paths=b)
        return b, out

    # ---- value: inputs resident in HBM, CUDA events, L2 flushed between steps ---------------
    for _ in range(max(3, args.warmup)):
        reset(); step_resident()
    barrier()
    clocks = ClockSampler(local)
    clocks.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(args.steps)]
    l0 = ctx.launches
    barrier()
    for k in range(args.steps):
        reset()
        flush.zero_()
        torch.cuda.synchronize()
        if world > 1:
            # reset() and the L2 flush are outside the event brackets on every rank; without a common start a
            # rank that finished them earlier would count its wait for the other ranks' host work as step time
            td.barrier()
        ev[k][0].record()
        step_resident()
        ev[k][1].record()
    barrier()
    launches = ctx.launches - l0
    clk = clocks.stop()
    ms = sum(a.elapsed_time(b) for a, b in ev)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        td.all_reduce(t, op=td.ReduceOp.MAX)
    ms_total = float(t.item())
    ms_step = ms_total / args.steps
    value = N * world / (ms_step * 1e-3)

    # ---- e2e: public API from pinned host buffers, H2D + D2H inside the timed region ----------
    for _ in range(2):
        reset(); step_e2e()
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        reset()
        b, _ = step_e2e()
        torch.cuda.synchronize()
    barrier()
    te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        td.all_reduce(te, op=td.ReduceOp.MAX)
    e2e_s = float(te.item()) / args.steps
    d = int(pol.d)
    d2h = 4 * d + 8 * (2 + 1 + 4) + 8 * 4 + (4 * 18689 if use_mlp else 8 * (n + 4))
    e2e = {"value": N * world / e2e_s, "unit": "timesteps/s", "ms_per_step": 1e3 * e2e_s,
           "h2d_bytes_per_step": int(b.h2d_bytes), "d2h_bytes_per_step": int(d2h)}

    # ---- strong scaling (north_star: ONE batch of the named size sharded over the GPUs) -----------
    # The headline `value` is weak scaling (the named batch per GPU).  Here the same N timesteps are split
    # by whole paths over the ranks: N / world per GPU, same kernels, same exchanges.
    strong = None
    if world > 1 and not use_mlp:
        Ns = N // world
        sb_h = syn.make_flat_batch(Ns, n, m, horizon, 5000 + rank, flat0, hidden)
        sbatch = PathBatch(pin(sb_h["observations"]), pin(sb_h["actions"]), pin(sb_h["rewards"]),
                           sb_h["path_offsets"], sb_h["terminated"], device=local)

        def step_strong():
            sbatch.returns = sbatch.baseline = sbatch.adv = None
            return agent.train_step(Ns, gamma=GAMMA, gae_lambda=LAM, paths=sbatch)

        for _ in range(3):
            reset(); step_strong()
        barrier()
        sev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
               for _ in range(args.steps)]
        for k in range(args.steps):
            reset()
            flush.zero_()
            torch.cuda.synchronize()
            td.barrier()
            sev[k][0].record()
            step_strong()
            sev[k][1].record()
        barrier()
        ts_ = torch.tensor([sum(a.elapsed_time(bq) for a, bq in sev)], dtype=torch.float64, device=dev)
        td.all_reduce(ts_, op=td.ReduceOp.MAX)
        s_ms = float(ts_.item()) / args.steps
        strong = {"timesteps_total": Ns * world, "timesteps_per_gpu": Ns, "ms_per_step": s_ms,
                  "value": Ns * world / (s_ms * 1e-3), "unit": "timesteps/s",
                  "whole_update_hbm_frac": Ns * world * (56 * n + 16 * m + 88) / (s_ms * 1e-3) / 1e9 / (world * peaks()[0]),
                  "note": "same update on ONE batch of the named size split over the ranks by whole paths"}

    # ---- roofline of the dominant kernel: the FVP pass (10 of them per update) -----------------
    eng = agent._engine()
    reset()
    params = eng.dv.params("new")
    vec = torch.randn(d, dtype=torch.float32, device=dev) * 0.01
    out64 = torch.empty(d, dtype=torch.float64, device=dev)
    eng.cg_flags.zero_()

    def time_fvp(mode, reps=20):
        for _ in range(3):
            eng.fvp(params, resident.obs, vec, out64, cache_mode=mode)
        fe = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
        torch.cuda.synchronize()
        for a, bq in fe:
            a.record(); eng.fvp(params, resident.obs, vec, out64, cache_mode=mode); bq.record()
        torch.cuda.synchronize()
        return sorted(a.elapsed_time(bq) for a, bq in fe)[reps // 2]

    fvp_fill_ms = time_fvp(1)          # computes the forward pass, fills the activation cache
    cached = hidden and eng._fvp_cache(N) is not None
    fvp_rows_ms = time_fvp(2) if cached else fvp_fill_ms   # activation-cached product (fvp_rows.cuh)
    # small action dimensions: Jacobian-cached product (csrc/fvp_jac.cu) = what the CG solve runs
    jac = bool(cached and eng._jac_prepare(params, resident.obs))
    jac_fill_ms = None
    if jac:
        fe = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(5)]
        for a, bq in fe:
            eng._jac_key = None
            a.record(); eng._jac_prepare(params, resident.obs); bq.record()
        torch.cuda.synchronize()
        jac_fill_ms = sorted(a.elapsed_time(bq) for a, bq in fe)[2]
    fvp_ms = time_fvp(3) if jac else fvp_rows_ms
    hbm_peak, peak_src, sm_max = peaks()
    U = (n * hidden[0] + hidden[0] * hidden[1] + hidden[1] * m) if hidden else n * m
    # MACs per timestep: forward U, tangent U + (h1 h2 + h2 m), backward data h1 h2 + h2 m, weight grads U
    fill_flop = 2.0 * (5 * U - 2 * n * (hidden[0] if hidden else 0)) if hidden else 2.0 * 2 * U
    # `achieved` uses SURVEY 8(d)'s algorithmic figure of one product (forward + tangent + transpose =
    # fill_flop); the kernels issue fewer: the activation cache reads the forward pass back (- 2U), the
    # Jacobian cache also drops the tangent propagation and the backward data products (4U left)
    fvp_flop = fill_flop
    issued_flop = (4.0 * U if jac else fill_flop - 2.0 * U) if cached else fill_flop
    hp = (32 if max(hidden) <= 32 else 64) if hidden else 0
    fvp_bytes = 4.0 * n + (4.0 * 2 * hp if cached else 0.0) + (4.0 * m * hp if jac else 0.0)   # obs + caches
    ach = N * fvp_bytes / (fvp_ms * 1e-3) / 1e9
    fma_peak = 148 * 128 * 2 * sm_max * 1e6 / 1e12          # nominal fp32 FMA peak at max clock
    bf16_peak = None
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        bf16_peak = float(json.load(open(pk)).get("bf16_tflops", 0.0)) or None
    tensor_peak, tensor_src = (bf16_peak, peak_src) if bf16_peak else (1590.0, "fallback (B200_PROFILING.md)")
    traffic = None
    tp = os.path.join(ROOT, "profiles", "r01_fvp_traffic.json")
    if os.path.exists(tp) and args.workload == "C3-hopper":
        tj = json.load(open(tp))
        key = ("dram_bytes_per_timestep_jac" if jac else "dram_bytes_per_timestep_cached") if cached \
            else "dram_bytes_per_timestep"
        if key in tj:
            traffic = float(tj[key]) * N
    ach_tf = N * fvp_flop / (fvp_ms * 1e-3) / 1e12
    roofline = {
        "kernel": ("fvp_jac_kernel (analytic Fisher-vector product on cached activations and output-Jacobian "
                   "rows; all %d products of an update, after one jac_fill_kernel pass)" % CG_ITERS) if jac else
                  ("fvp_rows_kernel<use cache> (analytic Fisher-vector product; %d of the %d products of an "
                   "update, the first one also computes the forward pass and fills the cache)"
                   % (CG_ITERS - 1, CG_ITERS)) if cached else
                  "FVP kernel (analytic Fisher-vector product, %d per update)" % CG_ITERS,
        "bound": "tensor", "achieved": ach_tf, "peak": tensor_peak, "unit": "TFLOP/s",
        "frac": ach_tf / tensor_peak, "peak_source": tensor_src + ", dense bf16 burst",
        "traffic": traffic, "traffic_source": "ncu dram bytes of one launch (profiles/r01_fvp_traffic.json) scaled by N",
        "ms_per_launch": fvp_ms, "ms_per_launch_fill": fvp_fill_ms,
        "ms_per_launch_activation_cached": fvp_rows_ms, "ms_jac_fill": jac_fill_ms,
        "algorithmic_flop_per_timestep": fvp_flop, "issued_flop_per_timestep": issued_flop,
        "algorithmic_bytes_per_timestep": fvp_bytes,
        "note": "compute bound (%.0f flop per byte, SURVEY F13/H3).  The pass must hold fp32 accuracy "
                "(1e-4 parity gate, measured ~1e-6), so every algorithmic flop is 3 TF32 tensor flops "
                "(a_hi*b_hi + a_hi*b_lo + a_lo*b_hi) on mma.sync m16n8k8 (tcgen05 has no fp32-accurate "
                "kind and its per-instruction issue floor at N=64 gives no gain here, DESIGN.md).  "
                "`achieved` counts SURVEY 8(d)'s algorithmic flops of one product; the kernel issues "
                "`issued_flop_per_timestep` of them (the caches replace the rest by HBM reads), which "
                "`fp32_fma` and `tensor_flops_issued_tflops` are quoted on.  TF32 dense peak is half the "
                "bf16 peak used as denominator.  ncu: profiles/r01d_jac_ncu_summary.txt."
                % (fvp_flop / fvp_bytes),
        "tensor_flops_issued_tflops": 3.0 * N * issued_flop / (fvp_ms * 1e-3) / 1e12,
        "fp32_fma": {"achieved_tflops": N * issued_flop / (fvp_ms * 1e-3) / 1e12, "peak_tflops_nominal": fma_peak,
                     "frac": N * issued_flop / (fvp_ms * 1e-3) / 1e12 / fma_peak, "flop_per_timestep": issued_flop},
        "hbm": {"achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak},
        "whole_update_hbm": {"bytes_per_timestep": 56 * n + 16 * m + 88,
                             "achieved": N * (56 * n + 16 * m + 88) / (ms_step * 1e-3) / 1e9,
                             "frac": N * (56 * n + 16 * m + 88) / (ms_step * 1e-3) / 1e9 / hbm_peak},
    }

    if args.phases and rank == 0:
        phase_breakdown(agent, resident, eng, N, n, m, hbm_peak)

    mlp_extra = None
    if rank == 0 and world == 1 and not use_mlp:
        try:
            mlp_extra = mlpbase_timing(spec, resident, N)
        except Exception as e:   # e.g. obs_dim too wide for the fit kernels' shared-memory plan
            mlp_extra = {"unavailable": str(e)[:200]}
    cb = None
    if rank == 0 and not args.no_cpu_baseline:
        cb = cpu_baseline(args.workload)
    if rank == 0:
        print(json.dumps({
            "metric": "NPG update timesteps/sec", "value": value, "unit": "timesteps/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "%s: NPG + %s + %s, obs %d act %d, %d timesteps per GPU "
                                   "(%d paths), gamma %.3f lambda %.2f, %d CG iters, damping %g"
                                   % (args.workload, "MLP%s" % (hidden,) if hidden else "LinearPolicy",
                                      "MLPBaseline(epochs=1, batch 64: %d serial Adam steps)" % base.last_fit_steps
                                      if use_mlp else "LinearBaseline", n, m,
                                      N, len(off) - 1, GAMMA, LAM, CG_ITERS, DAMPING),
                       "parallelism": "dp%d (whole paths per GPU)" % world,
                       "exchange": ("none (1 GPU)" if world == 1 else
                                    "peer-memory kernels over NVLink, CG update fused with its all-reduce (csrc/peer.cu)"
                                    if agent._engine().dist.peer is not None else
                                    "NCCL all-reduces (%s)" % agent._engine().dist.peer_note),
                       "l2": "256 MB buffer written between timed steps (outside the event brackets)",
                       "policy_reset_each_step": True},
            "clocks": clk, "e2e": e2e, "gpu_launches": launches, "roofline": roofline,
            **({"strong_scaling": strong} if strong else {}),
            "cpu_baseline": cb, "mlp_baseline": mlp_extra}))
    if world > 1:
        td.destroy_process_group()


def mlpbase_timing(spec, batch, N):
    """MLPBaseline (BASELINE.json configs[2]) on the same batch: predict pass and one fit()
    (epochs=1, batch 64), device times.  Reported beside the headline, not inside it: the fit
    is a serial chain of int(N/64)-1 dependent Adam steps."""
    import torch
    from mjrl_b200.baselines.mlp_baseline import MLPBaseline
    from mjrl_b200.utils import process_samples as ps
    torch.manual_seed(7)
    np.random.seed(7)
    mb = MLPBaseline(spec, epochs=1)
    if batch.returns is None:
        ps.returns_on_device(batch, GAMMA)
    mb.predict_batch(batch); mb.fit_batch(batch, sync=False)        # warm-up
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    ev[0].record(); mb.predict_batch(batch); ev[1].record(); mb.fit_batch(batch, sync=False); ev[2].record()
    torch.cuda.synchronize()
    p_ms, f_ms = ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2])
    return {"predict_ms": p_ms, "predict_tflops_fp32": 2.0 * (15 * 128 + 128 * 128 + 128) * N / p_ms / 1e9
            if spec.observation_dim == 11 else None,
            "fit_ms": f_ms, "fit_steps": mb.last_fit_steps,
            "fit_us_per_step": 1e3 * f_ms / max(mb.last_fit_steps, 1),
            "kernel": "fit_cluster_kernel: one 8-CTA cluster, model parallel over hidden units, DSMEM exchanges"}


def phase_breakdown(agent, batch, eng, N, n, m, hbm_peak):
    """Per-phase device times of one update, each against its binding roofline (stderr)."""
    import torch
    from mjrl_b200.utils import process_samples as ps

    def timed(fn, reps=10):
        fn(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record(); torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    base = agent.baseline
    rows = []
    rows.append(("returns scan", timed(lambda: ps.returns_on_device(batch, GAMMA)), 16))
    rows.append(("linbase predict", timed(lambda: base.predict_batch(batch)), 4 * n + 8))
    bvals = base.predict_batch(batch)
    rows.append(("GAE scan", timed(lambda: ps.advantages_on_device(batch, bvals, GAMMA, LAM)), 24))
    rows.append(("whiten (moments + cast)", timed(lambda: eng.whiten(batch.adv)), 20))
    w = eng.whiten(batch.adv)
    p = eng.dv.params("new")
    rows.append(("policy gradient", timed(lambda: eng.gradient(p, batch.obs, batch.act, w)), 4 * n + 8 * m + 12))
    ll, mean = eng.old_cache(p, batch.obs, batch.act)
    rows.append(("surrogate+KL eval", timed(lambda: eng.evaluate(p, batch.obs, batch.act, w, ll, mean, p[-m:], float(N))),
                 4 * n + 8 * m + 12))
    eng.cg_flags.zero_()
    rows.append(("FVP", timed(lambda: eng.fvp(p, batch.obs, eng.cg_p32, eng.cg_z)), 4 * n))
    rows.append(("linbase fit (Gram+Cholesky)", timed(lambda: base.fit_batch(batch, sync=False)), 4 * n + 8))
    for name, ms, bpt in rows:
        gbs = N * bpt / (ms * 1e-3) / 1e9
        sys.stderr.write("phase %-28s %8.3f ms  %7.1f GB/s algorithmic  (%.1f%% of measured HBM)\n"
                         % (name, ms, gbs, 100 * gbs / hbm_peak))


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
