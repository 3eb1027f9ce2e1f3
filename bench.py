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
describes the dominant kernel (the FVP pass); `cpu_baseline` is the reference's own update
(staged in oracle/_ref; the oracle port when nothing is staged) timed on this box's host cores on a
bounded 100k-timestep sample; `parity` is the CUDA path against it on that same sample.
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
    ap.add_argument("--algo", default="npg", choices=["npg", "trpo", "dapg"],
                    help="update rule inside the timed step: NPG (default, the metric), TRPO with its KL line search "
                         "(BASELINE.json configs[3]) or DAPG with 25 synthetic demonstration paths (configs[4])")
    ap.add_argument("--phases", action="store_true", help="also print a per-phase breakdown to stderr")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


# ------------------------------------------------------------------------------------------
# CPU baseline / reference arm.  `kind: "reference"`: the reference's OWN modules (process_samples,
# NPG.train_from_paths, LinearBaseline.fit) staged in oracle/_ref by oracle/build_ref.py (the package is pure
# Python; the staged copy travels to the GPU box, /root/reference does not).  `kind: "port"`: the oracle
# restatement (oracle/mjrl_oracle.py, pinned bit-exact to the reference by tests/golden), used when nothing is
# staged.
# ------------------------------------------------------------------------------------------
def cpu_update_setup(shape, N, seed):
    import torch
    from mjrl_b200.utils import synthetic as syn
    from oracle import mjrl_oracle as O
    from oracle import ref_runner as RR
    n, m, hidden, horizon, _ = syn.SHAPES[shape]
    torch.set_num_threads(os.cpu_count())
    flat = O.init_policy_params(n, m, hidden, seed=123)
    b0 = syn.make_flat_batch(N, n, m, horizon, seed, flat, hidden)
    b1 = syn.make_flat_batch(N, n, m, horizon, seed + 1, flat, hidden)
    kind = "reference" if RR.available() else "port"
    if kind == "reference":
        R = RR.load()
        base = R.lb.LinearBaseline(R.EnvSpec(n, m, horizon, 1))
        p0 = syn.flat_to_paths(b0)
        R.ps.compute_returns(p0, GAMMA)
    else:
        base = O.LinearBaseline()
        p0 = syn.flat_to_paths(b0)
        O.compute_returns(p0, GAMMA)
    base.fit(p0)                                   # steady state: predict() is non-trivial (F8)
    return dict(kind=kind, O=O, RR=RR, syn=syn, dims=(n, m, hidden, horizon), flat=flat, b0=b0, b1=b1, base=base)


def cpu_update_once(c, keep=False):
    import copy
    n, m, hidden, horizon = c["dims"]
    paths = c["syn"].flat_to_paths(c["b1"])
    bl = copy.deepcopy(c["base"])
    if c["kind"] == "reference":
        pol, _, agent = c["RR"].make_agent(n, m, hidden, horizon, c["flat"], step=STEP, cg_iters=CG_ITERS, damping=DAMPING)
        agent.baseline = bl
        t = time.perf_counter()
        c["RR"].full_iteration(agent, paths, GAMMA, LAM)
        dt = time.perf_counter() - t
        if keep:
            c["last"] = dict(new_params=pol.get_param_values(), coeffs=np.asarray(bl._coeffs), paths=paths)
    else:
        O = c["O"]
        pol = O.GaussianPolicy(n, m, hidden, c["flat"])
        t = time.perf_counter()
        out = O.full_npg_iteration(pol, bl, paths, GAMMA, LAM, normalized_step_size=STEP,
                                   cg_iters=CG_ITERS, damping=DAMPING)
        dt = time.perf_counter() - t
        if keep:
            c["last"] = dict(new_params=out["new_params"], coeffs=np.asarray(bl._coeffs), paths=paths)
    return dt


def _cpu_desc(kind):
    return ("the reference's own modules (mjrl.utils.process_samples, mjrl.algos.npg_cg.NPG.train_from_paths, "
            "mjrl.baselines.linear_baseline.LinearBaseline.fit; numpy + CPU torch autograd, python-loop scans are "
            "single-threaded)" if kind == "reference" else
            "oracle port of the reference update (numpy + CPU torch autograd, python-loop scans are single-threaded)")


def cpu_baseline(shape, reps=5):
    import torch
    c = cpu_update_setup(shape, CPU_SAMPLE_N, 900)
    cpu_update_once(c)                             # warm-up (first torch call is slow)
    ts = sorted(cpu_update_once(c, keep=True) for _ in range(reps))
    med = ts[len(ts) // 2]
    cb = {"value": CPU_SAMPLE_N / med, "unit": "timesteps/s", "cores": torch.get_num_threads(),
          "kind": c["kind"],
          "sample": "%s, %d timesteps of %s, median of %d after 1 warm-up"
                    % (_cpu_desc(c["kind"]), CPU_SAMPLE_N, shape, reps),
          "versions": {"numpy": np.__version__, "torch": torch.__version__}}
    return cb, c


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    c = cpu_update_setup(args.workload, CPU_SAMPLE_N, 900)
    for _ in range(max(1, args.warmup)):
        cpu_update_once(c)
    ts = [cpu_update_once(c) for _ in range(args.steps)]
    tot = sum(ts)
    val = CPU_SAMPLE_N * args.steps / tot
    cb = {"value": val, "unit": "timesteps/s", "cores": torch.get_num_threads(), "kind": c["kind"],
          "sample": "%d-timestep sample of %s per step: %s" % (CPU_SAMPLE_N, args.workload, _cpu_desc(c["kind"]))}
    print(json.dumps({
        "impl": "reference", "metric": "NPG update timesteps/sec", "value": val,
        "unit": "timesteps/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * tot / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "%s NPG + MLP(64,64) + LinearBaseline, %s on a bounded "
                               "sample of %d timesteps/step" % (args.workload,
                                                                "the reference's own CPU update" if c["kind"] == "reference"
                                                                else "CPU oracle port", CPU_SAMPLE_N)},
        "cpu_baseline": cb,
        "e2e": {"value": val, "unit": "timesteps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def parity_at_sample_size(c, device):
    """The CUDA path on the SAME 100k-timestep batch the CPU arm just pushed through the reference: parity at a
    size where every CTA of the policy kernels accumulates over thousands of timesteps (the unit tests run
    N <= 40k).  Gates (north_star): returns / advantages / baseline predictions 1e-6, gradient / FVP / step 1e-4;
    the 10-iteration step is compared through the fp64 truth (the reference's own fp32 CG is only
    reproducible to ~1e-3 on this shape, so both are measured against the all-fp64 solve)."""
    import torch
    from mjrl_b200 import _lib
    from mjrl_b200.utils.gym_env import EnvSpec
    from mjrl_b200.policies.gaussian_mlp import MLP
    from mjrl_b200.policies.gaussian_linear import LinearPolicy
    from mjrl_b200.baselines.linear_baseline import LinearBaseline
    from mjrl_b200.algos.npg_cg import NPG
    O, syn = c["O"], c["syn"]
    n, m, hidden, horizon = c["dims"]
    flat = c["flat"]
    rel = lambda a, b: float(np.linalg.norm(np.asarray(a, np.float64).ravel() - np.asarray(b, np.float64).ravel()) /
                             max(np.linalg.norm(np.asarray(b, np.float64).ravel()), 1e-300))
    spec = EnvSpec(n, m, horizon, 1)
    pol = MLP(spec, hidden_sizes=hidden, seed=123) if hidden else LinearPolicy(spec, seed=123)
    pol.set_param_values(flat.copy(), True, True)
    base = LinearBaseline(spec)
    p0 = syn.flat_to_paths(c["b0"])
    from mjrl_b200.utils import process_samples as ps
    ps.compute_returns(p0, GAMMA)
    base.fit(p0)
    agent = NPG(None, pol, base, normalized_step_size=STEP, FIM_invert_args={"iters": CG_ITERS, "damping": DAMPING})
    paths = syn.flat_to_paths(c["b1"])
    # the reference's own call sequence (batch_reinforce.py:86-101) through the drop-in modules: the path dicts
    # receive returns / baseline / advantages like the reference's
    ps.compute_returns(paths, GAMMA)
    ps.compute_advantages(paths, base, GAMMA, LAM)
    agent.train_from_paths(paths)
    base.fit(paths)
    ref = c["last"]
    cat = lambda ps_, k: np.concatenate([p[k] for p in ps_])
    out = {"timesteps": int(c["b1"]["observations"].shape[0]), "against": c["kind"],
           "returns_rel": rel(cat(paths, "returns"), cat(ref["paths"], "returns")),
           "baseline_pred_rel": rel(cat(paths, "baseline"), cat(ref["paths"], "baseline")),
           "advantages_rel": rel(cat(paths, "advantages"), cat(ref["paths"], "advantages")),
           "baseline_coeffs_pred_rel": None, "kl_abs": None}
    obs, act = c["b1"]["observations"], c["b1"]["actions"]
    adv = cat(ref["paths"], "advantages")
    adv_w = (adv - np.mean(adv)) / (np.std(adv) + 1e-6)
    # gradient and Fisher-vector product of the starting policy, CUDA path vs oracle autograd (fp32) and fp64 truth
    opol = O.GaussianPolicy(n, m, hidden, flat)
    g_ref = O.flat_vpg(opol, obs, act, adv_w)
    pol2 = MLP(spec, hidden_sizes=hidden, seed=123) if hidden else LinearPolicy(spec, seed=123)
    pol2.set_param_values(flat.copy(), True, True)
    ag2 = NPG(None, pol2, LinearBaseline(spec), normalized_step_size=STEP,
              FIM_invert_args={"iters": CG_ITERS, "damping": DAMPING})
    out["grad_rel"] = rel(ag2.flat_vpg(obs, act, adv_w), g_ref)
    hv_ref = O.hvp(opol, obs, act, g_ref, DAMPING)
    out["fvp_rel"] = rel(ag2.HVP(obs, act, g_ref, DAMPING), hv_ref)
    # the product variant the CG solve runs at this size (tensor-memory kernel when the shape has it)
    eng = ag2._engine()
    d_obs = torch.from_numpy(np.ascontiguousarray(obs, dtype=np.float32)).to(eng.device)
    params = eng.dv.params("new")
    v32 = torch.from_numpy(g_ref.astype(np.float32)).to(eng.device)
    o64 = torch.empty(eng.dv.d, dtype=torch.float64, device=eng.device)
    eng.cg_flags.zero_()
    mode = 0
    if hidden and eng._fvp_cache(d_obs.shape[0]) is not None:
        eng.fvp(params, d_obs, v32, o64, cache_mode=1)
        eng._cache_key = eng._hold(params, d_obs)
        mode = eng._cached_mode(params, d_obs)
    eng.fvp(params, d_obs, v32, o64, cache_mode=mode)
    hv_dev = (o64 / d_obs.shape[0]).float().cpu().numpy() + np.float32(DAMPING) * g_ref
    out["fvp_rel_cg_kernel"] = rel(hv_dev, hv_ref)
    out["fvp_cg_kernel_mode"] = {0: "uncached", 2: "activation cache (rows kernel)", 3: "Jacobian rows (rows kernel)",
                                 4: "tensor memory (tcgen05)"}.get(mode, str(mode))
    if hidden:
        tp = O.TruthPolicy(opol)
        gt = tp.vpg(obs, act, adv_w)
        out["grad_rel_vs_truth"] = rel(ag2.flat_vpg(obs, act, adv_w), gt)
        out["fvp_rel_vs_truth"] = rel(hv_dev, tp.fvp(obs, g_ref.astype(np.float64), DAMPING))
        xs = O.truth_cg(tp, obs, gt, cg_iters=CG_ITERS, damping=DAMPING)
        step_true = np.sqrt(abs(STEP / (gt @ xs + 1e-20))) * xs
        out["step_rel_vs_truth"] = rel(pol.get_param_values().astype(np.float64) - flat, step_true)
        out["reference_step_rel_vs_truth"] = rel(ref["new_params"].astype(np.float64) - flat, step_true)
    out["step_rel_vs_reference"] = rel(pol.get_param_values().astype(np.float64) - flat,
                                       ref["new_params"].astype(np.float64) - flat)
    # baseline coefficients through their predictions (cond(A) up to 1e9: compare X c, not c)
    X = np.concatenate([O.linear_features(p) for p in ref["paths"]])
    out["baseline_coeffs_pred_rel"] = rel(X @ np.asarray(base._coeffs), X @ ref["coeffs"])
    out["kl_abs_vs_target"] = abs(agent.last_update["kl_dist"] - 0.5 * STEP)
    out["kl_cuda"] = agent.last_update["kl_dist"]
    out.pop("kl_abs")
    gates = {"returns_rel": 1e-6, "baseline_pred_rel": 1e-6, "advantages_rel": 1e-6, "grad_rel": 1e-4,
             "fvp_rel": 1e-4, "fvp_rel_cg_kernel": 1e-4, "baseline_coeffs_pred_rel": 1e-6}
    out["gates"] = gates
    ok = all(out[k] <= v for k, v in gates.items())
    if hidden:   # H1 protocol for the 10-iteration step
        ok = ok and out["step_rel_vs_truth"] <= max(1e-4, 1.05 * out["reference_step_rel_vs_truth"]) + 1e-6
    out["pass"] = bool(ok)
    return out


def exchange_profile(agent, epoch0, steps):
    """Per-exchange timing of the peer-memory kernels (csrc/peer.cu) over the timed region, from the %globaltimer
    stamps each exchange kernel records on this rank (MJRL_B200_PEER_STAMPS=1): wait = entry -> every peer's flag seen
    (what a rank spends waiting for the slowest peer), pull = flags seen -> sums pulled over NVLink, rest = the fused
    CG arithmetic.  None when the exchange runs over NCCL."""
    import ctypes
    import numpy as np
    from mjrl_b200 import _lib
    pe = agent._engine().dist.peer
    if pe is None:
        return None
    n = min(pe.epoch - epoch0, 4096)
    if n <= 0:
        return None
    buf = (ctypes.c_uint64 * (4 * n))()
    rc = _lib.load().mjrl_peer_stamps(pe.h, buf, ctypes.c_int64(pe.epoch - n + 1), ctypes.c_int64(n))
    if rc != 0:
        return None
    a = np.frombuffer(buf, dtype=np.uint64).reshape(n, 4).astype(np.int64)
    a = a[(a > 0).all(axis=1)]
    if len(a) == 0:
        return None
    wait, pull, rest, tot = (a[:, 1] - a[:, 0]) / 1e3, (a[:, 2] - a[:, 1]) / 1e3, (a[:, 3] - a[:, 2]) / 1e3, (a[:, 3] - a[:, 0]) / 1e3
    q = lambda v: {"median_us": float(np.median(v)), "mean_us": float(np.mean(v)), "p90_us": float(np.percentile(v, 90))}
    per_update = len(a) / max(1, steps)
    return {"exchanges_per_update": per_update, "wait_for_peers": q(wait), "pull_over_nvlink": q(pull), "fused_arithmetic": q(rest),
            "kernel_total": q(tot), "ms_per_update_in_exchange_kernels": float(np.sum(tot)) / 1e3 / max(1, steps),
            "ms_per_update_waiting": float(np.sum(wait)) / 1e3 / max(1, steps),
            "source": "%globaltimer stamps inside the exchange kernels of rank 0 over the timed region (mjrl_peer_stamps)"}


def rank_invariance(dev, world, rank):
    """Multi-GPU correctness on the bench line itself (the 1-GPU test box skips the torchrun tests): the same small
    batch updated (a) sharded by whole paths over all ranks and (b) on one GPU with no exchange; every rank computes
    (b) itself.  Relative difference of the NPG step and of the refit baseline's predictions."""
    import torch.distributed as td
    from mjrl_b200.sharding import shard_paths
    from mjrl_b200.utils import synthetic as syn
    from mjrl_b200.utils.gym_env import EnvSpec
    from mjrl_b200.policies.gaussian_mlp import MLP
    from mjrl_b200.baselines.linear_baseline import LinearBaseline
    from mjrl_b200.algos.npg_cg import NPG
    from mjrl_b200.algos.engine import Dist
    from mjrl_b200.batch import drop_cache
    from oracle import mjrl_oracle as O
    n, m, hidden, horizon = 11, 3, (64, 64), 250
    Nt = 40000 * world       # >= 32768 per rank: the sharded run uses the same product kernel as the headline
    spec = EnvSpec(n, m, horizon, 1)
    flat = MLP(spec, hidden_sizes=hidden, seed=123).get_param_values()
    hb = syn.make_flat_batch(Nt, n, m, horizon, 7000, flat, hidden)
    res = {}
    for mode in ("sharded", "single"):
        pol = MLP(spec, hidden_sizes=hidden, seed=123)
        base = LinearBaseline(spec)
        ag = NPG(None, pol, base, normalized_step_size=STEP, FIM_invert_args={"iters": CG_ITERS, "damping": DAMPING})
        paths = syn.flat_to_paths(hb)
        if mode == "sharded":
            ag.dist = Dist()
            paths = shard_paths(paths, rank, world)
        else:
            ag.dist = Dist.__new__(Dist); ag.dist.td = None; ag.dist.world = 1; ag.dist.rank = 0
            ag.dist.calls = 0; ag.dist.group = None; ag.dist.peer = None; ag.dist.peer_note = "single"
        drop_cache()
        ag.train_step(len(paths), gamma=GAMMA, gae_lambda=LAM, paths=paths)
        res[mode] = (pol.get_param_values().astype(np.float64) - flat, np.asarray(base._coeffs, dtype=np.float64))
    X = np.concatenate([O.linear_features(p) for p in syn.flat_to_paths(hb)[:50]])
    rel = lambda a, b: float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))
    return {"timesteps": Nt, "step_rel": rel(res["sharded"][0], res["single"][0]),
            "coef_rel": rel(X @ res["sharded"][1], X @ res["single"][1]),
            "note": "sharded over %d ranks vs the same batch on one GPU without any exchange (10 CG iterations)" % world}


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
        self.nvml, self.stop_flag = None, False

    def _nvml_loop(self):
        nv, h = self.nvml
        # NVML clocks-event (throttle) reason bits
        bits = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))
        while not self.stop_flag:
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
                try:
                    rs = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                row = [str(self.index), str(sm), str(mx), "", ""] + ["Active" if rs & b else "Not Active" for _, b in bits]
                self.rows.append(row)
            except Exception:
                pass
            time.sleep(0.005)

    def start(self):
        # NVML in-process (a query takes microseconds: 8 ranks each spawning nvidia-smi got no sample inside a 0.1 s timed region),
        # nvidia-smi as the fallback
        try:
            import pynvml as nv
            nv.nvmlInit()
            self.nvml = (nv, nv.nvmlDeviceGetHandleByIndex(self.index))
            self.th = threading.Thread(target=self._nvml_loop, daemon=True)
            self.th.start()
            return
        except Exception:
            self.nvml = None
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
        if self.nvml is not None:
            self.stop_flag = True
            self.th.join(timeout=1)
        elif not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        else:
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
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvml" if self.nvml is not None else "nvidia-smi"}


def run_b200(args):
    import torch
    import torch.distributed as td
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        os.environ.setdefault("MJRL_B200_PEER_STAMPS", "1")    # exchange kernels stamp %globaltimer (exchange_profile below)
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
    fim = {"iters": CG_ITERS, "damping": DAMPING}
    if args.algo == "trpo":
        from mjrl_b200.algos.trpo import TRPO
        agent = TRPO(None, pol, base, kl_dist=0.5 * STEP, FIM_invert_args=fim)
        _tfp = agent.train_from_paths
        agent.train_from_paths = lambda paths, **kw: _tfp(paths, verbose=False, **kw)
    elif args.algo == "dapg":
        from mjrl_b200.algos.dapg import DAPG
        demo = syn.flat_to_paths(syn.make_flat_batch(25 * min(horizon, 200), n, m, min(horizon, 200), 4242, flat0, hidden))
        agent = DAPG(None, pol, base, demo_paths=demo, kl_dist=0.5 * STEP, lam_0=1.0, lam_1=0.95, FIM_invert_args=fim)
    else:
        agent = NPG(None, pol, base, normalized_step_size=STEP, FIM_invert_args=fim)

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
            if args.algo == "dapg":
                agent.iter_count = 0.0

    resident = PathBatch(h_obs, h_act, h_rew, off, term, device=local)

    def step_resident():
        resident.returns = resident.baseline = resident.adv = None
        return agent.train_step(N, gamma=GAMMA, gae_lambda=LAM, paths=resident)

    # e2e inputs: pinned HOST buffers as a producer of this package leaves them -- observations and actions in
    # fp32 (the cast the reference applies on every call, gaussian_mlp.py:103-104, done once by the producer; the
    # values are fp32-representable), rewards in fp64.  `e2e_fp64_paths` times the same step from fp64
    # observation / action buffers, i.e. arrays laid out like the reference's list of path dicts.
    h_obs32, h_act32 = pin(hb["observations"].astype(np.float32)), pin(hb["actions"].astype(np.float32))

    def step_e2e(fp64=False):
        b = PathBatch(h_obs if fp64 else h_obs32, h_act if fp64 else h_act32, h_rew, off, term,
                      device=local)                                # H2D from pinned host memory
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
    _pe = agent._engine().dist.peer if world > 1 else None
    e0_peer = _pe.epoch if _pe is not None else 0
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
    xprof = exchange_profile(agent, e0_peer, args.steps) if world > 1 else None
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
    d2h0 = _lib.D2H_BYTES[0]
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
    d2h = (_lib.D2H_BYTES[0] - d2h0) // args.steps      # counted where the update reads results back (_lib.to_host)
    e2e = {"value": N * world / e2e_s, "unit": "timesteps/s", "ms_per_step": 1e3 * e2e_s,
           "h2d_bytes_per_step": int(b.h2d_bytes), "d2h_bytes_per_step": int(d2h),
           "host_buffers": "pinned; observations / actions fp32, rewards fp64"}
    # the same from fp64 observation / action buffers (the reference's `paths` dtype)
    for _ in range(2):
        reset(); step_e2e(True)
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        reset()
        b64, _ = step_e2e(True)
        torch.cuda.synchronize()
    barrier()
    te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        td.all_reduce(te, op=td.ReduceOp.MAX)
    e2e64_s = float(te.item()) / args.steps
    e2e["fp64_paths"] = {"value": N * world / e2e64_s, "ms_per_step": 1e3 * e2e64_s,
                         "h2d_bytes_per_step": int(b64.h2d_bytes)}

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
    fvp_jac_ms = time_fvp(3) if jac else None
    # hidden 64 x 64: tensor-memory product (csrc/fvp_tc.cu, tcgen05) = what the CG solve runs when the shape has it
    tc = bool(cached and eng._tc_prepare(params, resident.obs))
    tc_fill_ms = None
    if tc:
        fe = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(5)]
        for a, bq in fe:
            eng._tc_key = None
            a.record(); eng._tc_prepare(params, resident.obs); bq.record()
        torch.cuda.synchronize()
        tc_fill_ms = sorted(a.elapsed_time(bq) for a, bq in fe)[2]
    fvp_tc_ms = time_fvp(4) if tc else None
    fvp_ms = fvp_tc_ms if tc else (fvp_jac_ms if jac else fvp_rows_ms)
    jac_used = jac and not tc
    hbm_peak, peak_src, sm_max = peaks()
    U = (n * hidden[0] + hidden[0] * hidden[1] + hidden[1] * m) if hidden else n * m
    # MACs per timestep: forward U, tangent U + (h1 h2 + h2 m), backward data h1 h2 + h2 m, weight grads U
    fill_flop = 2.0 * (5 * U - 2 * n * (hidden[0] if hidden else 0)) if hidden else 2.0 * 2 * U
    # `achieved` uses SURVEY 8(d)'s algorithmic figure of one product (forward + tangent + transpose =
    # fill_flop); the kernels issue fewer: the activation cache reads the forward pass back (- 2U), the
    # Jacobian cache also drops the tangent propagation and the backward data products (4U left)
    fvp_flop = fill_flop
    issued_flop = (4.0 * U if jac_used else fill_flop - 2.0 * U) if cached else fill_flop
    hp = (32 if max(hidden) <= 32 else 64) if hidden else 0
    fvp_bytes = 4.0 * n + (4.0 * 2 * hp if cached else 0.0) + (4.0 * m * hp if jac_used else 0.0)   # obs + caches
    ach = N * fvp_bytes / (fvp_ms * 1e-3) / 1e9
    fma_peak = 148 * 128 * 2 * sm_max * 1e6 / 1e12          # nominal fp32 FMA peak at max clock
    bf16_peak = None
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        bf16_peak = float(json.load(open(pk)).get("bf16_tflops", 0.0)) or None
    tensor_peak, tensor_src = (bf16_peak, peak_src) if bf16_peak else (1590.0, "fallback (B200_PROFILING.md)")
    traffic = None
    tp = os.path.join(ROOT, "profiles", "r02_fvp_traffic.json" if tc else "r01_fvp_traffic.json")
    if os.path.exists(tp) and args.workload == "C3-hopper":
        tj = json.load(open(tp))
        key = "dram_bytes_per_timestep_tc" if tc else \
            ("dram_bytes_per_timestep_jac" if jac else "dram_bytes_per_timestep_cached") if cached \
            else "dram_bytes_per_timestep"
        if key in tj:
            traffic = float(tj[key]) * N
    # measured non-bf16 peaks of this pool's B200 (profiles/peaks_measured.json: probes under profiles/probes)
    pm = {}
    pmf = os.path.join(ROOT, "profiles", "peaks_measured.json")
    if os.path.exists(pmf):
        pm = json.load(open(pmf))
    fma_peak_meas = pm.get("fp32_fma_tflops")
    tf32_tc_peak = pm.get("tf32_tcgen05_tflops")
    ach_tf = N * fvp_flop / (fvp_ms * 1e-3) / 1e12
    tf32_issued = 3.0 * N * issued_flop / (fvp_ms * 1e-3) / 1e12
    roofline = {
        "kernel": ("fvp_tc_kernel (analytic Fisher-vector product on tcgen05 tensor cores: tangent + backward pass per "
                   "96-timestep tile, A operands and accumulators in tensor memory, kind::tf32 with hi/lo error "
                   "compensation; %d of the %d products of an update, after the cache-filling first product and one "
                   "tc_fill relayout pass)" % (CG_ITERS - 1, CG_ITERS)) if tc else
                  ("fvp_jac_kernel (analytic Fisher-vector product on cached activations and output-Jacobian "
                   "rows; all %d products of an update, after one jac_fill_kernel pass)" % CG_ITERS) if jac else
                  ("fvp_rows_kernel<use cache> (analytic Fisher-vector product; %d of the %d products of an "
                   "update, the first one also computes the forward pass and fills the cache)"
                   % (CG_ITERS - 1, CG_ITERS)) if cached else
                  "FVP kernel (analytic Fisher-vector product, %d per update)" % CG_ITERS,
        "bound": "tensor", "achieved": ach_tf, "peak": tensor_peak, "unit": "TFLOP/s",
        "frac": ach_tf / tensor_peak, "peak_source": tensor_src + ", dense bf16 burst",
        "traffic": traffic,
        "traffic_source": "ncu dram bytes of one launch (profiles/%s) scaled by N" % os.path.basename(tp),
        "ms_per_launch": fvp_ms, "ms_per_launch_fill": fvp_fill_ms,
        "ms_per_launch_activation_cached": fvp_rows_ms, "ms_per_launch_jacobian_cached": fvp_jac_ms,
        "ms_per_launch_tensor_memory": fvp_tc_ms, "ms_jac_fill": jac_fill_ms, "ms_tc_fill": tc_fill_ms,
        "algorithmic_flop_per_timestep": fvp_flop, "issued_flop_per_timestep": issued_flop,
        "algorithmic_bytes_per_timestep": fvp_bytes,
        "note": "compute bound (%.0f flop per byte, SURVEY F13/H3).  The pass must hold fp32 accuracy (1e-4 parity "
                "gate, measured ~1e-6) and no tensor-core kind is fp32-accurate, so every algorithmic flop is 3 "
                "kind::tf32 flops (hi*hi + lo*hi + hi*lo; the hi plane is the raw fp32 word, which the tensor core "
                "truncates).  `achieved` counts SURVEY 8(d)'s algorithmic flops of one product against the measured "
                "dense-bf16 peak; the kernel issues `issued_flop_per_timestep` of them (h1, h2 come from the cache), "
                "i.e. `tensor_flops_issued_tflops` of TF32 work, quoted next to the measured tcgen05 TF32 peak in "
                "`tf32`.  ncu: profiles/r02_tc_ncu_summary.txt." % (fvp_flop / fvp_bytes),
        "tensor_flops_issued_tflops": tf32_issued,
        "tf32": {"achieved_tflops": tf32_issued, "peak_tflops_measured": tf32_tc_peak,
                 "frac": (tf32_issued / tf32_tc_peak) if tf32_tc_peak else None,
                 "peak_source": pm.get("tf32_tcgen05_source")},
        "fp32_fma": {"achieved_tflops": N * issued_flop / (fvp_ms * 1e-3) / 1e12, "peak_tflops_nominal": fma_peak,
                     "peak_tflops_measured": fma_peak_meas,
                     "frac": N * issued_flop / (fvp_ms * 1e-3) / 1e12 / (fma_peak_meas or fma_peak),
                     "flop_per_timestep": issued_flop},
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
    configs2 = None
    if rank == 0 and world == 1 and not use_mlp and args.algo == "npg" and args.workload == "C3-hopper":
        try:
            configs2 = configs2_value(spec, hidden, flat0, (h_obs32, h_act32, h_rew, off, term), local, N)
        except Exception as e:
            configs2 = {"unavailable": repr(e)[:200]}
    cb = parity = None
    if rank == 0 and not args.no_cpu_baseline:
        cb, cctx = cpu_baseline(args.workload)
        if world == 1:
            try:
                parity = parity_at_sample_size(cctx, dev)
            except Exception as e:
                parity = {"error": repr(e)[:300], "pass": False}
    rinv = None
    if world > 1 and not use_mlp:
        rinv = rank_invariance(dev, world, rank)
    if rank == 0:
        print(json.dumps({
            "metric": "NPG update timesteps/sec", "value": value, "unit": "timesteps/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "%s: %s + %s + %s, obs %d act %d, %d timesteps per GPU "
                                   "(%d paths), gamma %.3f lambda %.2f, %d CG iters, damping %g"
                                   % (args.workload, {"npg": "NPG", "trpo": "TRPO (KL line search)",
                                                      "dapg": "DAPG (25 demonstration paths)"}[args.algo],
                                      "MLP%s" % (hidden,) if hidden else "LinearPolicy",
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
            **({"value_configs2": configs2.get("value"), "configs2": configs2} if configs2 else {}),
            **({"strong_scaling": strong} if strong else {}),
            **({"rank_invariance": rinv} if rinv else {}),
            **({"exchange_profile": xprof} if xprof else {}),
            **({"parity": parity} if parity else {}),
            "cpu_baseline": cb, "mlp_baseline": mlp_extra}))
    if world > 1:
        td.destroy_process_group()


def configs2_value(spec, hidden, flat0, host, local, N):
    """BASELINE.json configs[2] VERBATIM: the same 1M-timestep update with `MLPBaseline` (epochs 1, batch 64) inside
    the timed region -- returns, MLP predict, GAE, gradient, CG, step, evaluation and the baseline's fit, a serial
    chain of int(N/64)-1 dependent Adam steps (mlp_baseline.py:83-98) that no amount of data parallelism shortens.
    CUDA events, 1 warm-up + 3 timed steps, policy and baseline reset before each."""
    import torch
    from mjrl_b200.policies.gaussian_mlp import MLP
    from mjrl_b200.baselines.mlp_baseline import MLPBaseline
    from mjrl_b200.algos.npg_cg import NPG
    from mjrl_b200.batch import PathBatch
    from mjrl_b200.utils import process_samples as ps
    torch.manual_seed(7); np.random.seed(7)
    pol = MLP(spec, hidden_sizes=hidden, seed=123)
    base = MLPBaseline(spec, epochs=1)
    agent = NPG(None, pol, base, normalized_step_size=STEP, FIM_invert_args={"iters": CG_ITERS, "damping": DAMPING})
    batch = PathBatch(*host, device=local)
    dev = batch.device
    ps.returns_on_device(batch, GAMMA)
    base.fit_batch(batch)
    st0 = {k: v.clone() for k, v in base._state(dev).items()}
    step0 = base.adam_step

    def one():
        pol.set_param_values(flat0.copy(), True, True)
        for k, v in base._state(dev).items():
            v.copy_(st0[k])
        base.adam_step = step0
        batch.returns = batch.baseline = batch.adv = None
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        agent.train_step(N, gamma=GAMMA, gae_lambda=LAM, paths=batch)
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b)

    one()
    ts = sorted(one() for _ in range(3))
    ms = ts[1]
    return {"value": N / (ms * 1e-3), "unit": "timesteps/s", "ms_per_step": ms, "fit_steps": base.last_fit_steps,
            "workload": "NPG + MLP(64,64) + MLPBaseline(epochs=1, batch 64) on the same batch: BASELINE.json configs[2] "
                        "verbatim; the fit is %d SERIAL Adam steps" % base.last_fit_steps}


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
