"""GPU parity: MLPBaseline (csrc/mlpbase.cu) vs the live-reference goldens and the oracle.

Tolerances (fp32 network): features bit-exact; predictions 1e-5 relative; one Adam
step and the full 44-step chain compared weight-for-weight (2e-4 relative to the parameter
scale: Adam's update is ~lr * sign(g) for small |g|, so a gradient entry whose sign flips
inside fp32 round-off moves one weight by 2*lr); error ratios 1e-4 relative."""
import copy
import os
import pickle

import numpy as np
import pytest
import torch

from conftest import GOLDEN_DIR, rel_l2

pytestmark = pytest.mark.gpu
CASES = ["mlpbase_1step", "mlpbase_hopper"]


def load(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    return {k: z[k] for k in z.files}


def paths_of(g):
    off = g["path_offsets"]
    return [dict(observations=g["observations"][off[i]:off[i + 1]],
                 rewards=g["rewards"][off[i]:off[i + 1]],
                 returns=g["returns"][off[i]:off[i + 1]],
                 terminated=bool(g["terminated"][i])) for i in range(len(off) - 1)]


def make(g):
    from mjrl_b200.baselines.mlp_baseline import MLPBaseline
    from mjrl_b200.utils.gym_env import EnvSpec
    torch.manual_seed(int(g["seed"]))
    return MLPBaseline(EnvSpec(int(g["n"]), 3, int(g["horizon"]), 1), reg_coef=float(g["reg_coef"]),
                       batch_size=int(g["batch_size"]), epochs=int(g["epochs"]))


@pytest.mark.parametrize("name", CASES)
def test_init_features_predict(name):
    g = load(name)
    base = make(g)
    np.testing.assert_array_equal(base.get_param_values(), g["init_params"])   # seed-for-seed init
    paths = paths_of(g)
    from mjrl_b200.batch import batch_for
    from mjrl_b200._lib import c_i64, c_int, ptr, stream_ptr
    b = batch_for(paths)
    feat = torch.empty((b.N, b.n + 4), dtype=torch.float32, device=b.device)
    b.ctx.call("mjrl_mlpbase_features", ptr(b.obs), ptr(None), ptr(b.offsets), c_int(b.P), c_i64(b.N),
               c_int(b.n), ptr(feat), ptr(None), stream_ptr())
    np.testing.assert_array_equal(feat.cpu().numpy(), g["features"])
    pred = np.concatenate([base.predict(p) for p in paths])
    assert pred.dtype == np.float32
    scale = np.abs(g["pred_before"]).max()
    assert np.abs(pred - g["pred_before"]).max() < 1e-5 * max(scale, 1.0)
    pb = base.predict_batch(b).cpu().numpy()
    np.testing.assert_array_equal(pb.astype(np.float32), pred)                # batched == per path


@pytest.mark.parametrize("name", CASES)
def test_fit_chain_matches_reference(name):
    g = load(name)
    base = make(g)
    paths = paths_of(g)
    np.random.seed(int(g["seed"]) + 1)
    eb, ea = base.fit(paths, return_errors=True)
    steps = int(g["epochs"]) * (int(int(g["N"]) / 64) - 1)
    assert base.last_fit_steps == steps and base.adam_step == steps
    p = base.get_param_values()
    scale = np.abs(g["fit_params"]).max()
    d = np.abs(p - g["fit_params"]).max()
    moved = np.abs(g["fit_params"] - g["init_params"]).max()
    print("%s: %d steps, max|dp|=%.3e (params moved %.3e), rel_l2=%.3e" %
          (name, steps, d, moved, rel_l2(p, g["fit_params"])))
    assert d < 2e-4 * scale
    assert rel_l2(p - g["init_params"], g["fit_params"] - g["init_params"]) < 2e-2
    assert abs(eb - g["errors"][0]) < 1e-4 * g["errors"][0]
    assert abs(ea - g["errors"][1]) < 1e-3 * g["errors"][1]
    pred = np.concatenate([base.predict(q) for q in paths])
    assert np.abs(pred - g["pred_after"]).max() < 1e-3 * max(np.abs(g["pred_after"]).max(), 1.0)
    # Adam moments and the step count persist across fit() calls and through pickling
    base2 = pickle.loads(pickle.dumps(base))
    assert base2._dev is None and base2.adam_step == steps
    np.testing.assert_array_equal(base2.get_param_values(), p)
    base2.fit(paths)
    d2 = np.abs(base2.get_param_values() - g["fit2_params"]).max()
    print("   second fit (continued chain): max|dp|=%.3e" % d2)
    assert d2 < 5e-4 * scale


def test_single_step_is_tight():
    """One Adam step from identical weights on the reference's minibatch: the step is
    lr*m_hat/(sqrt(v_hat)+eps), so each weight must land within fp32 round-off."""
    g = load("mlpbase_1step")
    base = make(g)
    np.random.seed(int(g["seed"]) + 1)
    base.fit(paths_of(g))
    p = base.get_param_values()
    step = g["fit_params"] - g["init_params"]
    mine = p - g["init_params"]
    big = np.abs(step) > 0.5e-3          # entries whose gradient is well away from zero
    assert big.sum() > 1000
    assert np.abs(mine[big] - step[big]).max() < 2e-6
    assert (np.sign(mine[big]) == np.sign(step[big])).all()


def test_small_batch_and_explicit_permutation_vs_oracle():
    """batch_size < 64 (masked rows), weight decay, explicit permutation: GPU vs the oracle."""
    from oracle import mjrl_oracle as O
    from mjrl_b200.baselines.mlp_baseline import MLPBaseline
    from mjrl_b200.utils.gym_env import EnvSpec
    from mjrl_b200.batch import batch_for, drop_cache
    g = load("mlpbase_hopper")
    paths = paths_of(g)
    N = int(g["N"])
    rng = np.random.default_rng(5)
    perm = rng.permutation(N)[:24 * 10]
    torch.manual_seed(3)
    base = MLPBaseline(EnvSpec(int(g["n"]), 3, 200, 1), reg_coef=1e-2, batch_size=24, epochs=1)
    torch.manual_seed(3)

    class Spec:
        observation_dim = int(g["n"])
    ob = O.MLPBaseline(Spec(), reg_coef=1e-2, batch_size=24, epochs=1,
                       perm_source=lambda n: np.concatenate([perm, np.zeros(24 * 3, dtype=np.int64)]))
    drop_cache()
    base.fit_batch(batch_for(paths), perm=perm)
    # the oracle runs int(N/bs)-1 steps of its permutation; give it exactly 10 by truncating N
    X = torch.from_numpy(O.mlp_baseline_features(paths).astype("float32"))
    y = torch.from_numpy(g["returns"].reshape(-1, 1).astype("float32"))
    for mb in range(10):
        idx = torch.from_numpy(perm[mb * 24:(mb + 1) * 24].astype(np.int64))
        ob.optimizer.zero_grad()
        torch.nn.functional.mse_loss(ob.model(X[idx]), y[idx]).backward()
        ob.optimizer.step()
    ref = np.concatenate([q.data.numpy().ravel() for q in ob.model.parameters()])
    p = base.get_param_values()
    assert np.abs(p - ref).max() < 2e-4 * np.abs(ref).max()
