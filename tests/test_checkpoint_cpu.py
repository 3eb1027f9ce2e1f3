"""Checkpoint interchange with stock mjrl (mjrl_b200/utils/checkpoint.py).

tests/golden/stock_*.pickle were written by `pickle.dump` of the LIVE reference's objects exactly like
train_agent.py:61-66 (generator: make_golden.py --checkpoint-only).  Loading them needs neither mjrl
nor a GPU.  The reverse direction is checked structurally everywhere and against the live reference
where /root/reference exists (the build container)."""
import os
import pickle
import pickletools
import subprocess
import sys

import numpy as np
import pytest

from conftest import GOLDEN_DIR
from mjrl_b200.utils import checkpoint as ck

G = lambda f: os.path.join(GOLDEN_DIR, f)
REF = "/root/reference"


@pytest.fixture(scope="module")
def gold():
    z = np.load(G("stock_checkpoints.npz"))
    return {k: z[k] for k in z.files}


def test_load_stock_mlp_policy(gold):
    from mjrl_b200.policies.gaussian_mlp import MLP, MuNet
    pol = ck.load(G("stock_policy_mlp.pickle"))
    assert type(pol) is MLP and type(pol.model) is MuNet and type(pol.old_model) is MuNet
    np.testing.assert_array_equal(pol.get_param_values(), gold["mlp_params"])
    np.testing.assert_array_equal(pol.get_old_param_values(), gold["mlp_params"])
    assert pol.hidden_sizes == (64, 64) and pol._dev is None
    # written after set_param_values(float32, True, True), i.e. from an ALIASED policy (F14) -- but every
    # from_numpy view has its own Storage object, so pickling un-shares them: a loaded policy is not aliased
    assert pol._old_aliases_new is False
    assert pol.trainable_params[0].data_ptr() != pol.old_params[0].data_ptr()
    assert "mjrl_b200" not in sys.modules.get("mjrl", type(sys)("x")).__dict__
    # host forward of the loaded network == the reference's means
    import torch
    mean = pol.model(torch.from_numpy(gold["mlp_obs"]).float()).detach().numpy()
    np.testing.assert_allclose(mean, gold["mlp_mean"], rtol=0, atol=1e-6)
    # a deepcopy / pickle of the loaded object is again one of ours
    again = pickle.loads(pickle.dumps(pol))
    np.testing.assert_array_equal(again.get_param_values(), gold["mlp_params"])


def test_load_stock_linear_policy_not_aliased(gold):
    from mjrl_b200.policies.gaussian_linear import LinearPolicy
    pol = ck.load(G("stock_policy_linear.pickle"))
    assert type(pol) is LinearPolicy and pol._old_aliases_new is False
    np.testing.assert_array_equal(pol.get_param_values(), gold["lin_params"])


def test_load_stock_baselines(gold):
    from mjrl_b200.baselines.linear_baseline import LinearBaseline
    from mjrl_b200.baselines.mlp_baseline import MLPBaseline
    lb = ck.load(G("stock_baseline_linear.pickle"))
    assert type(lb) is LinearBaseline and lb.n == 11
    np.testing.assert_array_equal(np.asarray(lb._coeffs), gold["lb_coeffs"])
    mb = ck.load(G("stock_baseline_mlp.pickle"))
    assert type(mb) is MLPBaseline and not hasattr(mb, "optimizer")
    np.testing.assert_array_equal(mb.get_param_values(), gold["mb_params"])
    assert mb.adam_step == 1500 // 64 - 1 and mb.reg_coef == 1e-3 and mb.learn_rate == 1e-3
    assert mb._host_m.shape == gold["mb_params"].shape and np.abs(mb._host_v).max() > 0


def _globals_of(data):
    out, stack = set(), []
    for op, arg, _ in pickletools.genops(data):
        if op.name == "GLOBAL":
            out.add(arg.replace(" ", "."))
        elif op.name in ("SHORT_BINUNICODE", "BINUNICODE", "UNICODE"):
            stack.append(arg)
        elif op.name == "STACK_GLOBAL":
            out.add(stack[-2] + "." + stack[-1])
    return out


@pytest.mark.parametrize("fname", ["stock_policy_mlp", "stock_policy_linear", "stock_baseline_linear",
                                   "stock_baseline_mlp"])
def test_dump_writes_stock_class_paths(fname):
    obj = ck.load(G(fname + ".pickle"))
    before = set(k for k in sys.modules if k == "mjrl" or k.startswith("mjrl."))
    data = ck.dumps(obj)
    assert set(k for k in sys.modules if k == "mjrl" or k.startswith("mjrl.")) == before   # stand-ins removed
    ours = _globals_of(data)
    stock = _globals_of(open(G(fname + ".pickle"), "rb").read())
    assert not any(g.startswith("mjrl_b200") for g in ours), ours
    assert {g for g in ours if g.startswith("mjrl.")} == {g for g in stock if g.startswith("mjrl.")}
    # and the round trip through our own loader is lossless
    back = ck.loads(data)
    if hasattr(obj, "get_param_values"):
        np.testing.assert_array_equal(back.get_param_values(), obj.get_param_values())
    else:
        np.testing.assert_array_equal(np.asarray(back._coeffs), np.asarray(obj._coeffs))
    if fname == "stock_policy_mlp":
        assert back._old_aliases_new is False
    if fname == "stock_baseline_mlp":
        np.testing.assert_array_equal(back._host_m, obj._host_m)
        assert back.adam_step == obj.adam_step


_STOCK_SIDE = r"""
import sys, types, pickle, numpy as np
for name in ("gym", "matplotlib", "matplotlib.pyplot"):
    sys.modules.setdefault(name, types.ModuleType(name))
sys.modules["matplotlib"].use = lambda *a, **k: None
sys.path.insert(0, "/root/reference")
import torch
z = np.load(sys.argv[1]); d = sys.argv[2]
pol = pickle.load(open(d + "/p.pickle", "rb"))
import mjrl.policies.gaussian_mlp as gm, mjrl.baselines.mlp_baseline as mb, mjrl.baselines.linear_baseline as lb
assert type(pol) is gm.MLP and type(pol.model) is gm.MuNet, type(pol)
assert np.array_equal(pol.get_param_values(), z["mlp_params"])
assert np.array_equal(pol.log_likelihood(z["mlp_obs"], z["mlp_act"]), z["mlp_ll"])   # stock code, our pickle
a = pol.get_action(z["mlp_obs"][0])
pol.set_param_values(pol.get_param_values() * 1.0, set_new=True, set_old=True)
l = pickle.load(open(d + "/l.pickle", "rb"))
assert type(l) is lb.LinearBaseline and np.array_equal(l._coeffs, z["lb_coeffs"])
m = pickle.load(open(d + "/m.pickle", "rb"))
assert type(m) is mb.MLPBaseline and isinstance(m.optimizer, torch.optim.Adam)
paths = []
off = z["b_path_offsets"]
for i in range(len(off) - 1):
    s = slice(off[i], off[i + 1])
    paths.append(dict(observations=z["b_observations"][s], rewards=z["b_rewards"][s], returns=z["returns"][s]))
pred = np.concatenate([m.predict(p) for p in paths])
assert np.array_equal(pred, z["mb_pred"])
np.random.seed(13)
m.fit(paths)        # continues the Adam chain from the moments we wrote
flat = np.concatenate([p.data.numpy().ravel() for p in m.model.parameters()]).astype(np.float32)
assert np.array_equal(flat, z["mb_params_2"]), np.abs(flat - z["mb_params_2"]).max()
print("STOCK-OK")
"""


@pytest.mark.skipif(not os.path.isdir(REF), reason="live reference only exists in the build container")
def test_stock_mjrl_loads_our_dumps(tmp_path):
    ck.dump(ck.load(G("stock_policy_mlp.pickle")), str(tmp_path / "p.pickle"))
    ck.dump(ck.load(G("stock_baseline_linear.pickle")), str(tmp_path / "l.pickle"))
    ck.dump(ck.load(G("stock_baseline_mlp.pickle")), str(tmp_path / "m.pickle"))
    env = {k: v for k, v in os.environ.items() if k != "PYTHONPATH"}
    r = subprocess.run([sys.executable, "-c", _STOCK_SIDE, G("stock_checkpoints.npz"), str(tmp_path)],
                       capture_output=True, text=True, cwd=str(tmp_path), env=env, timeout=300)
    assert "STOCK-OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def test_dump_load_roundtrip_keeps_hidden_sizes():
    """A mirror-built MLP with non-default widths survives dumps -> loads: stock MuNet stores `hidden_sizes`
    (gaussian_mlp.py:153), stock MLP does not; the loader recovers the policy-level field from the net."""
    from mjrl_b200.policies.gaussian_mlp import MLP
    from mjrl_b200.utils.gym_env import EnvSpec
    pol = MLP(EnvSpec(7, 2, 50, 1), hidden_sizes=(32, 16), seed=3)
    back = ck.loads(ck.dumps(pol))
    assert tuple(back.hidden_sizes) == (32, 16)
    assert tuple(back.model.hidden_sizes) == (32, 16) and tuple(back.old_model.hidden_sizes) == (32, 16)
    np.testing.assert_array_equal(back.get_param_values(), pol.get_param_values())
    assert back.d == pol.d
    # a stock-side pickle without the policy-level field and without MuNet.hidden_sizes still resolves
    del back.__dict__["hidden_sizes"]; del back.model.__dict__["hidden_sizes"]
    assert tuple(ck._adopt_policy(back).hidden_sizes) == (32, 16)
