"""CPU-side checks of the drop-in boundary: the C-ABI library builds, loads and exports every
entry point that include/mjrl_b200.h declares (no compute calls -- there is no GPU here)."""
import os
import subprocess

import pytest

from mjrl_b200 import _lib, build


def test_library_builds_and_exports_every_declared_symbol():
    lib_path = build.build()
    assert os.path.exists(lib_path)
    out = subprocess.run(["nm", "-D", "--defined-only", lib_path], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    declared = _lib.declared_symbols()
    assert len(declared) >= 15
    missing = [s for s in declared if s not in exported]
    assert not missing, "declared in include/mjrl_b200.h but not exported: %s" % missing
    lib = _lib.load()
    assert lib.mjrl_abi_version() == 1


def test_library_is_sm100a_only():
    out = subprocess.run(["cuobjdump", "--list-elf", build.LIB], capture_output=True, text=True).stdout
    archs = {tok for l in out.splitlines() for tok in l.replace(".", " ").split() if tok.startswith("sm_")}
    assert archs == {"sm_100a"}, archs


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(_lib.MjrlError):
        _lib.context()


def test_product_package_never_imports_oracle():
    root = os.path.dirname(os.path.abspath(build.__file__))
    for dp, _, files in os.walk(root):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dp, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, os.path.join(dp, f)


def test_policy_init_matches_reference_rng_order(golden):
    """Seed-for-seed identical initial parameters (gaussian_mlp.py:27-57); host only."""
    import numpy as np
    from mjrl_b200.policies.gaussian_mlp import MLP
    from mjrl_b200.policies.gaussian_linear import LinearPolicy
    from mjrl_b200.utils.gym_env import EnvSpec
    g = golden
    spec = EnvSpec(g["n"], g["m"], g["horizon"], 1)
    pol = MLP(spec, hidden_sizes=g["hidden"], seed=123) if g["hidden"] else LinearPolicy(spec, seed=123)
    np.testing.assert_array_equal(pol.get_param_values(), g["init_params"])
    assert pol.d == g["init_params"].size
    # clamp semantics of set_param_values (gaussian_mlp.py:75-76) and pickling without handles
    import pickle, copy
    p = pol.get_param_values(); p[-1] = -7.0
    pol.set_param_values(p)
    assert pol.get_param_values()[-1] == -3.0 and pol.get_old_param_values()[-1] == -3.0
    q = pickle.loads(pickle.dumps(pol)); c = copy.deepcopy(pol)
    np.testing.assert_array_equal(q.get_param_values(), pol.get_param_values())
    np.testing.assert_array_equal(c.get_param_values(), pol.get_param_values())
