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
