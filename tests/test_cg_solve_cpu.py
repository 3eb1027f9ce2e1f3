"""utils.cg_solve keeps the reference's signature and arithmetic (mjrl/utils/cg_solve.py:3-22)."""
import inspect
import os

import numpy as np
import pytest

from mjrl_b200.utils.cg_solve import cg_solve
from oracle import mjrl_oracle as O


def _spd(d, seed):
    rng = np.random.default_rng(seed)
    A = rng.standard_normal((d, d)).astype(np.float32)
    return A @ A.T + np.eye(d, dtype=np.float32), rng.standard_normal(d).astype(np.float32)


def test_signature_is_the_references():
    sig = inspect.signature(cg_solve)
    assert list(sig.parameters) == ["f_Ax", "b", "x_0", "cg_iters", "residual_tol"]
    assert sig.parameters["cg_iters"].default == 10 and sig.parameters["residual_tol"].default == 1e-10
    assert sig.parameters["x_0"].default is None


@pytest.mark.parametrize("d,iters", [(4, 1), (16, 3), (200, 10), (16, 100)])
def test_bit_equal_to_the_oracle_and_ignores_x0(d, iters):
    A, b = _spd(d, d + iters)
    op = lambda v: A @ v
    x = cg_solve(op, b, x_0=123.0 * b, cg_iters=iters)
    assert x.dtype == np.float32
    np.testing.assert_array_equal(x, O.cg_solve(op, b, x_0=b.copy(), cg_iters=iters))
    np.testing.assert_array_equal(x, cg_solve(op, b, cg_iters=iters))
    assert not np.shares_memory(x, b)


def test_stops_after_the_update_that_meets_the_tolerance():
    A, b = _spd(8, 1)
    calls = []
    def op(v):
        calls.append(1)
        return A @ v
    cg_solve(op, b, cg_iters=10, residual_tol=1e30)
    assert len(calls) == 1              # the test comes after the update: one operator call, never zero
    calls.clear()
    x = cg_solve(op, b.astype(np.float64), cg_iters=50, residual_tol=1e-20)
    assert len(calls) <= 9 and np.allclose(A.astype(np.float64) @ x, b, atol=1e-6)


@pytest.mark.skipif(not os.path.isdir("/root/reference/mjrl"), reason="live reference only in the build container")
def test_bit_equal_to_the_live_reference():
    import importlib.util
    spec = importlib.util.spec_from_file_location("_ref_cg", "/root/reference/mjrl/utils/cg_solve.py")
    ref = importlib.util.module_from_spec(spec); spec.loader.exec_module(ref)
    for d, iters in ((5, 2), (64, 10), (300, 25)):
        A, b = _spd(d, 7 * d)
        np.testing.assert_array_equal(cg_solve(lambda v: A @ v, b, x_0=b.copy(), cg_iters=iters),
                                      ref.cg_solve(lambda v: A @ v, b, x_0=b.copy(), cg_iters=iters))
