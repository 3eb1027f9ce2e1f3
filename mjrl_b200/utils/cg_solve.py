"""Drop-in for mjrl/utils/cg_solve.py:3-22 -- conjugate gradient around a caller-supplied operator.

`NPG.train_from_paths` does not come through here: it runs the whole solve on the device through one
C-ABI call (`mjrl_cg_solve`, algos/engine.py).  This function keeps the reference's public signature for
callers that bring their own `f_Ax` (e.g. `agent.build_Hvp_eval(...)`, npg_cg.py:76-81,122-125): the operator
is a host callable numpy [d] -> numpy [d], so the loop runs where the operator's results arrive.

Reference behaviour kept on purpose:
  * `x_0` is ignored: the solve starts from x = 0 with r = p = b (cg_solve.py:4-7 overwrite x before use);
  * the dtype of every vector follows `b` (fp32 in the reference's call, npg_cg.py:124);
  * the residual test happens AFTER the update of an iteration (`:19-20`), so at least one operator call
    is made, and `cg_iters` bounds them.
"""
import numpy as np


def cg_solve(f_Ax, b, x_0=None, cg_iters=10, residual_tol=1e-10):
    """Returns x ~ A^-1 b after at most `cg_iters` applications of `f_Ax` (arithmetic in b's dtype)."""
    b = np.asarray(b)
    solution = np.zeros_like(b)                 # :4  -- x_0 is not used
    resid = np.array(b, copy=True)              # :5  -- b - A 0
    direction = np.array(resid, copy=True)      # :6
    rs = np.dot(resid, resid)                   # :7
    done = 0
    while done < cg_iters:                      # :9
        Ad = f_Ax(direction)                    # :10
        step = rs / np.dot(direction, Ad)       # :11
        solution += step * direction            # :12
        resid -= step * Ad                      # :13
        rs_next = np.dot(resid, resid)          # :14
        direction = resid + (rs_next / rs) * direction      # :15-16
        rs = rs_next                            # :18
        done += 1
        if rs < residual_tol:                   # :19-20, tested after the update
            break
    return solution
