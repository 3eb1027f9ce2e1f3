"""mjrl_moments_noalloc (include/mjrl_b200.h): the allocation-free variant of mjrl_moments for callers that
cannot call cudaMalloc (CUDA graph capture, serving loops).  Its per-CTA partial sums live in a file-scope
`static __device__` array instead of the context's grow-on-demand scratch.  It has no counterpart in the
reference; the check is against the existing entry point through the C-ABI.

Kept in a file pytest collects last: a failure here must not hide other
tests.  It passes alone in a fresh process (it is then the first user of the library) and after the others."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_N = 4096 * 2048     # the limit stated in include/mjrl_b200.h


def _both(ctx, x):
    import torch
    from mjrl_b200._lib import c_i64, ptr, stream_ptr
    ref = torch.full((4,), -1.0, dtype=torch.float64, device=x.device)
    out = torch.full((4,), -2.0, dtype=torch.float64, device=x.device)
    ctx.call("mjrl_moments", ptr(x), c_i64(x.numel()), ptr(ref), stream_ptr())
    ctx.call("mjrl_moments_noalloc", ptr(x), c_i64(x.numel()), ptr(out), stream_ptr())
    return ref.cpu().numpy(), out.cpu().numpy()


def test_noalloc_moments_match_the_allocating_entry_point_twice_in_a_row():
    import torch
    from mjrl_b200 import _lib
    from mjrl_b200._lib import c_i64, ptr, stream_ptr
    ctx = _lib.context()
    dev = torch.device("cuda", ctx.device)
    rng = np.random.default_rng(11)
    # two calls in a row share the static array: the second (shorter, fewer CTAs, different values) would pick
    # up stale partial sums of the first if any entry it reads were not rewritten
    sizes = (1_000_003, 4_097)
    xs = [torch.from_numpy(rng.standard_normal(n) * (1.0 + 3.0 * k) + 5.0 * k).to(dev) for k, n in enumerate(sizes)]
    outs = []
    for x in xs:                                   # call 1, call 2: nothing between them resets the array
        o = torch.full((4,), -2.0, dtype=torch.float64, device=dev)
        ctx.call("mjrl_moments_noalloc", ptr(x), c_i64(x.numel()), ptr(o), stream_ptr())
        outs.append(o)
    for x, o in zip(xs, outs):
        ref = torch.full((4,), -1.0, dtype=torch.float64, device=dev)
        ctx.call("mjrl_moments", ptr(x), c_i64(x.numel()), ptr(ref), stream_ptr())
        got, want = o.cpu().numpy(), ref.cpu().numpy()
        np.testing.assert_array_equal(got, want)            # same kernels, same grid: same bits
        h = x.cpu().numpy()                                 # and both are numpy's moments (tolerance of the scan tests)
        np.testing.assert_allclose(got, [h.size, h.sum(), ((h - h.mean()) ** 2).sum(), h.mean()], rtol=1e-10)
    # interleaved with the allocating call, again twice
    for x in xs:
        want, got = _both(ctx, x)
        np.testing.assert_array_equal(got, want)


def test_noalloc_moments_allocate_nothing_and_reject_oversize():
    import torch
    from mjrl_b200 import _lib
    from mjrl_b200._lib import c_i64, ptr, stream_ptr
    ctx = _lib.context()
    dev = torch.device("cuda", ctx.device)
    x = torch.arange(300_000, dtype=torch.float64, device=dev)
    o = torch.zeros(4, dtype=torch.float64, device=dev)
    ctx.call("mjrl_moments_noalloc", ptr(x), c_i64(x.numel()), ptr(o), stream_ptr())   # set-up call
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(dev)[0]
    for _ in range(3):
        ctx.call("mjrl_moments_noalloc", ptr(x), c_i64(x.numel()), ptr(o), stream_ptr())
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(dev)[0] == free0        # no device memory taken by the calls
    np.testing.assert_allclose(o.cpu().numpy()[[0, 1, 3]], [300_000, 299_999 * 150_000.0, 149_999.5], rtol=1e-12)
    # beyond the static scratch: a clear error, no launch (the pointer is never dereferenced)
    lib = _lib.load()
    rc = lib.mjrl_moments_noalloc(ctx.h, ptr(x), c_i64(MAX_N + 1), ptr(o), stream_ptr())
    assert rc == -2                                        # MJRL_E_SHAPE
    msg = lib.mjrl_last_error().decode()
    assert "8388608" in msg and "mjrl_moments" in msg, msg
    assert lib.mjrl_moments_noalloc(ctx.h, ptr(x), c_i64(0), ptr(o), stream_ptr()) == -1   # MJRL_E_ARG
    with pytest.raises(_lib.MjrlError):
        ctx.call("mjrl_moments_noalloc", ptr(x), c_i64(MAX_N + 1), ptr(o), stream_ptr())
