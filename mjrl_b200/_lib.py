"""ctypes binding of the C-ABI library (include/mjrl_b200.h).

There is no CPU fallback: if the library is missing or no sm_100a device is present the
first call raises.  Loading the library (symbol table only) works without a GPU so that the
CPU test-suite can verify that every declared entry point is exported.
"""
import ctypes
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libmjrl_b200.so")
HEADER = os.path.join(HERE, "..", "include", "mjrl_b200.h")

_lib = None


class MjrlError(RuntimeError):
    pass


def declared_symbols():
    """Names of every `int|const char* mjrl_*(` entry point declared in the header."""
    txt = open(HEADER).read()
    return sorted(set(re.findall(r"\b(mjrl_[a-z0-9_]+)\s*\(", txt)))


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise MjrlError(
            "mjrl_b200: %s is missing -- run `python -m mjrl_b200.build` (no CPU fallback exists)"
            % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    lib.mjrl_last_error.restype = ctypes.c_char_p
    for name in declared_symbols():
        fn = getattr(lib, name)  # raises AttributeError if a declared symbol is not exported
        if name.endswith("_count"):   # size queries return int64, not a status
            fn.restype = ctypes.c_int64
        elif name in ("mjrl_peer_slot", "mjrl_peer_base"):   # device pointers
            fn.restype = ctypes.c_void_p
        elif name != "mjrl_last_error":
            fn.restype = ctypes.c_int
    _lib = lib
    return lib


def check(rc, what=""):
    if rc != 0:
        msg = load().mjrl_last_error().decode(errors="replace")
        raise MjrlError("mjrl_b200 %s failed (rc=%d): %s" % (what, rc, msg))


# ---- argument marshalling helpers ------------------------------------------------------
c_void_p, c_int, c_i64, c_double = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64, ctypes.c_double
c_float = ctypes.c_float


def ptr(t):
    """Device (or host) pointer of a torch tensor / None as c_void_p."""
    if t is None:
        return c_void_p(0)
    if isinstance(t, int):
        return c_void_p(t)
    if isinstance(t, ctypes.c_void_p):
        return t
    return c_void_p(t.data_ptr())


class Context:
    """One per process / GPU (mjrl_ctx_create)."""

    def __init__(self, device=0):
        self.lib = load()
        h = c_void_p()
        check(self.lib.mjrl_ctx_create(ctypes.byref(h), c_int(int(device))), "ctx_create")
        self.h = h
        self.device = int(device)
        self.sm_count = self.lib.mjrl_ctx_sm_count(self.h)
        self.launches = 0   # kernels launched through this context (bench.py's gpu_launches)

    def close(self):
        if getattr(self, "h", None):
            self.lib.mjrl_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def call(self, name, *args):
        check(getattr(self.lib, name)(self.h, *args), name)
        self.launches += KERNELS_PER_CALL.get(name, 1)


# kernels each entry point launches (csrc/*.cu), for the launch count bench.py reports
KERNELS_PER_CALL = {"mjrl_moments": 4, "mjrl_moments_noalloc": 4, "mjrl_path_return_stats": 2, "mjrl_linbase_gram": 2,
                    "mjrl_quadbase_gram": 2,
                    "mjrl_policy_eval": 2, "mjrl_policy_eval_tc": 2, "mjrl_policy_grad": 3, "mjrl_policy_fvp": 2,
                    "mjrl_policy_fvp_cached": 2, "mjrl_policy_grad_cached": 3, "mjrl_policy_grad_cached_tc": 3,
                    "mjrl_policy_act": 2, "mjrl_rollout_act": 2, "mjrl_mlpbase_features": 2, "mjrl_mlpbase_sqerr": 2}

_ctx = {}


def context(device=None):
    """Process-wide context for `device` (default: torch's current CUDA device)."""
    import torch
    if not torch.cuda.is_available():
        raise MjrlError("mjrl_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    if device is None:
        device = torch.cuda.current_device()
    device = int(device)
    if device not in _ctx:
        _ctx[device] = Context(device)
    return _ctx[device]


# bytes brought back from the device by the update path (bench.py reports the per-step count)
D2H_BYTES = [0]


def to_host(t):
    """Device tensor -> numpy (synchronises the stream), counted in D2H_BYTES."""
    D2H_BYTES[0] += int(t.numel() * t.element_size())
    return t.cpu().numpy()


_raw_stream = None


def stream_ptr():
    """cudaStream_t of torch's current stream on the current device.  torch.cuda.current_stream() builds a Stream
    object and resolves the device index in Python (15 us per call under a profiler, a dozen calls per update: a fifth
    of the host time of a small update); the raw getter is the same lookup without the object."""
    global _raw_stream
    import torch
    if _raw_stream is None:
        raw = getattr(torch._C, "_cuda_getCurrentRawStream", None)
        _raw_stream = raw if raw is not None else (lambda dev: torch.cuda.current_stream(dev).cuda_stream)
    return c_void_p(_raw_stream(torch.cuda.current_device()))


# ---- C structs of include/mjrl_b200.h ---------------------------------------------------
class PolicyDesc(ctypes.Structure):
    _fields_ = [("n", ctypes.c_int32), ("m", ctypes.c_int32), ("h1", ctypes.c_int32),
                ("h2", ctypes.c_int32), ("in_shift", c_void_p), ("in_scale", c_void_p),
                ("out_shift", c_void_p), ("out_scale", c_void_p)]


class CGState(ctypes.Structure):
    _fields_ = [("x", c_void_p), ("r", c_void_p), ("p", c_void_p), ("z", c_void_p),
                ("scal", c_void_p), ("p32", c_void_p), ("flags", c_void_p)]


c_u64 = ctypes.c_uint64
