"""Device bridge of a Gaussian policy: builds the `mjrl_policy_desc` of include/mjrl_b200.h
from the host object and wraps the policy entry points of the C-ABI."""
import ctypes

import numpy as np
import torch

from .. import _lib
from .._lib import PolicyDesc, c_i64, c_u64, c_void_p, ptr, stream_ptr


def _f32_dev(x, device):
    return torch.from_numpy(np.ascontiguousarray(np.asarray(x, dtype=np.float32))).to(device)


class PolicyOnDevice:
    def __init__(self, policy, device=None):
        self.ctx = _lib.context(device)
        self.device = torch.device("cuda", self.ctx.device)
        self.policy = policy
        hs = tuple(getattr(policy, "hidden_sizes", ()) or ())
        self.n, self.m = int(policy.n), int(policy.m)
        self.h1, self.h2 = (int(hs[0]), int(hs[1])) if hs else (0, 0)
        self._tr_key = None
        self._tr = None
        self.desc = None
        self.refresh_transforms()
        d = self.ctx.lib.mjrl_policy_param_count(ctypes.byref(self.desc))
        if d != int(policy.d):
            raise _lib.MjrlError("parameter count mismatch: kernel layout %d vs policy.d %d" % (d, policy.d))
        self.d = int(d)

    # in/out transforms can be changed at any time through model.set_transformations
    def refresh_transforms(self):
        mdl = self.policy.model
        vals = [mdl.in_shift, mdl.in_scale, mdl.out_shift, mdl.out_scale]
        # identity + in-place version of the four tensors first (no host work when nothing was touched: this runs before
        # every C call that takes the policy description); their bytes only when one of them changed
        ident = tuple((id(v), v._version) for v in vals)
        if ident == getattr(self, "_tr_ident", None):
            return self.desc
        self._tr_ident = ident
        self._tr_vals = vals          # keep the tensors alive: a freed one's id could be reused
        key = tuple(v.detach().numpy().tobytes() for v in vals)
        if key != self._tr_key:
            self._tr = [v.detach().float().contiguous().to(self.device) for v in vals]
            self._tr_key = key
            self.desc = PolicyDesc(self.n, self.m, self.h1, self.h2,
                                   *[c_void_p(t.data_ptr()) for t in self._tr])
        return self.desc

    def desc_ref(self):
        return ctypes.byref(self.refresh_transforms())

    def params(self, which="new"):
        # start of an update / a public call: compare the transforms byte for byte once (catches a tensor edited through a
        # numpy view, which bumps no version counter); the C calls that follow take the identity fast path
        self._tr_ident = None
        p = self.policy
        flat = p.get_param_values() if which == "new" else p.get_old_param_values()
        return _f32_dev(flat, self.device)

    # ---- C-ABI wrappers -------------------------------------------------------------------
    def forward(self, params, obs, act, want_mean=True, want_ll=True):
        N = obs.shape[0]
        mean = torch.empty((N, self.m), dtype=torch.float32, device=self.device) if want_mean else None
        ll = torch.empty(N, dtype=torch.float32, device=self.device) if want_ll else None
        self.ctx.call("mjrl_policy_forward", self.desc_ref(), ptr(params), ptr(obs), ptr(act),
                      c_i64(N), ptr(mean), ptr(ll), stream_ptr())
        return mean, ll

    def mean(self, observations):
        obs = _f32_dev(observations, self.device).reshape(-1, self.n)
        mean, _ = self.forward(self.params("new"), obs, None, True, False)
        return mean.cpu().numpy()

    def mean_ll(self, observations, actions, which="new"):
        obs = _f32_dev(observations, self.device).reshape(-1, self.n)
        act = _f32_dev(actions, self.device).reshape(-1, self.m)
        mean, ll = self.forward(self.params(which), obs, act)
        return mean.cpu().numpy(), ll.cpu().numpy()

    def act(self, observations, seed=0, offset=0, params=None):
        obs = observations if torch.is_tensor(observations) else _f32_dev(observations, self.device)
        obs = obs.reshape(-1, self.n)
        B = obs.shape[0]
        params = self.params("new") if params is None else params
        act = torch.empty((B, self.m), dtype=torch.float32, device=self.device)
        mean = torch.empty((B, self.m), dtype=torch.float32, device=self.device)
        self.ctx.call("mjrl_policy_act", self.desc_ref(), ptr(params), ptr(obs), c_i64(B),
                      c_u64(int(seed)), c_u64(int(offset)), ptr(act), ptr(mean), stream_ptr())
        return act, mean
