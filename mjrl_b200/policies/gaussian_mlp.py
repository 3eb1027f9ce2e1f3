"""Drop-in for mjrl/policies/gaussian_mlp.py (class MLP + MuNet) on the CUDA path.

State lives where the reference keeps it -- host torch tensors in `model`, `old_model`,
`log_std`, `old_log_std` -- so pickling, deepcopy, behaviour cloning on `trainable_params`
and `set_transformations` keep working.  All batch evaluation (log-likelihood, dist info,
batched action sampling) and every pass of the update run in csrc/policy.cu; the algorithms
pull the flat parameter vector at the start of an update and commit it back at the end.
"""
import numpy as np
import torch
import torch.nn as nn

from ._device import PolicyOnDevice


class MuNet(nn.Module):
    """Two-hidden-layer tanh mean network with affine in/out transforms
    (mjrl/policies/gaussian_mlp.py:143-182).  Host-side container of the parameters."""

    def __init__(self, obs_dim, act_dim, hidden_sizes=(64, 64), in_shift=None, in_scale=None,
                 out_shift=None, out_scale=None):
        super().__init__()
        self.obs_dim, self.act_dim, self.hidden_sizes = obs_dim, act_dim, hidden_sizes
        self.set_transformations(in_shift, in_scale, out_shift, out_scale)
        self.fc0 = nn.Linear(obs_dim, hidden_sizes[0])
        self.fc1 = nn.Linear(hidden_sizes[0], hidden_sizes[1])
        self.fc2 = nn.Linear(hidden_sizes[1], act_dim)

    def set_transformations(self, in_shift=None, in_scale=None, out_shift=None, out_scale=None):
        self.transformations = dict(in_shift=in_shift, in_scale=in_scale, out_shift=out_shift,
                                    out_scale=out_scale)
        as_t = lambda v, fill, k: (torch.from_numpy(np.float32(v)) if v is not None
                                   else fill(k))
        self.in_shift = as_t(in_shift, torch.zeros, self.obs_dim)
        self.in_scale = as_t(in_scale, torch.ones, self.obs_dim)
        self.out_shift = as_t(out_shift, torch.zeros, self.act_dim)
        self.out_scale = as_t(out_scale, torch.ones, self.act_dim)

    def forward(self, x):
        h = (x - self.in_shift) / (self.in_scale + 1e-8)
        h = torch.tanh(self.fc0(h))
        h = torch.tanh(self.fc1(h))
        return self.fc2(h) * self.out_scale + self.out_shift


class _GaussianPolicyBase:
    """Shared by MLP and LinearPolicy: flat-parameter plumbing (gaussian_mlp.py:59-88), the
    device bridge and the GPU-evaluated distribution queries."""

    hidden_sizes = ()

    def _finish_init(self, init_log_std):
        for p in list(self.model.parameters())[-2:]:      # make the last layer small, :35-36
            p.data = 1e-2 * p.data
        self.log_std = torch.ones(self.m) * init_log_std
        self.log_std.requires_grad_(True)
        self.trainable_params = list(self.model.parameters()) + [self.log_std]
        self.old_model = self._make_model()
        self.old_log_std = torch.ones(self.m) * init_log_std
        self.old_params = list(self.old_model.parameters()) + [self.old_log_std]
        for i, p in enumerate(self.old_params):
            p.data = self.trainable_params[i].data.clone()
        self.log_std_val = np.float64(self.log_std.data.numpy().ravel())
        self.param_shapes = [p.data.numpy().shape for p in self.trainable_params]
        self.param_sizes = [p.data.numpy().size for p in self.trainable_params]
        self.d = np.sum(self.param_sizes)
        self.obs_var = torch.randn(self.n)                  # RNG draw kept for seed parity, :57
        self._dev = None
        self._old_aliases_new = False

    # ---- flat parameters --------------------------------------------------------------
    def get_param_values(self):
        return np.concatenate([p.contiguous().view(-1).data.numpy()
                               for p in self.trainable_params]).copy()

    def get_old_param_values(self):
        return np.concatenate([p.contiguous().view(-1).data.numpy()
                               for p in self.old_params]).copy()

    def _assign(self, params, flat):
        k = 0
        for i, p in enumerate(params):
            sz = self.param_sizes[i]
            p.data = torch.from_numpy(
                np.asarray(flat[k:k + sz]).reshape(self.param_shapes[i]).copy()).float()
            k += sz
        params[-1].data = torch.clamp(params[-1].data, self.min_log_std)

    def set_param_values(self, new_params, set_new=True, set_old=True):
        # Reference quirk kept as STATE, not as shared memory: gaussian_mlp.py:66-88 assigns
        # `param.data = torch.from_numpy(vals).float()`, which does not copy a float32 array, so
        # after set_new=set_old=True on float32 input the new and old NETWORK tensors share one
        # buffer (log_std does not: the clamp re-creates it on both sides).  Only in-place
        # optimisers see it (PPO, ppo_clip.py:96); algos/ppo_clip.py reads this flag.
        if set_new and set_old:
            self._old_aliases_new = np.asarray(new_params).dtype == np.float32
        elif set_new or set_old:
            self._old_aliases_new = False
        if set_new:
            self._assign(self.trainable_params, new_params)
            self.log_std_val = np.float64(self.log_std.data.numpy().ravel())
        if set_old:
            self._assign(self.old_params, new_params)

    # ---- device bridge -------------------------------------------------------------------
    def device_view(self):
        if self._dev is None:
            self._dev = PolicyOnDevice(self)
        return self._dev

    def __getstate__(self):
        d = dict(self.__dict__)
        d["_dev"] = None
        # pickle / deepcopy give every tensor of the reference object its own storage again (each
        # torch.from_numpy view of gaussian_mlp.py:66-88 has its own Storage object), so a copy or a
        # checkpoint of an aliased policy is NOT aliased (pinned by tests/golden/stock_policy_mlp.pickle)
        d["_old_aliases_new"] = False
        return d

    # ---- queries -------------------------------------------------------------------------
    def get_action(self, observation):
        """gaussian_mlp.py:92-98.  One observation -> one action (noise from numpy's global
        stream, like the reference).  The batched GPU sampler is `get_action_batch`."""
        dv = self.device_view()
        mean = dv.mean(np.float32(observation).reshape(1, -1)).ravel()
        noise = np.exp(self.log_std_val) * np.random.randn(self.m)
        return [mean + noise, {"mean": mean, "log_std": self.log_std_val, "evaluation": mean}]

    def get_action_batch(self, observations, seed=0, offset=0):
        """(e) batched sampling with Philox noise: obs [B,n] -> (actions, means) fp32 [B,m]."""
        return self.device_view().act(observations, seed, offset)

    def mean_LL(self, observations, actions, model=None, log_std=None):
        which = "old" if (model is self.old_model and model is not None) else "new"
        mean, LL = self.device_view().mean_ll(observations, actions, which)
        return torch.from_numpy(mean), torch.from_numpy(LL)

    def log_likelihood(self, observations, actions, model=None, log_std=None):
        return self.mean_LL(observations, actions, model, log_std)[1].numpy()

    def old_dist_info(self, observations, actions):
        mean, LL = self.mean_LL(observations, actions, self.old_model, self.old_log_std)
        return [LL, mean, self.old_log_std.detach()]

    def new_dist_info(self, observations, actions):
        mean, LL = self.mean_LL(observations, actions, self.model, self.log_std)
        return [LL, mean, self.log_std.detach()]

    def likelihood_ratio(self, new_dist_info, old_dist_info):
        return torch.exp(new_dist_info[0] - old_dist_info[0])

    def mean_kl(self, new_dist_info, old_dist_info):
        """KL(old || new), gaussian_mlp.py:130-140, on already-evaluated dist infos."""
        old_std, new_std = torch.exp(old_dist_info[2]), torch.exp(new_dist_info[2])
        nr = (old_dist_info[1] - new_dist_info[1]) ** 2 + old_std ** 2 - new_std ** 2
        dr = 2 * new_std ** 2 + 1e-8
        return torch.mean(torch.sum(nr / dr + new_dist_info[2] - old_dist_info[2], dim=1))


class MLP(_GaussianPolicyBase):
    def __init__(self, env_spec, hidden_sizes=(64, 64), min_log_std=-3, init_log_std=0, seed=None):
        self.n = env_spec.observation_dim
        self.m = env_spec.action_dim
        self.min_log_std = min_log_std
        self.hidden_sizes = tuple(hidden_sizes)
        if len(self.hidden_sizes) != 2:
            raise ValueError("MLP supports exactly two hidden layers (gaussian_mlp.py:16)")
        if seed is not None:
            torch.manual_seed(seed)
            np.random.seed(seed)
        self.model = self._make_model()
        self._finish_init(init_log_std)

    def _make_model(self):
        return MuNet(self.n, self.m, self.hidden_sizes)
