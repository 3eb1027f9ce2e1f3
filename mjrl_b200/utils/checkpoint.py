"""Checkpoint interchange with stock mjrl.

`train_agent` pickles whole objects (mjrl/utils/train_agent.py:61-66: `policy_%i.pickle`,
`baseline_%i.pickle`, `best_policy.pickle`), so a checkpoint IS the attribute layout of
`mjrl.policies.gaussian_mlp.MLP` / `gaussian_linear.LinearPolicy` / the baselines plus the module
path of their classes.  The mirrors in this package keep that attribute layout; this module maps
the module paths, both ways, without needing stock mjrl to be importable:

  * `load(file)`   reads a pickle written by stock mjrl (or by this package) and returns objects of
                   THIS package: class references `mjrl.<x>` resolve to `mjrl_b200.<x>`, the private
                   device fields are (re)created empty, and the reference's new/old storage aliasing
                   (F14, gaussian_mlp.py:66-88) is recovered from the tensors' shared storage.
  * `dump(obj, file)` writes a pickle stock mjrl loads: class references point at `mjrl.<x>`, the
                   private fields of the mirrors are left out, `MLPBaseline` gets back the
                   `torch.optim.Adam` / `MSELoss` members stock code expects (mlp_baseline.py:34-35),
                   with the Adam moments of the device chain filled in.

Host-only: nothing here touches the GPU.
"""
import io
import pickle
import sys
import types

import numpy as np
import torch

# our module  <->  stock module, for the classes a checkpoint can contain
_MODULES = {
    "mjrl_b200.policies.gaussian_mlp": "mjrl.policies.gaussian_mlp",
    "mjrl_b200.policies.gaussian_linear": "mjrl.policies.gaussian_linear",
    "mjrl_b200.baselines.linear_baseline": "mjrl.baselines.linear_baseline",
    "mjrl_b200.baselines.quadratic_baseline": "mjrl.baselines.quadratic_baseline",
    "mjrl_b200.baselines.mlp_baseline": "mjrl.baselines.mlp_baseline",
    "mjrl_b200.utils.gym_env": "mjrl.utils.gym_env",
}
_TO_OURS = {v: k for k, v in _MODULES.items()}
_CLASSES = {"MLP", "MuNet", "LinearPolicy", "LinearModel", "LinearBaseline", "QuadraticBaseline",
            "MLPBaseline", "EnvSpec"}
# fields of the mirrors that stock objects do not have
_PRIVATE = {"_dev", "_old_aliases_new", "_dev_coeffs", "last_fit_info", "learn_rate",
            "adam_betas", "adam_eps", "adam_step", "_host_m", "_host_v", "last_fit_steps"}
# ... per class: stock MLP has no `hidden_sizes`, stock MuNet does (gaussian_mlp.py:153)
_PRIVATE_OF = {"MLP": {"hidden_sizes"}, "LinearPolicy": {"hidden_sizes"}}


# ------------------------------------------------------------------------------------ load
class _Unpickler(pickle.Unpickler):
    def find_class(self, module, name):
        if module in _TO_OURS and name in _CLASSES:
            import importlib
            return getattr(importlib.import_module(_TO_OURS[module]), name)
        return super().find_class(module, name)


def _adopt_policy(p):
    d = p.__dict__
    d.setdefault("_dev", None)
    d["_dev"] = None
    if "hidden_sizes" not in d:
        hs = getattr(p.model, "hidden_sizes", None)
        if hs is None and hasattr(p.model, "fc1"):
            hs = (p.model.fc0.out_features, p.model.fc1.out_features)
        if hs is not None:
            d["hidden_sizes"] = tuple(int(h) for h in hs)
    if "_old_aliases_new" not in d:
        # tensors that shared storage when pickled still do after unpickling
        new0, old0 = p.trainable_params[0], p.old_params[0]
        d["_old_aliases_new"] = bool(new0.data_ptr() == old0.data_ptr())
    # stock set_param_values leaves views into one flat buffer; give every tensor its own storage
    for plist in (p.trainable_params, p.old_params):
        for t in plist:
            t.data = t.data.clone().float()
    return p


def _adopt_mlp_baseline(b):
    d = b.__dict__
    opt = d.pop("optimizer", None)
    d.pop("loss_function", None)
    d.setdefault("learn_rate", 1e-3)
    d.setdefault("adam_betas", (0.9, 0.999))
    d.setdefault("adam_eps", 1e-8)
    d.setdefault("adam_step", 0)
    d.setdefault("_host_m", None)
    d.setdefault("_host_v", None)
    d.setdefault("last_fit_steps", 0)
    d["_dev"] = None
    if opt is not None:
        g = opt.param_groups[0]
        d["learn_rate"], d["adam_betas"], d["adam_eps"] = g["lr"], tuple(g["betas"]), g["eps"]
        d["reg_coef"] = g.get("weight_decay", d.get("reg_coef", 0.0))
        params = list(b.model.parameters())
        st = [opt.state.get(q, None) for q in g["params"]]
        if all(s is not None and "exp_avg" in s for s in st) and len(st) == len(params):
            d["_host_m"] = np.concatenate([s["exp_avg"].reshape(-1).numpy() for s in st]).astype(np.float32)
            d["_host_v"] = np.concatenate([s["exp_avg_sq"].reshape(-1).numpy() for s in st]).astype(np.float32)
            d["adam_step"] = int(st[0]["step"])
    return b


def _adopt(obj):
    name = type(obj).__name__
    if name in ("MLP", "LinearPolicy"):
        return _adopt_policy(obj)
    if name == "MLPBaseline":
        return _adopt_mlp_baseline(obj)
    if name in ("LinearBaseline", "QuadraticBaseline"):
        object.__setattr__(obj, "_dev_coeffs", None)
        if "last_fit_info" not in obj.__dict__:
            object.__setattr__(obj, "last_fit_info", None)
        if "n" not in obj.__dict__ and obj._coeffs is not None and name == "LinearBaseline":
            object.__setattr__(obj, "n", int(np.asarray(obj._coeffs).shape[0]) - 4)
    return obj


def load(file):
    """file: path or binary file object holding ONE pickled policy / baseline (train_agent.py:64-66)."""
    if isinstance(file, (str, bytes)):
        with open(file, "rb") as f:
            return load(f)
    return _adopt(_Unpickler(file).load())


def loads(data):
    return load(io.BytesIO(data))


# ------------------------------------------------------------------------------------ dump
def _stock_class(ours):
    """A class object that pickles as `mjrl.<module>.<name>`: the real stock class when mjrl is
    importable, else a stand-in registered under that module path for the duration of the dump."""
    smod, name = _MODULES[ours.__module__], ours.__name__
    mod = sys.modules.get(smod)
    if mod is not None and hasattr(mod, name):
        return getattr(mod, name), None
    parts, made = smod.split("."), []
    for i in range(1, len(parts) + 1):
        key = ".".join(parts[:i])
        if key not in sys.modules:
            sys.modules[key] = types.ModuleType(key)
            made.append(key)
    base = (torch.nn.Module,) if issubclass(ours, torch.nn.Module) else (object,)
    cls = type(name, base, {"__module__": smod})
    setattr(sys.modules[smod], name, cls)
    return cls, (made, smod, name)


def _release(token):
    if token is None:
        return
    made, smod, name = token
    if smod in sys.modules and hasattr(sys.modules[smod], name) and smod not in made:
        delattr(sys.modules[smod], name)
    for key in made:
        sys.modules.pop(key, None)


def _as_stock(obj, tokens):
    """Shallow re-labelling of one mirror object (and its `model` / `old_model`) as stock classes."""
    ours = type(obj)
    if ours.__module__ not in _MODULES or ours.__name__ not in _CLASSES:
        return obj
    cls, tok = _stock_class(ours)
    tokens.append(tok)
    state = obj.__getstate__() if hasattr(obj, "__getstate__") and not isinstance(obj, torch.nn.Module) \
        else dict(obj.__dict__)
    name = ours.__name__
    drop = _PRIVATE | _PRIVATE_OF.get(name, set())
    state = {k: v for k, v in dict(state).items() if k not in drop}
    if name in ("MLP", "LinearPolicy"):
        state["model"] = _as_stock(obj.model, tokens)
        state["old_model"] = _as_stock(obj.old_model, tokens)
    if name == "MLPBaseline":
        obj._sync_host() if hasattr(obj, "_sync_host") else None
        params = list(obj.model.parameters())
        opt = torch.optim.Adam(params, lr=obj.learn_rate, weight_decay=obj.reg_coef, betas=obj.adam_betas,
                               eps=obj.adam_eps)
        full = obj.__getstate__() if hasattr(obj, "__getstate__") else obj.__dict__
        m, v = full.get("_host_m"), full.get("_host_v")
        if m is not None and obj.adam_step > 0:
            i = 0
            for q in params:
                k = q.numel()
                opt.state[q] = dict(step=torch.tensor(float(obj.adam_step)),
                                    exp_avg=torch.from_numpy(np.array(m[i:i + k])).reshape(q.shape),
                                    exp_avg_sq=torch.from_numpy(np.array(v[i:i + k])).reshape(q.shape))
                i += k
        state["optimizer"] = opt
        state["loss_function"] = torch.nn.MSELoss()
    new = cls.__new__(cls)
    new.__dict__.update(state)
    return new


def dumps(obj, protocol=pickle.DEFAULT_PROTOCOL):
    tokens = []
    try:
        return pickle.dumps(_as_stock(obj, tokens), protocol=protocol)
    finally:
        for tok in reversed(tokens):
            _release(tok)


def dump(obj, file, protocol=pickle.DEFAULT_PROTOCOL):
    if isinstance(file, (str, bytes)):
        with open(file, "wb") as f:
            return dump(obj, f, protocol)
    file.write(dumps(obj, protocol))
