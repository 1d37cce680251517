"""Test infrastructure ONLY -- never imported by the product path.

Minimal stand-ins for the third-party packages the reference imports but this
image does not ship (lightning, torchmetrics, matplotlib), so that the
reference's own ``src`` package (/root/reference/src) can be imported
UNMODIFIED in the build container and used to (a) validate the restatement in
``oracle/lstm_oracle.py`` and (b) generate the golden vectors committed under
``tests/golden/`` (see ``oracle/make_golden.py``).

Nothing here travels to the GPU box as a dependency of tests, smoke() or
bench.py: /root/reference does not exist there.

What the stubs restate (third-party behaviour only):
  * lightning.LightningModule      -> torch.nn.Module + save_hyperparameters()/hparams/log()
  * lightning.LightningDataModule  -> plain object + save_hyperparameters()/hparams
  * torchmetrics.Metric            -> nn.Module with add_state(); forward() returns the
                                      value of the current batch and folds the batch
                                      into the running state (full_state_update=False path)
  * torchmetrics.MeanSquaredError / MeanAbsoluteError
  * matplotlib.pyplot              -> inert namespace (src/plots.py touches plt.style at import)
"""
from __future__ import annotations

import inspect
import sys
import types
from pathlib import Path

import torch

REFERENCE_ROOT = Path("/root/reference")


class _AttrDict(dict):
    __getattr__ = dict.__getitem__

    def __setattr__(self, k, v):
        self[k] = v


def _collect_init_args(frame, ignore=()):
    """Arguments of the __init__ frame that called save_hyperparameters()."""
    info = inspect.getargvalues(frame)
    out = _AttrDict()
    for name in info.args:
        if name == "self" or name in ignore:
            continue
        out[name] = info.locals[name]
    return out


class _HParamsMixin:
    def save_hyperparameters(self, ignore=()):
        if isinstance(ignore, str):
            ignore = (ignore,)
        frame = inspect.currentframe().f_back
        hp = _collect_init_args(frame, ignore)
        if not hasattr(self, "_hparams"):
            object.__setattr__(self, "_hparams", _AttrDict())
        self._hparams.update(hp)

    @property
    def hparams(self):
        if not hasattr(self, "_hparams"):
            object.__setattr__(self, "_hparams", _AttrDict())
        return self._hparams


class LightningModule(_HParamsMixin, torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.logged = {}

    def log(self, name, value, **kw):
        self.logged[name] = value


class LightningDataModule(_HParamsMixin):
    def __init__(self):
        pass


class Metric(torch.nn.Module):
    is_differentiable = None
    higher_is_better = None
    full_state_update = False

    def __init__(self):
        super().__init__()
        self._defaults = {}

    def add_state(self, name, default, dist_reduce_fx=None):
        self._defaults[name] = default.clone()
        setattr(self, name, default.clone())

    def reset(self):
        for k, v in self._defaults.items():
            setattr(self, k, v.clone())

    def forward(self, *args):
        # value of this batch alone, then merge into the running state (sum-type states)
        saved = {k: getattr(self, k) for k in self._defaults}
        self.reset()
        self.update(*args)
        batch_val = self.compute()
        for k in self._defaults:
            setattr(self, k, saved[k] + getattr(self, k))
        return batch_val


class MeanSquaredError(Metric):
    def __init__(self):
        super().__init__()
        self.add_state("sum_squared_error", torch.tensor(0.0))
        self.add_state("total", torch.tensor(0))

    def update(self, preds, target):
        diff = preds - target
        self.sum_squared_error = self.sum_squared_error + torch.sum(diff * diff)
        self.total = self.total + target.numel()

    def compute(self):
        return self.sum_squared_error / self.total


class MeanAbsoluteError(Metric):
    def __init__(self):
        super().__init__()
        self.add_state("sum_abs_error", torch.tensor(0.0))
        self.add_state("total", torch.tensor(0))

    def update(self, preds, target):
        self.sum_abs_error = self.sum_abs_error + torch.sum(torch.abs(preds - target))
        self.total = self.total + target.numel()

    def compute(self):
        return self.sum_abs_error / self.total


def install():
    """Put the stubs into sys.modules and make ``import src`` resolve to the reference."""
    if not REFERENCE_ROOT.exists():
        raise RuntimeError("reference tree not present (only available in the build container)")
    if "lightning" not in sys.modules:
        L = types.ModuleType("lightning")
        L.LightningModule = LightningModule
        L.LightningDataModule = LightningDataModule
        sys.modules["lightning"] = L
    if "torchmetrics" not in sys.modules:
        tm = types.ModuleType("torchmetrics")
        tm.Metric = Metric
        tm.MeanSquaredError = MeanSquaredError
        tm.MeanAbsoluteError = MeanAbsoluteError
        sys.modules["torchmetrics"] = tm
    if "matplotlib" not in sys.modules:
        mpl = types.ModuleType("matplotlib")
        plt = types.ModuleType("matplotlib.pyplot")
        plt.style = types.SimpleNamespace(use=lambda *_a, **_k: None)
        mpl.pyplot = plt
        sys.modules["matplotlib"] = mpl
        sys.modules["matplotlib.pyplot"] = plt
    if str(REFERENCE_ROOT) not in sys.path:
        sys.path.insert(0, str(REFERENCE_ROOT))
    import src  # noqa: F401  (the reference package, unmodified)
    return sys.modules["src"]
