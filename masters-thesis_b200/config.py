"""The subset of Hydra 1.3 composition that train.py and the sweep scripts use, on plain PyYAML.

Hydra/OmegaConf are not installed in this image, so `train.py` falls back to this
composer.  Supported, and nothing else: a `defaults` list of `group: option` entries (plus `_self_`
and `override hydra/...` entries, which are ignored), command-line overrides `group=option` and
`dotted.key=value`, `-m/--multirun` comma grids, `${a.b}` interpolation and the custom resolver
`${input_size_from_interaction:<bool>}` that train.py:39-42 registers.
"""
from __future__ import annotations

import itertools
import re
from pathlib import Path

import yaml


class Cfg(dict):
    """dict with attribute access (the part of DictConfig train.py relies on)."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v


def _wrap(obj):
    if isinstance(obj, dict):
        return Cfg({k: _wrap(v) for k, v in obj.items()})
    if isinstance(obj, list):
        return [_wrap(v) for v in obj]
    return obj


def to_container(cfg):
    if type(cfg).__module__.startswith("omegaconf"):          # a DictConfig composed by Hydra itself
        from omegaconf import OmegaConf
        return OmegaConf.to_container(cfg, resolve=True)
    if isinstance(cfg, dict):
        return {k: to_container(v) for k, v in cfg.items()}
    if isinstance(cfg, list):
        return [to_container(v) for v in cfg]
    return cfg


RESOLVERS = {"input_size_from_interaction": lambda interaction_only: 3 if interaction_only else 5}

_INTERP = re.compile(r"\$\{([^${}]+)\}")


def _lookup(root, dotted):
    node = root
    for part in dotted.split("."):
        node = node[part]
    return node


def _resolve_value(root, val, depth=0):
    if not isinstance(val, str) or "${" not in val:
        return val
    if depth > 20:
        raise ValueError(f"interpolation too deep: {val}")
    while True:
        m = _INTERP.search(val)
        if not m:
            return val
        expr = m.group(1)
        if ":" in expr:
            name, arg = expr.split(":", 1)
            arg = yaml.safe_load(arg) if isinstance(arg, str) else arg
            rep = RESOLVERS[name.strip()](arg)
        else:
            rep = _resolve_value(root, _lookup(root, expr.strip()), depth + 1)
        if m.start() == 0 and m.end() == len(val):
            return rep
        val = val[:m.start()] + str(rep) + val[m.end():]


def _resolve_tree(root, node):
    if isinstance(node, dict):
        for k in list(node):
            node[k] = _resolve_tree(root, node[k])
        return node
    if isinstance(node, list):
        return [_resolve_tree(root, v) for v in node]
    return _resolve_value(root, node)


def _set_dotted(cfg, dotted, value):
    parts = dotted.split(".")
    node = cfg
    for p in parts[:-1]:
        if p not in node or not isinstance(node[p], dict):
            node[p] = {}
        node = node[p]
    node[parts[-1]] = value


def expand_multirun(overrides):
    """['a=1,2', 'b=x'] -> [['a=1','b=x'], ['a=2','b=x']] (cartesian product, first key slowest)."""
    axes = []
    for ov in overrides:
        key, _, val = ov.partition("=")
        axes.append([f"{key}={v}" for v in val.split(",")] if "," in val else [ov])
    return [list(c) for c in itertools.product(*axes)]


def split_cli(argv):
    """argv -> (multirun flag, list of overrides)."""
    multirun = any(a in ("-m", "--multirun") for a in argv)
    return multirun, [a for a in argv if a not in ("-m", "--multirun") and "=" in a]


def compose_config(config_dir, config_name="config", overrides=()):
    config_dir = Path(config_dir)
    with open(config_dir / f"{config_name}.yaml") as f:
        root = yaml.safe_load(f) or {}
    defaults = root.pop("defaults", [])
    root.pop("hydra", None)
    choices = {}
    for entry in defaults:
        if isinstance(entry, dict):
            for k, v in entry.items():
                if k.startswith("override "):
                    continue                      # hydra/launcher etc.: not part of the job config
                choices[k] = v
    value_overrides = []
    for ov in overrides:
        key, _, val = ov.partition("=")
        key = key.lstrip("+")
        if (config_dir / key).is_dir() and "." not in key:
            choices[key] = val
        elif key.startswith("hydra"):
            continue
        else:
            value_overrides.append((key, yaml.safe_load(val)))
    cfg = {}
    for group, option in choices.items():
        path = config_dir / group / f"{option}.yaml"
        if not path.exists():
            avail = sorted(p.stem for p in (config_dir / group).glob("*.yaml"))
            raise ValueError(f"Could not find '{group}/{option}'. Available options in '{group}': {avail}")
        with open(path) as f:
            cfg[group] = yaml.safe_load(f) or {}
    for k, v in root.items():                     # _self_ last, as in the reference's defaults list
        if isinstance(v, dict) and isinstance(cfg.get(k), dict):
            cfg[k].update(v)
        else:
            cfg[k] = v
    for key, val in value_overrides:
        _set_dotted(cfg, key, val)
    _resolve_tree(cfg, cfg)
    return _wrap(cfg)
