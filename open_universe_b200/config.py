"""Minimal stand-in for the hydra / omegaconf scaffolding used by the reference's
`load_model` (open_universe/inference_utils/model_loader.py:54-62,115-117): yaml ->
attribute dict, `${a.b}` interpolation against the root, `_target_` instantiation.
hydra and omegaconf are not dependencies of this package."""
import re

import yaml


class AttrDict(dict):
    """dict with attribute access (the subset of omegaconf.DictConfig the path uses)."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v


_NUM = re.compile(r"^[-+]?(\d+\.?\d*|\.\d+)([eE][-+]?\d+)?$")


def _coerce(x):
    # PyYAML (YAML 1.1) reads "5e-4" as a string where OmegaConf's loader reads a float
    # (reference config/model/_old/universe_original.yaml:39).
    if isinstance(x, str) and _NUM.match(x):
        return float(x)
    return x


def resolve(root):
    """Returns a copy of `root` with every `${dotted.path}` replaced by the node it names
    (unresolvable references, e.g. into the absent datamodule config, become None)."""

    def lookup(path):
        node = root
        for k in path.split("."):
            node = node[k]
        return node

    def walk(x, depth=0):
        if depth > 32:
            raise ValueError("interpolation cycle in config")
        if isinstance(x, dict):
            return AttrDict({k: walk(v, depth) for k, v in x.items()})
        if isinstance(x, (list, tuple)):
            return [walk(v, depth) for v in x]
        if isinstance(x, str):
            m = re.fullmatch(r"\$\{([^}]+)\}", x.strip())
            if m:
                try:
                    return walk(lookup(m.group(1)), depth + 1)
                except (KeyError, TypeError, IndexError):
                    return None
            return _coerce(x)
        return x

    return walk(root)


def load_config(path):
    with open(path, "r") as f:
        return resolve(yaml.safe_load(f))


def to_attr(x):
    if isinstance(x, dict):
        return AttrDict({k: to_attr(v) for k, v in x.items()})
    if isinstance(x, (list, tuple)):
        return [to_attr(v) for v in x]
    return x


_TARGETS = {}


def register_target(name, cls):
    _TARGETS[name] = cls


def resolve_target(target):
    """Maps the reference's dotted `_target_` paths (any `open_universe.networks.universe*`
    package: config/model/default.yaml:2, default_orig.yaml:2, _old/*.yaml) to this package's classes."""
    name = target.rsplit(".", 1)[-1]
    if target.startswith("open_universe") and name in _TARGETS:
        return _TARGETS[name]
    raise ValueError(f"unsupported _target_ '{target}' (supported: open_universe.networks.universe*.{{{', '.join(_TARGETS)}}})")


def instantiate(node, **kwargs):
    """hydra.utils.instantiate(node, _recursive_=False) for the supported targets."""
    kwargs.pop("_recursive_", None)
    args = {k: v for k, v in node.items() if k != "_target_"}
    args.update(kwargs)
    return resolve_target(node["_target_"])(**args)
