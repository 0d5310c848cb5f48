"""TEST INFRASTRUCTURE ONLY -- never imported by the product path.

Imports the *real* reference implementation (kolyangg/open-universe mounted
read-only at /root/reference) so that golden vectors can be generated from the
reference's own `enhance()` and so that the CPU restatement in
`oracle/universe_oracle.py` can be validated against it.  The reference does
not exist on the GPU box, so nothing that runs there may import this module;
`oracle/gen_golden.py` is its only consumer.

The reference imports four packages that are absent from this image and used
only as scaffolding (hydra, omegaconf, pytorch_lightning, torch_ema).  They are
replaced by arithmetic-free stubs below.  The heavy
`open_universe/__init__.py` (datasets, metrics ...) is skipped by registering
path-only packages.
"""
import importlib
import os
import sys
import types

import torch
import yaml

REF_ROOT = os.environ.get("UV_REFERENCE_ROOT", "/root/reference")
REF_PKG = os.path.join(REF_ROOT, "open_universe")


class _AttrDict(dict):
    """Minimal omegaconf.DictConfig stand-in: attribute access + .get()."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v


def _wrap(x):
    if isinstance(x, dict):
        return _AttrDict({k: _wrap(v) for k, v in x.items()})
    if isinstance(x, list):
        return [_wrap(v) for v in x]
    return x


def _unwrap(x):
    if isinstance(x, dict):
        return {k: _unwrap(v) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return [_unwrap(v) for v in x]
    return x


def _resolve_target(path):
    mod, _, name = path.rpartition(".")
    return getattr(importlib.import_module(mod), name)


def _instantiate(cfg, *args, **kwargs):
    kwargs.pop("_recursive_", None)
    cfg = dict(cfg)
    target = cfg.pop("_target_")
    cfg.update(kwargs)
    cls = _resolve_target(target)
    return cls(*args, **{k: _wrap(v) for k, v in cfg.items()})


class _EMA:
    """torch_ema.ExponentialMovingAverage stand-in (no arithmetic on the path)."""

    def __init__(self, parameters, decay):
        self.decay = decay
        self.shadow_params = [p.clone().detach() for p in parameters]
        self.collected_params = None
        self.num_updates = 0

    def store(self, parameters):
        self.collected_params = [p.clone().detach() for p in parameters]

    def copy_to(self, parameters):
        for s, p in zip(self.shadow_params, parameters):
            p.data.copy_(s.data)

    def restore(self, parameters):
        for c, p in zip(self.collected_params, parameters):
            p.data.copy_(c.data)

    def to(self, *args, **kwargs):
        self.shadow_params = [p.to(*args, **kwargs) for p in self.shadow_params]

    def update(self, parameters=None):
        pass

    def state_dict(self):
        return {
            "decay": self.decay,
            "num_updates": self.num_updates,
            "shadow_params": self.shadow_params,
            "collected_params": self.collected_params,
        }

    def load_state_dict(self, sd):
        self.decay = sd["decay"]
        self.num_updates = sd["num_updates"]
        self.shadow_params = [p.clone() for p in sd["shadow_params"]]
        self.collected_params = sd.get("collected_params")


def _install_stubs():
    if "hydra" in sys.modules and getattr(sys.modules["hydra"], "_uv_stub", False):
        return
    hydra = types.ModuleType("hydra")
    hydra._uv_stub = True
    hydra_utils = types.ModuleType("hydra.utils")
    hydra_utils.instantiate = _instantiate
    hydra_utils.to_absolute_path = lambda p: p
    hydra.utils = hydra_utils
    sys.modules["hydra"] = hydra
    sys.modules["hydra.utils"] = hydra_utils

    oc = types.ModuleType("omegaconf")

    class OmegaConf:
        create = staticmethod(_wrap)
        to_container = staticmethod(lambda c, **kw: _unwrap(c))

    oc.OmegaConf = OmegaConf
    oc.DictConfig = _AttrDict
    sys.modules["omegaconf"] = oc

    pl = types.ModuleType("pytorch_lightning")

    class LightningModule(torch.nn.Module):
        def save_hyperparameters(self, *a, **k):
            pass

        def log(self, *a, **k):
            pass

    pl.LightningModule = LightningModule
    sys.modules["pytorch_lightning"] = pl

    te = types.ModuleType("torch_ema")
    te.ExponentialMovingAverage = _EMA
    sys.modules["torch_ema"] = te

    if importlib.util.find_spec("wandb") is None:
        sys.modules["wandb"] = types.ModuleType("wandb")

    for name, path in [
        ("open_universe", REF_PKG),
        ("open_universe.networks", REF_PKG + "/networks"),
        ("open_universe.networks.universe_orig", REF_PKG + "/networks/universe_orig"),
    ]:
        m = types.ModuleType(name)
        m.__path__ = [path]
        sys.modules[name] = m


_CLASSES = None


def reference_classes():
    """Returns the reference's (ScoreNetwork, ConditionerNetwork, UniverseGAN, Universe)."""
    global _CLASSES
    if _CLASSES is None:
        if not os.path.isdir(REF_PKG):
            raise RuntimeError(f"reference not mounted at {REF_ROOT}")
        _install_stubs()
        from open_universe.networks.universe_orig.condition import ConditionerNetwork
        from open_universe.networks.universe_orig.score import ScoreNetwork
        from open_universe.networks.universe_orig.universe import Universe
        from open_universe.networks.universe_orig.universe_gan_NS import UniverseGAN

        pkg = sys.modules["open_universe.networks.universe_orig"]
        pkg.ScoreNetwork = ScoreNetwork
        pkg.ConditionerNetwork = ConditionerNetwork
        pkg.UniverseGAN = UniverseGAN
        pkg.Universe = Universe
        _CLASSES = (ScoreNetwork, ConditionerNetwork, UniverseGAN, Universe)
    return _CLASSES


def resolve_interpolations(root):
    """Resolves ${a.b.c} against the root config (only what the model node needs)."""

    def lookup(path):
        node = root
        for k in path.split("."):
            node = node[k]
        return node

    def walk(x):
        if isinstance(x, dict):
            return {k: walk(v) for k, v in x.items()}
        if isinstance(x, list):
            return [walk(v) for v in x]
        if isinstance(x, str) and x.startswith("${") and x.endswith("}"):
            try:
                return walk(lookup(x[2:-1]))
            except (KeyError, TypeError):
                return None
        if isinstance(x, str):
            # PyYAML (YAML 1.1) reads "5e-4" as a string; OmegaConf's loader reads a float
            try:
                return float(x)
            except ValueError:
                return x
        return x

    return walk(root)


MODEL_YAMLS = {
    "universepp_16k": "config/model/default_orig.yaml",
    "universe_16k": "config/model/_old/universe_original.yaml",
    "universepp_24k": "config/model/_old/universepp_24k.yaml",
}


def load_model_cfg(name, overrides=None):
    with open(os.path.join(REF_ROOT, MODEL_YAMLS[name])) as f:
        model = yaml.safe_load(f)
    cfg = resolve_interpolations({"model": model})["model"]
    cfg["training"]["audio_len"] = 2.0
    cfg["validation"]["enh_losses"] = {}
    if overrides:
        for k, v in overrides.items():
            node = cfg
            ks = k.split(".")
            for kk in ks[:-1]:
                node = node[kk]
            node[ks[-1]] = v

    def retarget(x):
        if isinstance(x, dict):
            return {k: retarget(v) for k, v in x.items()}
        if isinstance(x, str):
            return x.replace("networks.universe.", "networks.universe_orig.")
        return x

    return retarget(cfg)


def build_reference_model(name, seed=0, overrides=None):
    """Instantiates the reference model `name` with random init under `seed`."""
    _, _, UniverseGAN, Universe = reference_classes()
    cfg = load_model_cfg(name, overrides)
    cls = UniverseGAN if cfg["_target_"].endswith("UniverseGAN") else Universe
    kw = {k: _wrap(v) for k, v in cfg.items() if k != "_target_"}
    torch.manual_seed(seed)
    model = cls(**kw)
    model.eval()
    return model, cfg


def randomize_prelu_and_film(model, seed=1):
    """Random-init leaves every PReLU slope at 0.25, every weight-norm g at |v|
    and the simple time embedding at 0; perturb them so that a slope / gain mixed
    up between layers changes the output (a stronger parity test)."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for n, p in model.named_parameters():
            if n.startswith("loss_"):
                continue
            if n.endswith("prelu.weight") or n.endswith(".prelu.weight") or n == "_edm_model.prelu.weight" or n == "score_model.prelu.weight":
                p.copy_(0.05 + 0.4 * torch.rand(p.shape, generator=g))
            elif n.endswith("weight_g"):
                p.mul_(0.8 + 0.4 * torch.rand(p.shape, generator=g))
            elif n.endswith("sigma_block.weight"):
                p.fill_(0.7)
            elif n.endswith("sigma_block.bias"):
                p.fill_(-0.3)
            elif n.endswith(".bias") and p.ndim == 1:
                p.add_(0.02 * torch.randn(p.shape, generator=g))


class NoiseReplay:
    """Replaces the reference's module-level `randn` with a replay of pre-drawn tensors."""

    def __init__(self, noises):
        self.noises = list(noises)
        self.i = 0

    def __call__(self, x, sigma, rng=None):
        z = self.noises[self.i].to(x.dtype)
        self.i += 1
        assert z.shape == x.shape, (z.shape, x.shape)
        return z * sigma[:, None, None]


def run_reference_enhance(model, mix, noises, **kw):
    """Runs the reference enhance() with injected noise; returns (out, per-step net outputs)."""
    mod = sys.modules[type(model).__mro__[1].__module__] if type(model).__name__ == "UniverseGAN" else sys.modules[type(model).__module__]
    net = model._edm_model if getattr(model, "with_edm", False) else model.score_model
    outs = []
    hook = net.register_forward_hook(lambda m, i, o: outs.append(o.detach().clone()))
    old = mod.randn
    mod.randn = NoiseReplay(noises)
    try:
        with torch.no_grad():
            y = model.enhance(mix, **kw)
    finally:
        mod.randn = old
        hook.remove()
    return y, outs
