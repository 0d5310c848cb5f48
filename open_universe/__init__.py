"""Import-path alias: `open_universe` IS `open_universe_b200`.

The reference's documented entry points (README.md:27-49: `from open_universe import inference_utils`,
`inference_utils.load_model(...)`, `model.enhance(noisy)`) and the `_target_` strings of its model configs
(`open_universe.networks.universe.UniverseGAN`, config/model/default.yaml) name the package `open_universe`.
With this directory on the path those imports resolve to the B200-native implementation: every
`open_universe.X` module is the very same module object as `open_universe_b200.X` (no second copy of any class).
Do not install it next to the reference package itself."""
import importlib
import importlib.abc
import importlib.util
import sys

_REAL = "open_universe_b200"


class _AliasFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, name, path=None, target=None):
        if not name.startswith(__name__ + "."):
            return None
        real = _REAL + name[len(__name__):]
        try:
            if importlib.util.find_spec(real) is None:
                return None
        except (ImportError, ValueError):
            return None
        return importlib.util.spec_from_loader(name, self, origin=real)

    def create_module(self, spec):
        return importlib.import_module(spec.origin)

    def exec_module(self, module):
        pass


sys.meta_path.insert(0, _AliasFinder())
_impl = importlib.import_module(_REAL)
__path__ = []   # submodules come from the finder above, never from this directory
from open_universe_b200 import inference_utils, networks  # noqa: E402,F401
