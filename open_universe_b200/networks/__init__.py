from . import universe
from . import universe as universe_orig  # the reference's pristine copy: same classes

__all__ = ["universe", "universe_orig"]
