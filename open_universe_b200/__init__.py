"""open_universe_b200: B200-native (sm_100a) implementation of the UNIVERSE / UNIVERSE++
`enhance()` hot path of kolyangg/open-universe, behind the reference's Python API."""
from . import inference_utils, networks

__all__ = ["inference_utils", "networks"]
__version__ = "0.1.0"
