"""Drop-in for `open_universe.networks.universe` (and `universe_orig`): same class names,
constructor arguments and state-dict layout; compute is native sm_100a."""
from ...config import register_target
from .condition import ConditionerNetwork
from .score import ScoreNetwork
from .universe import Universe, UniverseGAN

for _c in (ScoreNetwork, ConditionerNetwork, Universe, UniverseGAN):
    register_target(_c.__name__, _c)

__all__ = ["ScoreNetwork", "ConditionerNetwork", "Universe", "UniverseGAN"]
