"""B200-native ALS block-update path (drop-in for `scripts/als.py`).

The directory name follows the build contract (`collaborative-filtering_b200`);
because of the hyphen it is imported as `b200als` (alias package at the repo
root) or via `importlib.import_module("collaborative-filtering_b200")`.
"""
from .config import ALSConfig, BiasesConfig, CoreConfig, GraphConfig, GraphSimConfig
from .als import ALS, EPS, SCALE_FACTOR
from .dist import Shard
from . import _lib
from . import evaluate, hostrng

__all__ = ["ALS", "ALSConfig", "CoreConfig", "BiasesConfig", "GraphConfig", "GraphSimConfig",
           "Shard", "EPS", "SCALE_FACTOR"]
