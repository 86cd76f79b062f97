"""`scripts.als_config` — same module name as the reference; same dataclasses."""
from b200als import ALSConfig, BiasesConfig, CoreConfig, GraphConfig, GraphSimConfig

__all__ = ["ALSConfig", "CoreConfig", "BiasesConfig", "GraphConfig", "GraphSimConfig"]
