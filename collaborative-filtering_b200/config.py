"""Hyper-parameter surface of the ALS model.

Field names, order and defaults follow the reference's dataclasses
(`scripts/als_config.py:57-95`) exactly, because the reference's drivers build
them positionally / by keyword (`scripts/tune_params.py:281-322`,
`scripts/evaluate_models.py:229-230`).  No validation happens here, as in the
reference; `ALS.__init__` reads the fields.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Literal, Optional

__all__ = ["CoreConfig", "BiasesConfig", "GraphSimConfig", "GraphConfig", "ALSConfig"]


@dataclass
class CoreConfig:
    # latent size / sweeps / L2 on U and V           (als_config.py:57-66)
    n_factors: int
    n_iters: int
    lambda_u: float
    lambda_v: float
    # None → uniform lambda_v; "inverse_sqrt" → lambda_v / sqrt(count_i + 1)
    pop_reg_mode: Optional[Literal["inverse_sqrt"]] = None
    random_state: int = 42
    update_w_every: int = 5


@dataclass
class BiasesConfig:
    # None / 0.0 fall back to lambda_u / lambda_v in ALS.__init__ (als.py:166-167)
    lambda_bu: Optional[float] = None
    lambda_bi: Optional[float] = None


@dataclass
class GraphSimConfig:
    # item-item similarity used by the Laplacian term  (als_config.py:74-81)
    source: Literal["feature", "precomputed"] = "feature"
    feature_name: str = "genres"
    metric: Literal["cosine"] = "cosine"
    topk: Optional[int] = 50
    eps: float = 1e-8


@dataclass
class GraphConfig:
    alpha: float = 0.0                       # Laplacian strength; 0 disables
    sim: Optional[GraphSimConfig] = None     # None → no similarity graph


@dataclass
class ALSConfig:
    core: CoreConfig
    biases: BiasesConfig = field(default_factory=BiasesConfig)
    graph: GraphConfig = field(default_factory=GraphConfig)
