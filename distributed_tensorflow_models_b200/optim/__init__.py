"""Optimiser math of the reference (SGD, TF1 RMSProp, EMA, staircase LR decay) as flat-buffer updates."""
from .schedule import exponential_decay  # noqa: F401
from .updates import EmaConfig, OptimizerConfig, ShardOptimizer  # noqa: F401
