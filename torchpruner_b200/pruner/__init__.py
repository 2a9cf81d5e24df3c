"""Structured pruning under the reference's import path (torchpruner/pruner)."""
from . import opt_pruner, pruner

Pruner = pruner.Pruner
OptimizerPruner = opt_pruner.OptimizerPruner

__all__ = ["Pruner", "OptimizerPruner"]
