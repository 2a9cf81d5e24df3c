from .pruner import Pruner
from .opt_pruner import OptimizerPruner

__all__ = ["Pruner", "OptimizerPruner"]
