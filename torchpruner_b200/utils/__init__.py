from .graph import find_best_module_for_attributions
from .graph import get_vgg_pruning_graph

__all__ = ["find_best_module_for_attributions", "get_vgg_pruning_graph"]
