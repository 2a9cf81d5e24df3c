"""Graph helpers under the reference's import path (torchpruner/utils)."""
from . import graph

find_best_module_for_attributions = graph.find_best_module_for_attributions
get_vgg_pruning_graph = graph.get_vgg_pruning_graph
get_resnet_pruning_graph = graph.get_resnet_pruning_graph   # not in the reference

__all__ = ["find_best_module_for_attributions", "get_vgg_pruning_graph", "get_resnet_pruning_graph"]
