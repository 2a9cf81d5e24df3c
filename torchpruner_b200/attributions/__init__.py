"""Attribution metrics under the reference's import path (torchpruner/attributions/__init__.py
exports the same seven names)."""
from . import attributions as _base
from .methods import apoz, random, sensitivity, shapley_values, taylor, weight_norm

APoZAttributionMetric = apoz.APoZAttributionMetric
SensitivityAttributionMetric = sensitivity.SensitivityAttributionMetric
RandomAttributionMetric = random.RandomAttributionMetric
TaylorAttributionMetric = taylor.TaylorAttributionMetric
WeightNormAttributionMetric = weight_norm.WeightNormAttributionMetric
ShapleyAttributionMetric = shapley_values.ShapleyAttributionMetric
find_best_module_for_attributions = _base.find_best_module_for_attributions

__all__ = [name for name in dir() if name.endswith("AttributionMetric")] + \
    ["find_best_module_for_attributions"]
