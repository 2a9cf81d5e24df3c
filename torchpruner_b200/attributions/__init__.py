from .methods.apoz import APoZAttributionMetric
from .methods.sensitivity import SensitivityAttributionMetric
from .methods.random import RandomAttributionMetric
from .methods.taylor import TaylorAttributionMetric
from .methods.weight_norm import WeightNormAttributionMetric
from .methods.shapley_values import ShapleyAttributionMetric
from .attributions import find_best_module_for_attributions

__all__ = [
    "APoZAttributionMetric", "SensitivityAttributionMetric", "RandomAttributionMetric",
    "TaylorAttributionMetric", "WeightNormAttributionMetric", "ShapleyAttributionMetric",
    "find_best_module_for_attributions",
]
