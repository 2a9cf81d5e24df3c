from .. import attributions as _base
from ... import _lib


class WeightNormAttributionMetric(_base._AttributionMetric):
    """Sum of absolute weights per output unit (Li et al., ICLR 2017).

    Reference: attributions/methods/weight_norm.py — there the whole weight is copied to the
    host and reduced with numpy; here ``tp_weight_abs_sum`` reduces it where it lives and only
    the [Cout] result crosses PCIe.  Ignores data, criterion, ``reduction`` and the
    evaluation-module flag, like the reference.
    """

    def run(self, module, **kwargs):
        module = super().run(module, **kwargs)
        return _lib.weight_abs_sum(module.weight.detach()).cpu().numpy()

    def find_evaluation_module(self, module, find_best_evaluation_module=False):
        return module  # scores are a property of the weights of `module` itself
