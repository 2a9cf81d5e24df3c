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

    def run_many(self, modules, **kwargs):
        """Weight norms of several layers: one launch for all weights whose rows fit a tile, one
        device->host copy per layer."""
        modules = [_base._AttributionMetric.run(self, m, **kwargs) for m in modules]
        views = [m.weight.detach().reshape(m.weight.shape[0], 1, -1) for m in modules]
        outs = _lib.reduce_rows_multi("abs_sum", views)
        return [o.reshape(-1).cpu().numpy() for o in outs]

    def find_evaluation_module(self, module, find_best_evaluation_module=False):
        return module  # scores are a property of the weights of `module` itself
