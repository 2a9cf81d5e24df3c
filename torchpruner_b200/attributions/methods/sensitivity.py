import torch

from .. import attributions as _base
from ... import _lib


class SensitivityAttributionMetric(_base._AttributionMetric):
    """Average absolute gradient of the loss w.r.t. the activation (Mittal et al.).

    Reference: attributions/methods/sensitivity.py — ``sum_hw |dL/dz|`` per example and channel,
    one pass of ``tp_abs_sum`` over the gradient.
    """

    def run(self, module, **kwargs):
        module = super().run(module, **kwargs)
        rows = _base._RowCollector()

        def forward_hook(_mod, _inputs, output):
            if isinstance(output, torch.Tensor) and output.requires_grad:
                output.register_hook(lambda grad: rows.add(_lib.abs_sum(grad)))

        handle = module.register_forward_hook(forward_hook)
        try:
            self.run_all_forward_and_backward()
        finally:
            handle.remove()
        return self.aggregate_over_samples(rows.numpy("gradient"))

    def run_many(self, modules, **kwargs):
        """``sum |dL/dz|`` of several layers from one forward / backward per batch and one
        ``tp_reduce_rows_multi`` launch."""
        return _base._score_many(self, modules, kwargs, "abs_sum", True, "gradient")
