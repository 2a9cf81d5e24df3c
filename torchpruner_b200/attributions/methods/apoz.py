import torch

from .. import attributions as _base
from ... import _lib


class APoZAttributionMetric(_base._AttributionMetric):
    """1 - APoZ (Average Percentage of Zeros), Hu et al.

    Reference: attributions/methods/apoz.py.  Per example and channel: the number of strictly
    positive activations over the trailing dims (a count, not a fraction, for conv outputs —
    as in the reference).  ``tp_apoz_count`` counts with integers, so results are bit-exact.
    The reference also evaluates and concatenates the per-example loss during this pass and
    throws it away (attributions.py:50-54); that is skipped here.
    """

    def run(self, module, **kwargs):
        module = super().run(module, **kwargs)
        rows = _base._RowCollector()

        def forward_hook(_mod, _inputs, output):
            rows.add(_lib.apoz_count(output.detach()))

        with torch.no_grad():
            handle = module.register_forward_hook(forward_hook)
            try:
                self._run_all_forward_no_loss()
            finally:
                handle.remove()
            return self.aggregate_over_samples(rows.numpy("activation count"))

    def run_many(self, modules, **kwargs):
        """Activation counts of several layers from one forward pass per batch and one
        ``tp_reduce_rows_multi`` launch; integer counts, so identical to ``run``."""
        return _base._score_many(self, modules, kwargs, "apoz", False, "activation count")
