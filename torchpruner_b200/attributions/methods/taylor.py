import torch

from .. import attributions as _base
from ... import _lib


class TaylorAttributionMetric(_base._AttributionMetric):
    """Average absolute first-order Taylor expansion of the loss (Molchanov et al.).

    Reference: attributions/methods/taylor.py.  Per example and channel the score is
    ``| - sum_hw dL/dz * z |`` (no abs with ``signed=True``); the product, the spatial sum and
    the abs are one pass of ``tp_taylor_reduce`` over the activation and its gradient.
    """

    def __init__(self, *args, signed=False, **kwargs):
        super().__init__(*args, **kwargs)
        self.signed = signed

    def run(self, module, **kwargs):
        module = super().run(module, **kwargs)
        rows = _base._RowCollector()
        state = {"trust_no_clone": False, "batches": 0}

        def forward_hook(_mod, _inputs, output):
            if not (isinstance(output, torch.Tensor) and output.requires_grad):
                return
            live = output.detach()
            version = output._version
            # The reference always clones the activation (taylor.py:35): 8 bytes/element of
            # extra HBM traffic.  The clone is only needed if something modifies the tensor in
            # place before backward reaches it (e.g. an in-place ReLU after a BatchNorm).  The
            # first batch clones AND watches the version counter; once a whole batch went by
            # without an in-place write, later batches hold a reference instead.
            snapshot = None if state["trust_no_clone"] else live.clone()
            first = state["batches"] == 0
            state["batches"] += 1

            def grad_hook(grad):
                untouched = live._version == version
                if snapshot is None and not untouched:
                    raise RuntimeError(
                        "the activation of the evaluation module was modified in place between "
                        "forward and backward after an earlier batch left it untouched; "
                        "cannot recover the value the reference would have cloned")
                if first and untouched:
                    state["trust_no_clone"] = True
                z = snapshot if snapshot is not None else live
                rows.add(_lib.taylor_reduce(z, grad, signed=self.signed))

            output.register_hook(grad_hook)

        handle = module.register_forward_hook(forward_hook)
        try:
            self.run_all_forward_and_backward()
        finally:
            handle.remove()
        return self.aggregate_over_samples(rows.numpy("Taylor score"))

    def run_many(self, modules, **kwargs):
        """Taylor scores of several layers from one forward / backward per batch, every layer's
        ``| - sum g * z |`` in one ``tp_reduce_rows_multi`` launch.  Same numbers as ``run`` up to
        the order of the fp32 additions within a row; parameter gradients accumulate once per
        pass over the data instead of once per layer."""
        return _base._score_many(self, modules, kwargs, "taylor", True, "Taylor score",
                                 signed=self.signed)
