"""Common machinery of the attribution metrics (reference: attributions/attributions.py).

The public surface (constructor, ``run``, ``find_evaluation_module``, ``run_all_forward``,
``run_all_forward_and_backward``, ``run_forward_partial``, ``aggregate_over_samples``, the cuDNN
determinism toggles) is the reference's.  What differs is underneath: per-batch results stay on
the GPU until the data pass is over (one device->host copy per ``run`` instead of one per batch
or per evaluation), and the per-channel arithmetic is done by the sm_100a kernels behind
``torchpruner_b200._lib``.
"""
from abc import ABC, abstractmethod

import numpy as np
import torch

from .. import _lib
import torch.nn as nn
from torch.nn.modules.conv import _ConvNd

from ..utils import find_best_module_for_attributions
from ..utils.graph import ACTIVATIONS  # noqa: F401  (re-exported like the reference does)

SUPPORTED_OUT_PRUNING_MODULES = [nn.Linear, _ConvNd]


class _RowCollector:
    """Keeps the [B, C] result of every batch on the device; one concatenation and one
    device->host copy at the end (the reference does ``.cpu().numpy()`` + ``np.concatenate``
    per batch: taylor.py:45-48, apoz.py:34-38, sensitivity.py:30-33)."""

    def __init__(self):
        self.blocks = []
        self.fired = 0

    def add(self, rows):
        if self.blocks and rows.shape[1:] != self.blocks[0].shape[1:]:
            raise RuntimeError(
                "the evaluation module produced outputs with different channel counts "
                f"({tuple(self.blocks[0].shape[1:])} then {tuple(rows.shape[1:])}) during one "
                "data pass; it is probably shared between several layers (e.g. a single ReLU "
                "reused inside a ResNet block), for which per-channel attributions are undefined")
        self.blocks.append(rows)

    def numpy(self, what):
        if not self.blocks:
            raise AttributeError(
                f"no {what} was recorded: the evaluation module never ran (or, for gradient "
                "based metrics, its output does not require grad)")
        table = self.blocks[0] if len(self.blocks) == 1 else torch.cat(self.blocks, 0)
        return table.cpu().numpy()


class _ManyCapture:
    """Forward hooks on several evaluation modules for ONE pass over a batch: keeps a reference to
    each module's output (and, on request, to the gradient that reaches it) until ``flush``.

    Nothing is cloned.  A reference is only as good as the tensor stays unmodified, so the
    version counter is noted at hook time and compared at flush time: a module whose output was
    written in place in between (a BatchNorm followed by an in-place ReLU) is reported in
    ``spoiled`` and has to be scored on its own with ``run``, which snapshots."""

    def __init__(self, eval_modules, want_grad):
        self.modules = eval_modules
        self.want_grad = want_grad
        self.z, self.version, self.grad = {}, {}, {}
        self.spoiled = set()
        self.handles = []

    def __enter__(self):
        for k, mod in enumerate(self.modules):
            self.handles.append(mod.register_forward_hook(self._hook(k)))
        return self

    def __exit__(self, *exc):
        for h in self.handles:
            h.remove()
        return False

    def _hook(self, k):
        def forward_hook(_mod, _inputs, output):
            if not isinstance(output, torch.Tensor):
                return
            if k in self.z:   # fired twice in one forward: shared module, undefined per layer
                self.spoiled.add(k)
                return
            if self.want_grad:
                if not output.requires_grad:
                    return
                output.register_hook(lambda grad, k=k: self.grad.__setitem__(k, grad))
            self.z[k] = output.detach()
            self.version[k] = output._version
        return forward_hook

    def flush(self):
        """[(k, z, grad or None)] of the modules captured intact in this batch; clears."""
        items = []
        for k, z in self.z.items():
            if z._version != self.version[k]:
                self.spoiled.add(k)
            if k in self.spoiled or (self.want_grad and k not in self.grad):
                continue
            items.append((k, z, self.grad.get(k)))
        self.z, self.version, self.grad = {}, {}, {}
        return items


def _score_many(metric, modules, kwargs, op, want_grad, what, signed=False):
    """Shared body of the hook metrics' ``run_many``."""
    evals = [_AttributionMetric.run(metric, m, **kwargs) for m in modules]
    distinct = []
    for e in evals:
        if not any(e is d for d in distinct):
            distinct.append(e)
    rows = [_RowCollector() for _ in distinct]
    with _ManyCapture(distinct, want_grad) as cap:
        metric.set_deterministic()
        try:
            for x, y in metric._batches():
                if want_grad:
                    metric.criterion(metric.model(x), y).backward()
                else:
                    with torch.no_grad():
                        metric.model(x)
                items = cap.flush()
                tensors = [(z, g) if op == "taylor" else (g if want_grad else z) for _, z, g in items]
                for (k, _, _), out in zip(items, _lib.reduce_rows_multi(op, tensors, signed)):
                    rows[k].add(out)
        finally:
            metric.restore_deterministic()
    results = {}
    for k, e in enumerate(distinct):
        if k in cap.spoiled:   # needs the snapshot only a pass of its own can take
            results[id(e)] = None
        else:
            results[id(e)] = metric.aggregate_over_samples(rows[k].numpy(what))
    out = []
    for m, e in zip(modules, evals):
        if results[id(e)] is None:
            results[id(e)] = metric.run(m, **kwargs)
        out.append(results[id(e)])
    return out


class _AttributionMetric(ABC):
    def __init__(self, model, data_generator, criterion, device, reduction="mean"):
        assert reduction in ["mean", "none", "sum"] or callable(reduction), \
            'Reduction must be a string in ["mean", "none", "sum"] or a function'
        self.model = model
        self.data_gen = data_generator
        self.criterion = criterion
        self.device = device
        self.reduction = reduction
        self.deterministic = False
        self.benchmark = False
        self._tf32 = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)

    @abstractmethod
    def run(self, module, **kwargs):
        """Type-check ``module`` and resolve the module to evaluate at (attributions.py:28-32)."""
        assert isinstance(module, tuple(SUPPORTED_OUT_PRUNING_MODULES)), \
            f"Attributions can be computed only for the following modules {SUPPORTED_OUT_PRUNING_MODULES}"
        return self.find_evaluation_module(module, **kwargs)

    def run_many(self, modules, **kwargs):
        """Scores of several layers: ``[run(m, **kwargs) for m in modules]``.  Not in the
        reference.  The hook metrics override it to score every layer from ONE pass over the
        data (the reference, and ``run``, repeat the whole forward / backward per layer) and to
        reduce all the layers' activations in one launch."""
        return [self.run(m, **kwargs) for m in modules]

    def find_evaluation_module(self, module, find_best_evaluation_module=False):
        if find_best_evaluation_module is True:
            return find_best_module_for_attributions(self.model, module)
        return module

    # ------------------------------------------------------------------ data passes
    def _batches(self):
        for x, y in self.data_gen:
            yield x.to(self.device), y.to(self.device)

    def run_all_forward(self):
        """Forward pass over the whole loader; per-example losses, concatenated
        (attributions.py:40-56).  Returns None for an empty loader, like the reference."""
        self.set_deterministic()
        pieces = []
        with torch.no_grad():
            for x, y in self._batches():
                pieces.append(self.criterion(self.model(x), y, reduction="none"))
            self.restore_deterministic()
            if not pieces:
                return None
            return pieces[0] if len(pieces) == 1 else torch.cat(pieces, 0)

    def _run_all_forward_no_loss(self):
        """Forward-only data pass for metrics that read activations and ignore the loss."""
        self.set_deterministic()
        with torch.no_grad():
            for x, _ in self._batches():
                self.model(x)
        self.restore_deterministic()

    def run_all_forward_and_backward(self):
        """Forward + backward of the mean loss per batch (attributions.py:58-68).  Parameter
        gradients accumulate across batches and runs exactly as in the reference."""
        self.set_deterministic()
        for x, y in self._batches():
            loss = self.criterion(self.model(x), y)
            loss.backward()
        self.restore_deterministic()

    def run_forward_partial(self, x=None, y_true=None, to_module=None, from_module=None):
        """Partial forward through ``model.forward_partial``; per-example loss when targets are
        given and the pass runs to the end (attributions.py:70-89)."""
        self.set_deterministic()
        loss = None
        y = self.model.forward_partial(x, to_module=to_module, from_module=from_module)
        if y_true is not None and to_module is None:
            loss = self.criterion(y, y_true, reduction="none")
        self.restore_deterministic()
        return y, loss

    # ------------------------------------------------------------------ host aggregation
    def aggregate_over_samples(self, attributions):
        """Host-side numpy aggregation over the example axis (attributions.py:91-106); kept on
        the host so that callable reductions see the full [N, C] matrix and means are
        bit-identical to the reference's ``np.mean``."""
        if self.reduction == "mean":
            return np.mean(attributions, 0)
        if self.reduction == "sum":
            return np.sum(attributions, 0)
        if self.reduction == "none":
            return attributions
        return self.reduction(attributions)

    def set_deterministic(self):
        """cuDNN deterministic and not auto-tuned for the data pass, as the reference does
        (attributions.py:108-111).  Also: no TF32 in cuDNN / cuBLAS for its duration.  The
        scores are specified against the reference's fp32 arithmetic, and PyTorch's default
        (``cudnn.allow_tf32 = True``) would silently run the model's convolutions with 10-bit
        operands on this GPU, which moves Taylor / Sensitivity scores by 1e-2 and Shapley
        scores by more than their own size."""
        self.deterministic = torch.backends.cudnn.deterministic
        self.benchmark = torch.backends.cudnn.benchmark
        self._tf32 = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
        torch.backends.cudnn.deterministic = True
        torch.backends.cudnn.benchmark = False
        torch.backends.cudnn.allow_tf32 = False
        torch.backends.cuda.matmul.allow_tf32 = False

    def restore_deterministic(self):
        torch.backends.cudnn.deterministic = self.deterministic
        torch.backends.cudnn.benchmark = self.benchmark
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = self._tf32
