"""Structured pruning by physical slicing (reference: torchpruner/pruner/pruner.py).

The public API, the cascade discovery by "NaN trick", the printed lines and the numpy RNG
consumption are the reference's.  The data movement is not: every ``index_select`` the reference
issues for one ``prune_model`` call (weights, biases, their grads, BatchNorm statistics, SGD
momentum — one launch and one host->device index copy each) is recorded in a plan, validated as
a whole, and executed by ONE ``tp_multi_gather`` launch fed by ONE index upload.
"""
import contextlib
import os

import numpy as np
import torch
import torch.nn as nn
from torch.nn.modules.batchnorm import _BatchNorm
from torch.nn.modules.conv import _ConvNd
from torch.nn.modules.dropout import _DropoutNd

from .. import _lib
from . import cascade as _cascade
from .opt_pruner import OptimizerPruner, optimizer_state_tensors, sgd_momentum_buffer  # noqa: F401

SUPPORTED_IN_PRUNING_MODULES = [nn.Linear, _ConvNd, nn.Dropout, _BatchNorm]
SUPPORTED_OUT_PRUNING_MODULES = [nn.Linear, _ConvNd]

_NAN_BITS = 0x7FC00000  # what torch.tensor(np.nan) holds in fp32


class _GatherPlan:
    """Slices waiting to be executed together.

    ``add`` validates and allocates the destination; nothing is launched and no tensor is
    swapped until ``execute``, so an invalid request surfaces before the model is touched
    (the reference leaves a half-sliced model behind when it raises mid-way).
    """

    def __init__(self, device):
        self.device = torch.device(device)
        self.entries = []   # (holder tensor, src, dst, desc fields)
        self.keep_lists = []
        self.keep_len = 0

    def touches(self, tensor):
        return any(e[0] is tensor for e in self.entries)

    def add(self, tensor, axis, keep):
        """Schedule ``tensor.data = tensor.data.index_select(axis, keep)``."""
        src = tensor.data
        if not src.is_cuda:
            raise _lib.TorchPrunerB200Error(
                f"cannot slice a tensor on {src.device}: torchpruner_b200 has no CPU path")
        if not src.is_contiguous():
            src = src.contiguous()
        ndim = src.dim()
        if not -ndim <= axis < ndim:
            raise IndexError(f"Dimension out of range (got {axis} for a {ndim}-d tensor)")
        axis %= ndim
        n_src = src.shape[axis]
        keep = np.ascontiguousarray(keep, dtype=np.int64)
        if keep.size and (keep.min() < 0 or keep.max() >= n_src):
            raise IndexError(f"index out of range for an axis of length {n_src}")
        shape = list(src.shape)
        shape[axis] = int(keep.size)
        dst = torch.empty(shape, dtype=src.dtype, device=src.device)
        outer = int(np.prod(shape[:axis], dtype=np.int64))
        inner = int(np.prod(shape[axis + 1:], dtype=np.int64)) * src.element_size()
        # Identical keep lists (weight / grad / momentum of one parameter) are uploaded once.
        offset = None
        for off, arr in self.keep_lists:
            if arr.size == keep.size and np.array_equal(arr, keep):
                offset = off
                break
        if offset is None:
            offset = self.keep_len
            self.keep_lists.append((offset, keep))
            self.keep_len += int(keep.size)
        self.entries.append((tensor, src, dst, (outer, n_src, int(keep.size), inner, offset)))
        return dst

    def execute(self):
        if not self.entries:
            return
        flat = np.concatenate([arr for _, arr in self.keep_lists]) if self.keep_len else \
            np.zeros(0, np.int64)
        # One index upload and one launch per device that holds a tensor of the plan: a model
        # split over several GPUs is sliced where its tensors live, like the reference's
        # per-tensor index_select.
        by_device = {}
        for entry in self.entries:
            by_device.setdefault(entry[1].device, []).append(entry)
        for dev, entries in by_device.items():
            keep_dev = torch.from_numpy(flat).to(dev)
            descs = [
                _lib.GatherDesc(src.data_ptr(), dst.data_ptr(), outer, n_src, n_keep, inner, offset)
                for _, src, dst, (outer, n_src, n_keep, inner, offset) in entries
            ]
            _lib.multi_gather(descs, keep_dev, dev)
        for holder, _, dst, _ in self.entries:
            holder.data = dst
        self.entries = []
        self.keep_lists = []
        self.keep_len = 0


# Detection hooks are static in the reference and may be installed by hand (its tests do).  The
# results they leave on the device wait in a list until the probe forward is over: the list of the
# Pruner that installed the hooks, or, for hooks installed by hand, this one.
_PENDING_DETECTIONS = []


def _resolve_detections(pending_list=None):
    """One device->host copy for every cascading module of the probe forward (the reference
    synchronises once per module, pruner.py:160-162)."""
    pending = list(_PENDING_DETECTIONS)
    _PENDING_DETECTIONS.clear()
    if pending_list is not None:
        pending += pending_list
        pending_list.clear()
    if not pending:
        return
    packed = torch.cat([torch.cat([cnt, idx]) for _, idx, cnt in pending]).cpu().numpy()
    pos = 0
    for module, idx, _ in pending:
        count = int(packed[pos])
        found = packed[pos + 1: pos + 1 + count].copy()
        pos += 1 + idx.numel()
        if count > 0:
            setattr(module, "_nan_indices", found)


class Pruner:
    def __init__(self, model, input_size, device, optimizer=None, cascade=None):
        """Reference signature (pruner.py:15-19) plus ``cascade``: "analytic" records the module
        chain during the first probe forward and resolves every later cascade from that record
        whenever it is provably what the NaN trick would find (probe forward otherwise);
        "probe" always runs the reference's probe forward.  Default: $TORCHPRUNER_B200_CASCADE,
        else "analytic"."""
        self.model = model
        self.device = device
        self.input_size = input_size
        self.optimizer = optimizer
        self._plan = None
        self.cascade = cascade or os.environ.get("TORCHPRUNER_B200_CASCADE", "analytic")
        assert self.cascade in ("analytic", "probe"), "cascade should be 'analytic' or 'probe'"
        self._trace = None
        self.last_cascade = None  # "analytic" / "probe": how the last prune_model resolved it
        self._pending = []        # detections of this pruner's probe forward, still on the device

    @contextlib.contextmanager
    def _batched(self):
        if self._plan is not None:  # nested: the outer context executes
            yield
            return
        self._plan = _GatherPlan(self.device)
        try:
            yield
            self._plan.execute()
        finally:
            self._plan = None

    # ------------------------------------------------------------------ public API
    def prune_model(self, module, indices, cascading_modules=None):
        """Remove output units ``indices`` of ``module`` and everything downstream that depends
        on them (pruner.py:21-57).  Which input positions of each cascading module die is found
        with the reference's NaN trick: poison the units, run a 2-example probe forward, see
        where the NaNs arrive."""
        if cascading_modules is None:
            print("Warning: no cascading modules defined")
            cascading_modules = []
        if not self._resolve_analytically(module, indices, cascading_modules):
            self.last_cascade = "probe"
            module_handle = module.register_forward_hook(self._nanify_hook(indices))
            handles = [m.register_forward_hook(self._detect_nan_hook(self._pending))
                       for m in cascading_modules]
            # the first probe forward doubles as the structure trace later calls resolve from
            recorder = contextlib.nullcontext()
            if self.cascade == "analytic" and self._trace is None:
                recorder = _cascade.Recorder(self.model)
            try:
                with recorder as rec:
                    self._run_forward()
                if rec is not None:
                    self._trace = rec.trace(self.input_size)
            finally:
                module_handle.remove()
                for handle in handles:
                    handle.remove()
                self._pending.clear()

        with self._batched():
            for next_module in cascading_modules:
                if hasattr(next_module, "_nan_indices"):
                    self.prune_module(next_module, getattr(next_module, "_nan_indices"),
                                      direction="in",
                                      original_len=getattr(next_module, "_activation_len"))
                    delattr(next_module, "_nan_indices")
            self.prune_module(module, indices, direction="out")

    def _resolve_analytically(self, module, indices, cascading_modules):
        """Leave on the cascading modules what the probe forward would (``_activation_len``,
        ``_nan_indices``) without running it; False when that cannot be proven (cascade.py)."""
        if self.cascade != "analytic":
            return False
        if self._trace is not None and self._trace.signature != \
                _cascade.structure_signature(self.model, self.input_size):
            self._trace = None  # modules were replaced or changed mode: record again
        if self._trace is None:
            return False
        found = _cascade.resolve(self._trace, self.model, module, indices, cascading_modules)
        if found is None:
            return False
        # the probe's random batch comes from the global numpy RNG (pruner.py:178): consume the
        # same numbers so whatever the caller draws next is what it would have drawn
        np.random.random((2,) + tuple(self.input_size))
        for m in cascading_modules:
            if id(m) in found:
                length, dead = found[id(m)]
                setattr(m, "_activation_len", length)
                if dead.size:
                    setattr(m, "_nan_indices", dead)
        self.last_cascade = "analytic"
        return True

    def prune_module(self, module, indices, direction="out", original_len=None):
        """Slice the parameters of one module, knowing which ones its class has
        (pruner.py:59-92)."""
        assert direction in ["out", "in"], "direction should be 'out' or 'in'"
        if direction == "out":
            assert isinstance(module, tuple(SUPPORTED_OUT_PRUNING_MODULES)), \
                f"Cannot prune outgoing activations on this module. Only the following are supported {SUPPORTED_OUT_PRUNING_MODULES}"
        else:
            assert isinstance(module, tuple(SUPPORTED_IN_PRUNING_MODULES)), \
                f"Cannot prune incoming activations on this module. Only the following are supported {SUPPORTED_IN_PRUNING_MODULES}"

        print(f"Pruning {len(indices)} units from {module} ({direction})")
        with self._batched():
            if direction == "out":
                self.prune_parameter(module, "weight", indices, axis=0)
                self.prune_parameter(module, "bias", indices, axis=0)
            elif isinstance(module, (nn.Linear, _ConvNd)):
                self.prune_parameter(module, "weight", indices, axis=1)
            elif isinstance(module, _BatchNorm):
                self.prune_parameter(module, "weight", indices, axis=0)
                self.prune_parameter(module, "bias", indices, axis=0)
                self.prune_parameter(module, "running_mean", indices, axis=0)
                self.prune_parameter(module, "running_var", indices, axis=0)
            elif isinstance(module, _DropoutNd):
                self._adjust_dropout(module, indices, original_len)

    def prune_parameter(self, module, parameter_name, indices, axis=0):
        """Slice one tensor of a module, its gradient and its SGD momentum (pruner.py:94-115).
        The Parameter object keeps its identity; its ``.data`` becomes a fresh contiguous
        tensor, exactly what ``index_select`` returns."""
        param = getattr(module, parameter_name)
        if param is None:
            return
        n = param.data.shape[axis]
        mask = np.ones(n, dtype=bool)
        if isinstance(indices, torch.Tensor):
            indices = indices.detach().cpu().numpy()
        mask[indices] = False
        keep = np.arange(n)[mask]
        with self._batched():
            plan = self._plan
            targets = [param]
            if param.grad is not None:
                targets.append(param.grad)
            if self.optimizer is not None and isinstance(param, torch.nn.Parameter):
                targets.extend(optimizer_state_tensors(self.optimizer, param))
            if any(plan.touches(t) for t in targets):
                plan.execute()  # a second slice of the same tensor must see the first one
            for t in targets:
                plan.add(t, axis, keep)

    def _adjust_dropout(self, module, indices, original_len):
        """Lower the drop rate so the expected number of active units is unchanged
        (pruner.py:117-127)."""
        if original_len is None:
            raise RuntimeError("Cannot adjust Dropout rate with 'original_len=None'")
        module.p *= (1.0 - len(indices) / original_len)

    # ------------------------------------------------------------------ NaN trick
    def _nanify_hook(self, indices):
        """Forward hook that overwrites units ``indices`` (dim 1) of a module's output with NaN,
        in place (pruner.py:129-144)."""

        def _hook(_, __, output):
            idx = torch.tensor(indices)
            if idx.dtype != torch.int64:
                raise RuntimeError(
                    f"index_fill_(): Expected dtype int64 for index, got {idx.dtype}")
            C = output.shape[1]
            if idx.numel() and (int(idx.min()) < -C or int(idx.max()) >= C):
                raise IndexError(f"index out of range in self (dimension of size {C})")
            idx = (idx % C).to(output.device) if C else idx.to(output.device)
            if output.is_contiguous():
                return _lib.index_fill_dim1_(output, idx, _NAN_BITS)
            output.copy_(_lib.index_fill_dim1_(output.contiguous(), idx, _NAN_BITS))
            return output

        return _hook

    @staticmethod
    def _detect_nan_hook(sink=None):
        """Forward hook that records which positions along dim 1 of a module's input are NaN
        after summing over every other dim (pruner.py:146-168).  ``_activation_len`` is set at
        once; ``_nan_indices`` (ascending int64 numpy array, only when non-empty) appears when
        the probe forward of ``_run_forward`` has finished."""

        def _hook(module, input, __):
            x = input[0]
            setattr(module, "_activation_len", float(x.shape[1]))
            C = x.shape[1]
            flags = torch.empty(max(C, 1), dtype=torch.int32, device=x.device)
            idx = torch.empty(max(C, 1), dtype=torch.int64, device=x.device)
            cnt = torch.empty(1, dtype=torch.int64, device=x.device)
            _lib.nan_any_dim1(x.detach(), flags, idx, cnt)
            (_PENDING_DETECTIONS if sink is None else sink).append((module, idx, cnt))

        return _hook

    def _run_forward(self, x=None):
        """Probe forward on a random 2-example batch drawn from the global numpy RNG, in the
        model's current mode, with cuDNN forced deterministic (pruner.py:170-185)."""
        d, b = torch.backends.cudnn.deterministic, torch.backends.cudnn.benchmark
        torch.backends.cudnn.deterministic = True
        torch.backends.cudnn.benchmark = False
        try:
            if x is None:
                x = torch.tensor(np.random.random((2,) + tuple(self.input_size))).float().to(self.device)
            y = self.model(x)
            _resolve_detections(self._pending)
        finally:
            torch.backends.cudnn.deterministic = d
            torch.backends.cudnn.benchmark = b
        return y
