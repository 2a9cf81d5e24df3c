"""Analytic cascade resolution: which inputs of each cascading module die when output units of a
module are removed — without the probe forward of the NaN trick.

The reference answers that question empirically (pruner.py:21-54, 129-185): it overwrites the
units with NaN in a forward hook, runs a random 2-example batch through the whole model and has
every cascading module report the dim-1 positions of its input that arrive as NaN — one model
forward and one device->host synchronisation per cascading module for every ``prune_model``.

For the topologies the reference targets the answer is a matter of bookkeeping:

  * modules that act per channel (BatchNorm, activations, pooling, Dropout, Identity) hand the
    dead positions on unchanged;
  * flattening ``[B, C, *S]`` to ``[B, C*prod(S)]`` turns dead channel ``c`` into the contiguous
    block ``[c*prod(S), (c+1)*prod(S))`` (the mapping the reference tests in
    test_pruner.py:83-92);
  * a Linear / ungrouped convolution mixes every input into every output, so behind the first
    one that sees a dead input everything is dead.

What must be *known* for this to be exact is the order in which the leaf modules run, that each
one consumes exactly what the previous one produced (no functional op, residual add or in-place
update hidden between them) and the spatial size at the flatten.  That is established once per
model structure by a ``Recorder`` riding on the FIRST probe forward the pruner runs anyway: it
notes every leaf call with the identity and version counter of its input and output.  From the
second ``prune_model`` on, ``resolve`` walks that record instead of running the model — no
forward, no kernel launch, no synchronisation; whenever it meets
something it cannot prove (an unknown module type, a broken link, a version bump between two
modules, grouped convolutions, a module in training mode that has side effects) it returns
``None`` and the caller falls back to the NaN trick.  The results are asserted equal to the
NaN trick's in tests/test_cascade.py (CPU, against the oracle) and tests/test_gpu_api.py.
"""
import numpy as np
import torch
import torch.nn as nn
from torch.nn.modules.batchnorm import _BatchNorm
from torch.nn.modules.conv import _ConvNd
from torch.nn.modules.dropout import _DropoutNd

# element-wise activations: NaN in, NaN out, nothing crosses channels
_ELEMENTWISE = (nn.ReLU, nn.ReLU6, nn.LeakyReLU, nn.ELU, nn.SELU, nn.CELU, nn.GELU, nn.SiLU,
                nn.Mish, nn.Sigmoid, nn.Tanh, nn.Hardtanh, nn.Hardswish, nn.Hardsigmoid,
                nn.Softplus, nn.PReLU, nn.Identity)
_POOLING = (nn.MaxPool1d, nn.MaxPool2d, nn.MaxPool3d, nn.AvgPool1d, nn.AvgPool2d, nn.AvgPool3d,
            nn.AdaptiveAvgPool1d, nn.AdaptiveAvgPool2d, nn.AdaptiveAvgPool3d,
            nn.AdaptiveMaxPool1d, nn.AdaptiveMaxPool2d, nn.AdaptiveMaxPool3d)


class _Call:
    __slots__ = ("module", "inp", "in_shape", "in_version", "in_fn", "out", "out_shape",
                 "out_version", "out_fn")


def _same_storage(x, y):
    """x is y, or a view of it (or of the tensor y is a view of)."""
    if x is y:
        return True
    base = getattr(x, "_base", None)
    return base is not None and (base is y or base is getattr(y, "_base", None))


def _direct(a, b):
    """How call ``b`` receives the output of call ``a`` unchanged: 'same', 'flatten' or None."""
    if a.out is None or b.inp is None or not _same_storage(b.inp, a.out):
        return None
    if b.in_version != a.out_version:
        return None   # written in place in between: see _sum_of for the one case that is provable
    if b.inp is a.out or tuple(a.out_shape) == tuple(b.in_shape):
        return "same"
    if len(a.out_shape) > 2 and len(b.in_shape) == 2 and a.out_shape[0] == b.in_shape[0] \
            and int(np.prod(a.out_shape[1:])) == b.in_shape[1]:
        return "flatten"
    return None


def _sum_of(calls, i):
    """Indices (or -1 for the model input) of the two tensors whose elementwise SUM call ``i``
    receives — the residual add of a ResNet block, functional (``out = out + identity``) or in
    place (``out += identity``) — or None.  A hook cannot see the add itself; the autograd graph
    can: the input's grad_fn is an AddBackward0 node whose two operands are the grad_fns that
    two recorded outputs (or inputs) had when they were produced."""
    fn = calls[i].in_fn
    if fn is None or type(fn).__name__ != "AddBackward0" or len(fn.next_functions) != 2:
        return None
    found = []
    for node, _ in fn.next_functions:
        if node is None:
            return None
        src = None
        for j in range(i - 1, -1, -1):
            if calls[j].out_fn is node and tuple(calls[j].out_shape) == tuple(calls[i].in_shape):
                src = j
                break
        if src is None:
            return None
        found.append(src)
    return found


class ChainTrace:
    """Leaf-module calls of one forward, in order, and where each one's input comes from: the
    unchanged (or flattened) output of ONE earlier call, or the sum of two.  Only modules, shapes
    and sources are kept; the activations seen while recording are released."""

    def __init__(self, calls, signature):
        self.signature = signature
        self.calls = calls
        self.sources = []
        for i, c in enumerate(calls):
            src = None
            for j in range(i - 1, -1, -1):   # the latest producer wins
                kind = _direct(calls[j], c)
                if kind is not None:
                    src = (kind, j)
                    break
                if calls[j].out is not None and c.inp is not None and _same_storage(c.inp, calls[j].out):
                    break   # its tensor, but modified since: not a plain hand-over
            if src is None:
                pair = _sum_of(calls, i)
                if pair is not None:
                    src = ("add", pair[0], pair[1])
            self.sources.append(src)
        for c in calls:
            c.inp = c.out = c.in_fn = c.out_fn = None
        self.position = {}
        for i, c in enumerate(calls):
            # a module called twice has no single position: poison it
            self.position[id(c.module)] = i if id(c.module) not in self.position else -1

    def link(self, i):
        """Source of call i+1 if it is simply call i's output ('same' / 'flatten'), else None."""
        src = self.sources[i + 1]
        return src[0] if src is not None and src[0] != "add" and src[1] == i else None


def structure_signature(model, input_size):
    """Everything the trace depends on that pruning channels does not change."""
    sig = [tuple(input_size)]
    for m in model.modules():
        if next(m.children(), None) is not None:
            continue
        sig.append((id(m), type(m).__name__, bool(m.training)))
    return tuple(sig)


class Recorder:
    """Forward pre-/post-hooks on every leaf module that note the identity, shape and version
    counter of what goes in and what comes out.  The product installs it around the probe
    forward of the first ``prune_model`` (after the poisoning hook, so the output it sees is
    the one the next module receives); every later call on the same structure is bookkeeping."""

    def __init__(self, model):
        self.model = model
        self.calls, self._open, self._handles = [], {}, []

    def _pre(self, module, inputs):
        c = _Call()
        c.module = module
        x = inputs[0] if inputs else None
        if isinstance(x, torch.Tensor):
            c.inp, c.in_shape, c.in_version, c.in_fn = x, tuple(x.shape), x._version, x.grad_fn
        else:
            c.inp, c.in_shape, c.in_version, c.in_fn = None, (), -1, None
        self._open.setdefault(id(module), []).append(c)

    def _post(self, module, inputs, output):
        c = self._open[id(module)].pop()
        if isinstance(output, torch.Tensor):
            c.out, c.out_shape, c.out_version = output, tuple(output.shape), output._version
            c.out_fn = output.grad_fn
        else:
            c.out, c.out_shape, c.out_version, c.out_fn = None, (), -1, None
        self.calls.append(c)

    def __enter__(self):
        for m in self.model.modules():
            if next(m.children(), None) is None:
                self._handles.append(m.register_forward_pre_hook(self._pre))
                self._handles.append(m.register_forward_hook(self._post))
        return self

    def __exit__(self, *exc):
        for h in self._handles:
            h.remove()
        self._handles = []
        return False

    def trace(self, input_size):
        return ChainTrace(self.calls, structure_signature(self.model, input_size))


def trace_chain(model, input_size):
    """Record one forward of a zero batch of 2 (what the tests use; the product records its own
    probe forward instead).  Refused (None) when the forward would have side effects."""
    if _has_side_effects_in_training(model):
        return None
    param = next(model.parameters(), None)
    device = param.device if param is not None else torch.device("cpu")
    with Recorder(model) as rec:   # grad mode on, like the probe forward: adds show in grad_fn
        model(torch.zeros((2,) + tuple(input_size), dtype=torch.float32, device=device))
    return rec.trace(input_size)


def normalise_indices(indices, n_units):
    """The units ``index_fill_`` would poison: int64 required, negatives wrap, duplicates and
    order irrelevant (pruner.py:136-141)."""
    idx = torch.as_tensor(indices) if not isinstance(indices, torch.Tensor) else indices.detach().cpu()
    if idx.dtype != torch.int64:
        raise RuntimeError(f"index_fill_(): Expected dtype int64 for index, got {idx.dtype}")
    idx = idx.reshape(-1).numpy()
    if idx.size and (idx.min() < -n_units or idx.max() >= n_units):
        raise IndexError(f"index out of range in self (dimension of size {n_units})")
    return np.unique(idx % n_units) if n_units else idx


def _has_side_effects_in_training(model):
    for m in model.modules():
        if m.training and isinstance(m, (_BatchNorm, _DropoutNd)):
            return True
    return False


def resolve(trace, model, module, indices, cascading_modules):
    """Dead input positions of every cascading module, as the NaN trick would report them.

    Returns ``{id(cascading module): (activation_len, sorted int64 ndarray)}`` — modules that the
    forward reaches get an entry even when nothing dies in them (the reference sets
    ``_activation_len`` on every hooked module it reaches) — or None when the answer cannot be
    proven from the trace.
    """
    if trace is None or _has_side_effects_in_training(model):
        # train-mode BatchNorm / Dropout: the reference's probe forward moves running statistics
        # and draws from the torch RNG; only the real forward reproduces that
        return None
    start = trace.position.get(id(module), None)
    if start is None or start < 0:
        return None
    wanted = {}
    for m in cascading_modules:
        pos = trace.position.get(id(m), None)
        if pos is None:
            continue  # never called: the reference's hook never fires either
        if pos < 0:
            return None
        wanted[pos] = m
    first = trace.calls[start]
    if len(first.out_shape) < 2 or not isinstance(module, (nn.Linear, _ConvNd)):
        return None
    # Per call, what its OUTPUT carries: (recorded width, live width, dead positions).  ``rec`` is
    # what the trace saw when it was recorded and is only compared with the trace (is this still
    # the graph that was proven?).  ``live`` is what the model holds NOW, read from the
    # parameters: an earlier prune_model has sliced them since the recording (the structure
    # signature deliberately ignores sizes), and indices, Dropout's ``original_len`` and the
    # everything-dead range must refer to the current widths.  Calls that do not depend on the
    # poisoned module have no entry: they are clean.
    live0 = int(module.weight.shape[0])
    state = {start: (first.out_shape[1], live0, normalise_indices(indices, live0))}
    result = {}
    last = max(wanted) if wanted else start
    for pos, m in wanted.items():
        if pos <= start:  # runs before the poisoned module: reached, nothing dead
            width_now = _live_width_into(trace, pos)
            if width_now is None:
                return None
            result[id(m)] = (float(width_now), np.zeros(0, np.int64))
    for i in range(start + 1, last + 1):
        src = trace.sources[i]
        call = trace.calls[i]
        m = call.module
        if src is None:
            # Unknown origin.  Harmless only if nothing poisoned can have reached it: the first
            # call of the model (fed by the input) — anything else may hide an op on dead data.
            return None
        if src[0] == "add":
            parts = [state.get(j) for j in src[1:]]
            if all(p is None for p in parts):
                incoming = None
            else:
                known = [p for p in parts if p is not None]
                if any(p[0] != known[0][0] or p[1] != known[0][1] for p in known):
                    return None
                # NaN + anything = NaN: a position is dead if it is dead in either operand
                incoming = (known[0][0], known[0][1],
                            np.unique(np.concatenate([p[2] for p in known])))
        else:
            kind, j = src
            incoming = state.get(j)
            if incoming is not None and kind == "flatten":
                k = int(np.prod(trace.calls[j].out_shape[2:]))
                rec, live, dead = incoming
                dead = (dead[:, None] * k + np.arange(k, dtype=np.int64)[None, :]).reshape(-1)
                incoming = (rec * k, live * k, dead)
        if incoming is None:   # a branch the poisoned module does not feed
            if i in wanted:
                width_now = _live_width_into(trace, i)
                if width_now is None:
                    return None
                result[id(m)] = (float(width_now), np.zeros(0, np.int64))
            continue
        rec, live, dead = incoming
        if len(call.in_shape) < 2 or call.in_shape[1] != rec:
            return None
        width_now = _live_in_width(m)
        if width_now is not None and width_now != live:
            return None   # the model no longer fits together here: let the probe forward say so
        if i in wanted:
            result[id(m)] = (float(live), dead.astype(np.int64))
        # what the module hands on
        if isinstance(m, (nn.Linear, _ConvNd)):
            if isinstance(m, _ConvNd) and m.groups != 1:
                if i == last:
                    break
                return None
            if len(call.out_shape) < 2:
                return None
            rec, live = call.out_shape[1], int(m.weight.shape[0])
            dead = np.arange(live, dtype=np.int64) if dead.size else dead
        elif isinstance(m, nn.Flatten):
            if m.start_dim != 1 or m.end_dim not in (-1, len(call.in_shape) - 1):
                return None
            k = int(np.prod(call.in_shape[2:])) if len(call.in_shape) > 2 else 1
            dead = (dead[:, None] * k + np.arange(k, dtype=np.int64)[None, :]).reshape(-1)
            rec, live = rec * k, live * k
        elif isinstance(m, (_BatchNorm, _DropoutNd) + _ELEMENTWISE + _POOLING):
            if len(call.out_shape) < 2 or call.out_shape[1] != rec:
                return None
        else:
            if i == last:
                break
            return None
        state[i] = (rec, live, dead)
    return result


def _live_in_width(m):
    """Input positions along dim 1 that module ``m`` accepts right now, read from its tensors
    (None for modules that hold none: their width is whatever arrives)."""
    if isinstance(m, (nn.Linear, _ConvNd)):
        return int(m.weight.shape[1]) * (m.groups if isinstance(m, _ConvNd) else 1)
    if isinstance(m, _BatchNorm):
        for t in (m.weight, m.running_mean):
            if t is not None:
                return int(t.shape[0])
    return None


def _live_width_into(trace, pos):
    """Current dim-1 width of what call ``pos`` receives: from its own tensors when it has some,
    else followed back along the recorded sources to the nearest Linear / convolution (or the
    model input, which pruning never narrows).  None when the origin is unproven."""
    width = _live_in_width(trace.calls[pos].module)
    if width is not None:
        return width
    src = trace.sources[pos]
    if src is None:
        if pos == 0 and len(trace.calls[0].in_shape) >= 2:
            return trace.calls[0].in_shape[1]
        return None
    if src[0] == "add":
        return _live_width_out(trace, src[1])
    out = _live_width_out(trace, src[1])
    if out is None:
        return None
    if src[0] == "flatten":
        out *= int(np.prod(trace.calls[src[1]].out_shape[2:]))
    return out


def _live_width_out(trace, j):
    """Current dim-1 width of call ``j``'s output."""
    call = trace.calls[j]
    m = call.module
    if isinstance(m, (nn.Linear, _ConvNd)):
        return int(m.weight.shape[0])
    width = _live_width_into(trace, j)
    if width is None:
        return None
    if isinstance(m, nn.Flatten):
        if m.start_dim != 1 or m.end_dim not in (-1, len(call.in_shape) - 1):
            return None
        return width * (int(np.prod(call.in_shape[2:])) if len(call.in_shape) > 2 else 1)
    if isinstance(m, (_BatchNorm, _DropoutNd) + _ELEMENTWISE + _POOLING):
        return width
    return None
