"""fp32-accurate tensor-core execution of the layers after the evaluation module.

The reference evaluates every Shapley mask by calling the user's ``forward_partial`` — an opaque
callback (attributions.py:85).  When the layers it runs form a plain chain of
Conv2d(3x3, stride 1, pad 1) / BatchNorm2d / ReLU / LeakyReLU / MaxPool2d(2) / Flatten / Dropout /
Linear
modules in eval mode (every VGG-style tail does), the same arithmetic is executed here by
``tp_conv_f16x3`` / ``tp_conv_tf32x3`` (tcgen05, three MMAs per fp32 product on an 11-bit split of
the operands, NHWC) instead of by cuDNN/cuBLAS kernels:

* the chain is *discovered*, not assumed: one ``forward_partial`` call on the unmasked
  activation is traced with forward hooks; anything that is not a supported leaf module applied
  to the previous module's output disqualifies the tail and the generic path is used;
* the plan is *verified* before it is trusted: its logits for the first masks of the run are
  compared with the user's ``forward_partial`` on the same masked activations.

BatchNorm (eval) and the conv/linear bias are folded into one per-channel scale/shift applied in
the kernel's epilogue; ReLU / LeakyReLU are fused; a MaxPool that follows is fused into the
epilogue too where the layer runs on position-major tiles (``tp_conv_f16x3_pool2``) and is a
separate NHWC kernel otherwise.  Channel counts are zero-padded to the kernels' 64-channel
k-block inside the engine's own copies, so any layer width runs.  The first layer of the tail is
evaluated incrementally along each permutation (csrc/incremental.cu).  The fp16 flavour is the
default; it falls back to the TF32 one when verification fails or a score comes out non-finite.
"""
import logging
import math
import os
import weakref

import torch
import torch.nn as nn

from .. import _lib

VERIFY_RTOL = 2e-5   # on logits, relative to their largest magnitude
VERIFY_ARBITRATE_BELOW = 1e-3   # above this the difference is structural, not rounding
# Incremental first layer: "stored" computes the per-channel contributions once per batch into a
# [C, pixels*C1] scratch (4.3 GB for a VGG16 block at bs 256) and streams them back per mask;
# "direct" recomputes them where they are consumed (no scratch; measured within 1 % of "stored"
# on the VGG16 job, profiles/r01_prefix_direct_vs_stored.txt).  "auto": stored while the scratch
# stays under INCREMENTAL_MAX_BYTES, direct beyond it.
CONTRIB = os.environ.get("TORCHPRUNER_B200_CONTRIB", "auto")
FUSE_POOL = os.environ.get("TORCHPRUNER_B200_FUSE_POOL", "1") != "0"
F16_TARGET = 64.0    # per-tensor power-of-two scaling brings max|x| to ~64 (fp16 max is 65504)


def _pow2_scale(max_abs):
    """Power of two s such that max_abs * s is within a factor sqrt(2) of F16_TARGET."""
    if not (max_abs > 0.0) or not math.isfinite(max_abs):
        return 1.0
    return float(2.0 ** round(math.log2(F16_TARGET / max_abs)))
INCREMENTAL_MAX_BYTES = 8 << 30   # per-channel contributions of the first layer, per batch


class _Unsupported(Exception):
    pass


CHANNEL_GRANULE = 64   # one 128-byte k-block of fp16 operands; a multiple of the TF32 one (32)


def _pad(n):
    return -(-int(n) // CHANNEL_GRANULE) * CHANNEL_GRANULE


def _pad_to(t, dim, size):
    """``t`` zero-padded along ``dim`` to ``size`` entries."""
    extra = size - t.shape[dim]
    if extra <= 0:
        return t
    shape = list(t.shape)
    shape[dim] = extra
    return torch.cat([t, t.new_zeros(shape)], dim)


# model -> {(id(module), activation shape): (signature, plan or None, info)}.  A plan (traced,
# weight-prepared and verified against forward_partial) stays valid exactly as long as no
# parameter or buffer of the model has been written, replaced or resized and no BatchNorm/Dropout
# changed mode — which is what the signature captures (in-place writes bump `_version`, pruning
# replaces `.data`).
_PLANS = weakref.WeakKeyDictionary()


_MODE_MODULES = (nn.modules.batchnorm._BatchNorm, nn.modules.dropout._DropoutNd)


def _signature(model):
    # one walk over the modules' own parameter / buffer dicts (named_parameters() builds and
    # de-duplicates dotted names on every call: 0.48 ms against 0.21 ms for VGG16, once per layer
    # run with the GPU idle); a tensor shared by two modules is simply listed twice
    sig = []
    for m in model.modules():
        for t in m._parameters.values():
            if t is not None:
                sig.append((id(t), t._version, t.data_ptr(), tuple(t.shape)))
        for t in m._buffers.values():
            if t is not None:
                sig.append((id(t), t._version, t.data_ptr(), tuple(t.shape)))
        if isinstance(m, _MODE_MODULES):
            sig.append((id(m), m.training))
    return tuple(sig)


# Prepared (permuted, folded, TF32-split) weights of one layer, shared by every tail that runs
# through it: key = identities and versions of the tensors they were made from.
_LAYERS = weakref.WeakKeyDictionary()


def _tensor_key(*tensors):
    return tuple(None if t is None else (id(t), t._version, t.data_ptr(), tuple(t.shape))
                 for t in tensors)


def _digest(tensors):
    """Order-independent integer checksum of the bit patterns of ``tensors`` (left on the
    device): per dtype size, one concatenation and one integer sum — three launches for a whole
    tail instead of one per tensor."""
    groups = {}
    for t in tensors:
        groups.setdefault(t.element_size(), []).append(t.detach().reshape(-1))
    sums = []
    for size in sorted(groups):
        flat = groups[size][0] if len(groups[size]) == 1 else torch.cat(groups[size])
        view = {8: torch.int64, 4: torch.int32, 2: torch.int16}.get(size, torch.uint8)
        sums.append(flat.view(view).sum(dtype=torch.int64))
    return torch.stack(sums) if sums else None


def cached_plan(model, module, z, kind="tail", sig=None):
    """(found, plan, info) for this model state.

    The signature catches everything PyTorch itself can see: replaced or resized tensors
    (pruning), in-place writes that bump ``_version``, mode changes.  A write through ``.data``
    (``p.data.mul_(mask)``, ``p.data.copy_(w)`` in older masking / fine-tuning code) or from a
    raw kernel leaves all of that unchanged, and the plan holds its own folded, split copies of
    the weights — so a hit is additionally checked against a checksum of the bits of every
    tensor the plan was made from (one small reduction per tensor and one 8-byte-per-tensor
    comparison; the pipeline is idle at this point of a run anyway).  A mismatch drops the entry
    and the caller rebuilds and re-verifies."""
    key = (kind, id(module), tuple(z.shape[1:]))
    entry = _PLANS.get(model, {}).get(key)
    if entry is not None and entry[0] == (sig if sig is not None else _signature(model)):
        plan = entry[1]
        if plan is not None and plan.watched and \
                not torch.equal(_digest(plan.watched), plan.digest):
            forget_plans(model)
            return False, None, None
        return True, plan, entry[2]
    return False, None, None


# Activations the head plans have computed, per model and input batch: scoring the layers of a
# model one after the other on the same data (the reference's workflow: one `run` per layer, the
# same loader) asks for the same prefix of the network again and again — the head of layer l + 1
# is the head of layer l plus one block.  model -> {batch checksum: (signature, step keys,
# activations after each step)}, least recently used batches dropped beyond ACT_CACHE_BYTES.
_ACTS = weakref.WeakKeyDictionary()
ACT_CACHE_BYTES = int(os.environ.get("TORCHPRUNER_B200_ACT_CACHE_MB", "2048")) << 20


def _batch_checksum(x):
    """Position-dependent checksum of a batch's bits (two 64-bit sums, one weighted by the odd
    numbers 1, 3, 5, ...): the same examples in another order — a shuffling loader over a small
    dataset — must not look like the same batch.  One device->host copy of 16 bytes."""
    bits = x.detach().reshape(-1).view(torch.int32).to(torch.int64)
    weights = torch.arange(1, 2 * bits.numel(), 2, dtype=torch.int64, device=x.device)
    both = torch.stack([bits.sum(), (bits * weights).sum()])   # integer overflow wraps: fine
    return tuple(both.tolist())


def cached_activation(model, sig, head, x, compute):
    """The head's output for the batch ``x``, continuing from the longest prefix of its steps
    whose outputs for this very batch (same bits: the batch is identified by a checksum of its
    contents) are cached.  ``compute(z, start)`` returns the outputs of steps ``start`` .. last
    given the output ``z`` of step ``start - 1``.  Entries are only valid under the model
    signature they were made with, and ``forget_plans`` (which every detected change of a weight
    goes through) drops them."""
    if ACT_CACHE_BYTES <= 0:
        return compute(x, 0)[-1]
    chain = tuple(e["key"] for e in head.exact)
    batch = (tuple(x.shape), _batch_checksum(x))
    entries = _ACTS.setdefault(model, {})
    entry = entries.pop(batch, None)
    acts, start = [], 0
    if entry is not None and entry[0] == sig:
        while start < min(len(chain), len(entry[1])) and chain[start] == entry[1][start]:
            start += 1
        if start == len(chain):   # everything asked for is there (the entry may go further)
            entries[batch] = entry
            return entry[2][start - 1]
        acts = list(entry[2][:start])
    acts += compute(acts[-1] if acts else x, start)
    entries[batch] = (sig, chain, acts)   # most recently used last
    total = 0
    for key in reversed(list(entries)):
        total += sum(a.numel() * a.element_size() for a in entries[key][2])
        if total > ACT_CACHE_BYTES:
            del entries[key]
    return acts[len(chain) - 1]


def note_agreement(model, module, z, kind, key, outcome):
    """Remember, with the cached plan, what the ranks of a sharded job agreed on for it (see
    shapley_values.agree_on_plan): the next run of the same job applies it without a collective.
    Every rank stores the same outcome, so every rank skips — or does not skip — together."""
    entries = _PLANS.get(model, {})
    slot = (kind, id(module), tuple(z.shape[1:]))
    entry = entries.get(slot)
    if entry is None:
        return
    info = dict(entry[2] or {})
    agreed = dict(info.get("agreed", {}))
    agreed[key] = tuple(bool(v) for v in outcome)
    info["agreed"] = agreed
    entries[slot] = (entry[0], entry[1], info)


def forget_plans(model=None):
    """Drop every cached plan and prepared weight copy of ``model`` (of all models when None)."""
    if model is None:
        _PLANS.clear()
        _LAYERS.clear()
        _ACTS.clear()
        _TWINS.clear()
    else:
        _PLANS.pop(model, None)
        _LAYERS.pop(model, None)
        _ACTS.pop(model, None)
        _TWINS.pop(model, None)


def remember_plan(model, module, z, plan, info, kind="tail", sig=None):
    _PLANS.setdefault(model, {})[(kind, id(module), tuple(z.shape[1:]))] = \
        (sig if sig is not None else _signature(model), plan, info)


def model_signature(model):
    """What ``cached_plan`` compares (pass it as ``sig`` to look several plans up at the cost of
    one walk over the model's tensors)."""
    return _signature(model)


def _trace_chain(model, z, module, head=False):
    """Leaf modules executed by ``model.forward_partial(z, from_module=module)`` (or, for
    ``head``, by ``model.forward_partial(z, to_module=module)``), in order, with the tensors they
    consumed and produced."""
    calls = []
    handles = []

    def hook(mod, inputs, output):
        if len(inputs) != 1 or not isinstance(inputs[0], torch.Tensor) or \
                not isinstance(output, torch.Tensor):
            raise _Unsupported(f"{type(mod).__name__} is not a single-tensor layer")
        calls.append((mod, inputs[0], output))

    for m in model.modules():
        if len(list(m.children())) == 0:
            handles.append(m.register_forward_hook(hook))
    try:
        with torch.no_grad():
            if head:
                logits = model.forward_partial(z, to_module=module)
            else:
                logits = model.forward_partial(z, to_module=None, from_module=module)
    finally:
        for h in handles:
            h.remove()
    return calls, logits


def _is_pool2(m):
    def pair(v):
        return tuple(v) if isinstance(v, (tuple, list)) else (v, v)
    return (isinstance(m, nn.MaxPool2d) and pair(m.kernel_size) == (2, 2) and
            pair(m.stride if m.stride is not None else m.kernel_size) == (2, 2) and
            pair(m.padding) == (0, 0) and pair(m.dilation) == (1, 1) and not m.ceil_mode and
            not m.return_indices)


class NativeTail:
    """Execution plan for one evaluation module of one model."""

    def __init__(self, steps, pool_first, in_shape, n_out, first=None):
        self.steps = steps            # list of dicts
        self.pool_first = pool_first
        self.in_shape = in_shape      # (C, H, W) of the evaluation module's output
        self.n_out = n_out
        self.flops_per_example = sum(s["flops"] for s in steps)
        # Incremental first layer (csrc/incremental.cu): the masks of one permutation are
        # nested and the first layer is linear in the activation, so its output for mask k is
        # the previous one minus the contribution of one channel.
        self.first = first            # dict or None
        self.incremental = False      # decided per batch in prepare_batch()
        # fp16x3 flavour: every step has fp16 weights and per-tensor scales, and fits the kernel's
        # limits.  use_f16 may be switched off at run time (verification failure / overflow).
        self.f16_ok = all(st.get("w16_hi") is not None for st in steps)
        self.use_f16 = self.f16_ok and os.environ.get("TORCHPRUNER_B200_F16", "1") != "0"
        self._T = self._s_all = self._base = self._zc = None
        self._perm = None
        self._pos = 0
        self.watched, self.digest = [], None   # see cached_plan()
        self.head_out = None   # (C, H, W) when this plan is a head (input -> evaluation module)
        self.exact = []        # head plans: per-layer operands of the exact form (activation())

    # ------------------------------------------------------------------ construction
    @classmethod
    def build(cls, model, module, z, head=False):
        """Plan for the layers after ``module`` given a sample activation ``z`` — or, with
        ``head``, for the layers from the model input ``z`` up to and including ``module`` — or
        None when they are not a supported chain.  Never raises for an unsupported model."""
        try:
            return cls._build(model, module, z, head)
        except _Unsupported:
            return None
        except Exception as exc:   # a model this tracer misreads must not break scoring
            logging.getLogger(__name__).debug("native tail not built: %r", exc)
            return None

    @classmethod
    def _build(cls, model, module, z, head=False):
        if z.dim() not in (2, 4) or z.dtype != torch.float32:
            raise _Unsupported("activation rank/dtype")
        calls, logits = _trace_chain(model, z, module, head)
        if not calls or logits.dim() != (4 if head else 2):
            raise _Unsupported("empty chain or unexpected output rank")
        # must be a chain: each layer consumes the previous layer's output (possibly reshaped)
        prev = z
        for mod, inp, out in calls:
            same = inp is prev or (inp.data_ptr() == prev.data_ptr() and inp.numel() == prev.numel())
            if not same:
                raise _Unsupported("tail is not a simple chain")
            if mod.training and isinstance(mod, (nn.modules.batchnorm._BatchNorm,
                                                 nn.modules.dropout._DropoutNd)):
                raise _Unsupported("train-mode BatchNorm/Dropout")
            prev = out
        if prev is not logits and prev.data_ptr() != logits.data_ptr():
            raise _Unsupported("output is not the last layer's output")

        mods = [c[0] for c in calls]
        # C: channels of the activation as the model sees it; Cp: channels of the engine's NHWC
        # copy of it, zero-padded to the kernels' granule.  Padded weight rows/columns are zero
        # and padded outputs get scale 1 / shift 0, so they stay exactly 0 through ReLU and
        # pooling: any width (48 or 445 channels after pruning) runs on the same tiles.
        C = z.shape[1]
        Cp = _pad(C)
        H, W = (z.shape[2], z.shape[3]) if z.dim() == 4 else (1, 1)
        i = 0
        pool_first = False
        if mods and _is_pool2(mods[0]):
            if H % 2 or W % 2:
                raise _Unsupported("odd spatial size before pooling")
            pool_first, H, W, i = True, H // 2, W // 2, 1
        steps = []
        first = None
        flat = z.dim() == 2
        dev = z.device
        watched = []   # every model tensor the plan's own copies are derived from
        exact, pool_pending = [], False   # head plans: see activation()
        while i < len(mods):
            m = mods[i]
            if isinstance(m, (nn.Dropout, nn.Dropout2d, nn.Identity)):
                i += 1
                continue
            if isinstance(m, nn.Flatten):
                if m.start_dim != 1 or m.end_dim not in (-1, 3):
                    raise _Unsupported("Flatten dims")
                flat = True
                i += 1
                continue
            layer_call = i   # index of this layer's module in the trace (for calibration)
            if isinstance(m, nn.Conv2d):
                if flat or m.kernel_size != (3, 3) or m.stride != (1, 1) or m.padding != (1, 1) or \
                        m.dilation != (1, 1) or m.groups != 1 or m.padding_mode != "zeros" or \
                        m.weight.shape[1] != C:
                    raise _Unsupported("conv geometry")
                # sizes from the weight itself: the reference's pruner (and ours, for parity) slices
                # tensors but leaves out_channels / in_features & co. at their old values
                cout = m.weight.shape[0]
                w4 = _pad_to(_pad_to(m.weight.detach().float(), 1, Cp), 0, _pad(cout))
                w = w4.permute(0, 2, 3, 1).reshape(w4.shape[0], 9 * Cp)
                taps = 9
                bias = m.bias.detach().float() if m.bias is not None else torch.zeros(cout, device=dev)
            elif isinstance(m, nn.Linear):
                calls_in = calls[i][1]
                if calls_in.dim() != 2 or m.weight.shape[1] != C * H * W:
                    raise _Unsupported("linear geometry")
                cout = m.weight.shape[0]
                # NCHW flatten order -> NHWC flatten order of the padded activation; the layer that
                # produces the logits keeps its true width (its output is the unsplit fp32 result)
                produces_logits = not head and \
                    all(isinstance(r, (nn.Dropout, nn.Identity)) for r in mods[i + 1:])
                w = _pad_to(m.weight.detach().float().view(cout, C, H * W), 1, Cp)
                w = _pad_to(w, 0, cout if produces_logits else _pad(cout))
                w = w.permute(0, 2, 1).reshape(w.shape[0], -1)
                w4 = None
                C, Cp, H, W = C * H * W, Cp * H * W, 1, 1
                flat = True
                taps = 1
                bias = m.bias.detach().float() if m.bias is not None else torch.zeros(cout, device=dev)
            else:
                raise _Unsupported(f"unsupported layer {type(m).__name__}")
            if 128 % W or not ((128 // W) % H == 0 or H % (128 // W) == 0):
                raise _Unsupported("spatial size not tileable")
            cout_p = w.shape[0]
            scale = torch.ones(cout, device=dev)
            shift = bias.clone()
            scale64, shift64 = scale.double(), shift.double()
            bn_used = None
            i += 1
            if i < len(mods) and isinstance(mods[i], nn.BatchNorm2d) and taps == 9:
                bn = bn_used = mods[i]
                if not bn.track_running_stats or bn.running_mean is None:
                    raise _Unsupported("BatchNorm without running statistics")
                a = (bn.running_var.detach().double() + bn.eps).rsqrt()
                if bn.affine:
                    a = a * bn.weight.detach().double()
                shift = (bias.double() - bn.running_mean.detach().double()) * a
                if bn.affine:
                    shift = shift + bn.bias.detach().double()
                scale64, shift64 = a, shift
                scale, shift = a.float(), shift.float()
                i += 1
            relu = False   # activation spec: False, True (ReLU) or ("leaky", slope)
            if i < len(mods) and isinstance(mods[i], nn.ReLU):
                relu = True
                i += 1
            elif i < len(mods) and isinstance(mods[i], nn.LeakyReLU):
                relu = ("leaky", float(mods[i].negative_slope))
                i += 1
            pool = False
            if i < len(mods) and _is_pool2(mods[i]) and taps == 9:
                if H % 2 or W % 2:
                    raise _Unsupported("odd spatial size before pooling")
                pool = True
                i += 1
            # A 3x3 / pad 1 convolution on a 2x2 image touches 4 of its 9 taps per output pixel
            # (the other 5 fall in the padding): as an implicit GEMM 56 % of the MMAs multiply
            # zeros.  It is exactly a dense map from the 4*Cin inputs of the image to its 4*Cout
            # outputs, and NHWC [N,2,2,C] already is that [N, 4C] matrix, so the layer runs as a
            # linear one with the weight re-indexed (no data movement, 16/36 of the MMAs).
            scale = torch.cat([scale, torch.ones(cout_p - cout, device=dev)])
            shift = torch.cat([shift, torch.zeros(cout_p - cout, device=dev)])
            dense2x2 = taps == 9 and H == 2 and W == 2
            k_cin, k_cout, k_taps = Cp, cout_p, taps
            if dense2x2:
                wd = torch.zeros((2, 2, cout_p, 2, 2, Cp), device=dev)
                for h in range(2):
                    for ww in range(2):
                        for h2 in range(2):
                            for w2 in range(2):
                                wd[h, ww, :, h2, w2, :] = w4[:, :, h2 - h + 1, w2 - ww + 1]
                w = wd.reshape(4 * cout_p, 4 * Cp)
                scale, shift = scale.repeat(4), shift.repeat(4)
                k_cin, k_cout, k_taps = 4 * Cp, 4 * cout_p, 1
            watched += [t for t in (m.weight, m.bias) if t is not None]
            if bn_used is not None:
                watched += [t for t in (bn_used.weight, bn_used.bias, bn_used.running_mean,
                                        bn_used.running_var) if t is not None]
            key = (_tensor_key(m.weight, m.bias) + (
                _tensor_key(bn_used.weight, bn_used.bias, bn_used.running_mean, bn_used.running_var)
                if bn_used is not None else ()), w.shape[1], H * W if taps == 1 else 0, dense2x2)
            store = _LAYERS.setdefault(model, {})
            prep = store.get(id(m))
            if prep is None or prep[0] != key:
                w_hi, w_lo = _lib.tf32_split(w.contiguous())
                prep = (key, w_hi, w_lo, scale.contiguous(), shift.contiguous())
                store[id(m)] = prep
            _, w_hi, w_lo, scale_k, shift_k = prep[:5]
            # fp16x3 flavour of the same layer: fp16 pair of w * s_w, calibrated input scale s_in
            f16_fits = (k_cin % 64 == 0) if k_taps == 9 else (k_cin % 8 == 0)
            w16_hi = w16_lo = None
            s_w = s_in = 1.0
            if f16_fits:
                if len(prep) < 8:
                    s_w = _pow2_scale(float(w.abs().max()))
                    w16_hi, w16_lo = _lib.f16_split(w.contiguous(), s_w)
                    prep = prep[:5] + (w16_hi, w16_lo, s_w)
                    store[id(m)] = prep
                w16_hi, w16_lo, s_w = prep[5:8]
                s_in = _pow2_scale(float(calls[layer_call][1].detach().abs().max()))
            steps.append({"taps": k_taps, "cin": k_cin, "cout": k_cout, "H": H, "W": W, "relu": relu,
                          "pool": pool, "w_hi": w_hi, "w_lo": w_lo, "scale": scale_k,
                          "shift": shift_k, "out_view": (H, W, cout_p) if dense2x2 else None,
                          "w16_hi": w16_hi, "w16_lo": w16_lo, "s_in": s_in, "s_w": s_w,
                          "scale16": (scale_k / (s_in * s_w)).contiguous() if f16_fits else None,
                          "flops": 2 * k_cout * k_taps * k_cin * (1 if dense2x2 else H * W)})
            if dense2x2:   # per-channel vectors for the incremental form of this layer
                scale, shift = scale_k[:cout_p], shift_k[:cout_p]
            else:
                scale, shift = scale_k, shift_k
            if head:
                if taps != 9 or cout_p % 4:
                    raise _Unsupported("head layer is not a 3x3 convolution")
                # this layer in the exact form of csrc/incremental.cu: per-input-channel fp32
                # contributions summed in float64 (wt over the REAL input channels)
                c_in = m.weight.shape[1]
                exact.append({"wt": w4[:, :c_in].permute(1, 2, 3, 0).reshape(c_in, 9, cout_p).contiguous(),
                              "C1": cout_p, "cout": cout, "scale64": scale64, "shift64": shift64,
                              "relu": relu, "pool_before": pool_pending,
                              # what this step computes, for the activation cache: the tensors
                              # it was built from, its activation and whether it pools its input
                              "key": (_tensor_key(m.weight, m.bias),
                                      _tensor_key(bn_used.weight, bn_used.bias, bn_used.running_mean,
                                                  bn_used.running_var) if bn_used is not None else None,
                                      relu, pool_pending),
                              "unit": torch.ones(cout_p, device=dev),
                              "zero": torch.zeros(cout_p, device=dev)})
                pool_pending = pool
            if len(steps) == 1 and cout_p % 4 == 0 and not head:
                # contributions of the REAL channels of z (the permutation indexes those) to the
                # padded outputs of the layer
                if taps == 9:   # [Cout_p, Cin, 3, 3] -> [Cin, 9, Cout_p]
                    wt = w4[:, :z.shape[1]].permute(1, 2, 3, 0).reshape(z.shape[1], 9, cout_p)
                    first = {"mode": 0, "Hp": H, "Wp": W}
                else:           # [Cout, Cz*S] (NCHW flatten order) -> [Cz, S, Cout_p]
                    cz = z.shape[1]
                    sp = m.weight.shape[1] // cz
                    wt = _pad_to(m.weight.detach().float(), 0, cout_p).view(cout_p, cz, sp).permute(1, 2, 0)
                    first = {"mode": 1, "Hp": 1, "Wp": 1}
                first.update(wt=wt.contiguous(), C1=cout_p, scale=scale, shift=shift, relu=relu,
                             pool=pool)
            C, Cp = cout, cout_p
            if pool:
                H, W = H // 2, W // 2
        in_shape = (z.shape[1],) + ((z.shape[2], z.shape[3]) if z.dim() == 4 else (1, 1))
        if head:
            # ends at the evaluation module: a conv (+BN) (+activation) whose fp32 NHWC output,
            # cut back to its real channels, is the activation the scores are computed on
            if not exact or len(exact) != len(steps) or steps[-1]["pool"] or pool_first or \
                    tuple(logits.shape[1:]) != (C, H, W):
                raise _Unsupported("head does not end in a convolution block")
            plan = cls(steps, False, in_shape, C, None)
            plan.head_out = (C, H, W)
            plan.exact = exact
            plan.use_f16 = False   # the head never runs on the fp16 operand path
            plan.watched = watched
            plan.digest = _digest(watched)
            return plan
        if not steps or steps[-1]["relu"] or steps[-1]["pool"] or steps[-1]["cout"] != logits.shape[1] \
                or H * W != 1:
            raise _Unsupported("tail does not end in a linear read-out")
        if len(steps) < 2:
            first = None   # the only layer produces the logits: nothing to chain after it
        plan = cls(steps, pool_first, in_shape, logits.shape[1], first)
        plan.watched = watched
        plan.digest = _digest(watched)
        return plan

    # ------------------------------------------------------------------ execution
    def _first_out_geometry(self, B):
        """(pixels of the first layer's output, elements of its post-pool output) per mask."""
        f = self.first
        pixels = B * f["Hp"] * f["Wp"]
        return pixels, (pixels // 4 if f["pool"] else pixels) * f["C1"]

    def prepare_batch(self, z, allow_incremental=True):
        """Once per batch: per-channel contributions of the first layer and their float64 sum."""
        self.incremental = False
        if self.first is None or not allow_incremental:
            return
        f = self.first
        z4 = z if z.dim() == 4 else z.view(z.shape[0], z.shape[1], 1, 1)
        B, C = z4.shape[0], z4.shape[1]
        pixels, _ = self._first_out_geometry(B)
        M1 = pixels * f["C1"]
        if CONTRIB == "direct" or (CONTRIB != "stored" and C * M1 * 4 > INCREMENTAL_MAX_BYTES):
            if self._s_all is None or self._s_all.numel() < M1 or self._s_all.device != z.device:
                self._s_all = torch.empty(M1, dtype=torch.float64, device=z.device)
                self._base = torch.empty(M1, dtype=torch.float64, device=z.device)
            self._zc = _lib.channel_major(z4, self.pool_first)
            H, W = z4.shape[2], z4.shape[3]
            # spatial size of the layer's input: its output grid (conv) or all positions (dense)
            self._zgeom = (B, C) + ((H // 2, W // 2) if self.pool_first else (H, W))
            self._s_all[:M1].zero_()
            _lib.prefix_contrib_direct(self._zc, f["wt"], None, 0, C, self._s_all, f["scale"],
                                       f["shift"], *self._zgeom, f["mode"], False, False, sign=1.0)
            self._T = None
            self._M1, self._pixels = M1, pixels
            self.incremental = True
            return
        self._zc = None
        if C * M1 * 4 > INCREMENTAL_MAX_BYTES:
            return
        if self._T is None or self._T.numel() < C * M1 or self._T.device != z.device:
            self._T = torch.empty(C * M1, dtype=torch.float32, device=z.device)
            self._s_all = torch.empty(M1, dtype=torch.float64, device=z.device)
            self._base = torch.empty(M1, dtype=torch.float64, device=z.device)
        _lib.channel_contrib(z4, f["wt"], self.pool_first, f["mode"], self._T)
        _lib.contrib_sum(self._T, C, M1, self._s_all)
        self._M1, self._pixels = M1, pixels
        self.incremental = True

    def release(self):
        """Drop the per-batch scratch (contributions can be several GB); the plan itself, which is
        cached across runs, keeps only the prepared weights."""
        self._T = self._s_all = self._base = self._perm = self._zc = None
        self.incremental = False

    def begin(self, perm, k_start):
        """Start (or restart) a run of consecutive prefixes of permutation `perm` at k_start."""
        self._perm, self._pos = perm, int(k_start)
        if not self.incremental:
            return
        f = self.first
        base = self._base[:self._M1]
        base.copy_(self._s_all[:self._M1])
        if k_start > 0:   # a sharded rank starts mid-permutation: skip the prefix without output
            if self._zc is not None:
                _lib.prefix_contrib_direct(self._zc, f["wt"], perm, 0, int(k_start), base, f["scale"],
                                           f["shift"], *self._zgeom, f["mode"], f["pool"], f["relu"])
            else:
                _lib.prefix_contrib_stack(self._T, perm, 0, int(k_start), base, f["scale"],
                                          f["shift"], self._pixels, f["C1"], f["Hp"], f["Wp"],
                                          f["pool"], f["relu"])

    @property
    def stack_dtype(self):
        return torch.float16 if self.use_f16 else torch.float32

    def stack_elems(self, B, P):
        """Floats (per hi / lo buffer) of the stacked input of the chained layers for P masks."""
        if self.first is not None:
            direct = self._direct_stack_elems(B, P)
            return max(direct, P * self._first_out_geometry(B)[1])
        return self._direct_stack_elems(B, P)

    def _direct_stack_elems(self, B, P):
        C, H, W = self.in_shape
        if self.pool_first:
            H, W = H // 2, W // 2
        return P * B * _pad(C) * H * W

    def logits(self, z, inv_perm, k0, P, buf_hi, buf_lo):
        """[P*B, n_out] fp32 logits of the P prefix masks k0 .. k0+P-1 of the current
        permutation.  In incremental mode calls must walk the prefixes consecutively from the
        position given to begin()."""
        B = z.shape[0]
        f16 = self.use_f16
        if buf_hi.dtype != self.stack_dtype:
            raise RuntimeError("stack buffers were allocated for the other operand format")
        if self.incremental:
            if k0 != self._pos:
                raise RuntimeError("incremental tail evaluated out of order")
            f = self.first
            if self._zc is not None:
                m_out = _lib.prefix_contrib_direct(
                    self._zc, f["wt"], self._perm, k0, P, self._base[:self._M1], f["scale"],
                    f["shift"], *self._zgeom, f["mode"], f["pool"], f["relu"], out_hi=buf_hi,
                    out_lo=buf_lo, out_scale=self.steps[1]["s_in"] if f16 else 1.0)
            else:
                m_out = _lib.prefix_contrib_stack(
                    self._T, self._perm, k0, P, self._base[:self._M1], f["scale"], f["shift"],
                    self._pixels, f["C1"], f["Hp"], f["Wp"], f["pool"], f["relu"], buf_hi, buf_lo,
                    out_scale=self.steps[1]["s_in"] if f16 else 1.0)
            self._pos += P
            Ho, Wo = (f["Hp"] // 2, f["Wp"] // 2) if f["pool"] else (f["Hp"], f["Wp"])
            shape = (P * B, Ho, Wo, f["C1"])
            x_hi, x_lo = buf_hi[:P * m_out].view(shape), buf_lo[:P * m_out].view(shape)
            steps = self.steps[1:]
        else:
            z4 = z if z.dim() == 4 else z.view(z.shape[0], z.shape[1], 1, 1)
            x_hi, x_lo = _lib.mask_stack_nhwc(z4, inv_perm, k0, P, self.pool_first, buf_hi, buf_lo,
                                              scale=self.steps[0]["s_in"] if f16 else 1.0,
                                              c_pad=_pad(z4.shape[1]))
            steps = self.steps
        last = len(steps) - 1
        for idx, s in enumerate(steps):
            n = x_hi.shape[0]
            if s["taps"] == 1 and x_hi.dim() == 4 and x_hi.shape[1] * x_hi.shape[2] != 1:
                x_hi = x_hi.reshape(n, 1, 1, -1)
                x_lo = x_lo.reshape(n, 1, 1, -1)
            split = not (idx == last or s["pool"])
            next_scale = steps[idx + 1]["s_in"] if (f16 and idx < last) else 1.0
            if f16 and s["pool"] and s["taps"] == 9 and s["out_view"] is None and idx < last \
                    and FUSE_POOL and _lib.conv_pool2_fusable(n, x_hi.shape[1], x_hi.shape[2],
                                                              x_hi.shape[3], s["w16_hi"].shape[0]):
                # conv + BN + ReLU + MaxPool2d(2) as one kernel (window maximum kept in TMEM)
                x_hi, x_lo = _lib.conv_tf32x3(x_hi, x_lo, s["w16_hi"], s["w16_lo"], 9, s["scale16"],
                                              s["shift"], s["relu"], split=True,
                                              out_scale=next_scale, pool=True)
                continue
            if f16 and s["pool"] and s["out_view"] is not None and idx < last and FUSE_POOL \
                    and s["out_view"][2] % 64 == 0:
                # dense 2x2 layer + MaxPool2d(2): the four output pixels are four groups of output
                # features, maximum kept in TMEM (scale/shift are the per-channel vectors x 4)
                cp = s["out_view"][2]
                x_hi, x_lo = _lib.linear_pool4(x_hi, x_lo, s["w16_hi"], s["w16_lo"],
                                               s["scale16"][:cp], s["shift"][:cp], s["relu"],
                                               out_scale=next_scale)
                continue
            if f16:
                y_hi, y_lo = _lib.conv_tf32x3(x_hi, x_lo, s["w16_hi"], s["w16_lo"], s["taps"],
                                              s["scale16"], s["shift"], s["relu"], split=split,
                                              out_scale=next_scale)
            else:
                y_hi, y_lo = _lib.conv_tf32x3(x_hi, x_lo, s["w_hi"], s["w_lo"], s["taps"],
                                              s["scale"], s["shift"], s["relu"], split=split)
            if s["out_view"] is not None:   # dense form of a conv on a 2x2 image: back to NHWC
                y_hi = y_hi.view((n,) + s["out_view"])
                y_lo = y_lo.view((n,) + s["out_view"]) if y_lo is not None else None
            if s["pool"]:
                y_hi, y_lo = _lib.maxpool2_split_nhwc(y_hi, half=f16, scale=next_scale)
            x_hi, x_lo = y_hi, y_lo
        if self.head_out is not None:
            return x_hi   # [N, H, W, Cout padded] fp32
        return x_hi.view(x_hi.shape[0], self.n_out)

    def activation(self, x):
        """Head plans: the evaluation module's output for the input batch ``x`` as a contiguous
        NCHW fp32 tensor.

        Every convolution runs in the exact form the incremental first layer uses
        (``tp_channel_major`` + ``tp_prefix_contrib_direct`` with all channels added): fp32
        products, the 9 taps of one input channel summed in fp32, the channels summed in
        FLOAT64; BatchNorm (folded in float64) and the activation are applied to that float64
        sum, which is rounded to fp32 once per layer — like the reference's fp32 activations
        between layers, but without its layer-internal rounding.  This pass runs once per batch
        on B examples (the masked evaluations run 25*C times on the tensor-core engine), so its
        cost is noise, and its accuracy sets the floor for every score: with the model's own
        fp32 forward on the GPU (cuDNN) supplying this activation, VGG16 Shapley tables were
        0.6-1.3e-4 of the largest score from a float64 evaluation; with the tensor-core engine's
        22-bit operands 3-7e-5; with this, 0.3-3e-5, the reference's own CPU fp32 distance
        (profiles/r02_parity_sources.txt)."""
        return self.run_steps(x, 0)[-1]

    def run_steps(self, cur, start, stop=None):
        """Outputs of the head's steps ``start`` .. ``stop - 1`` (default: the last; each a
        contiguous NCHW fp32 tensor), given the output of step ``start - 1`` (the input batch for
        ``start == 0``)."""
        outs = []
        for e in self.exact[start:stop]:
            zc = _lib.channel_major(cur, e["pool_before"])
            B, C = cur.shape[0], cur.shape[1]
            H, W = (cur.shape[2] // 2, cur.shape[3] // 2) if e["pool_before"] else \
                (cur.shape[2], cur.shape[3])
            acc = torch.zeros(B * H * W * e["C1"], dtype=torch.float64, device=cur.device)
            _lib.prefix_contrib_direct(zc, e["wt"], None, 0, C, acc, e["unit"], e["zero"], B, C, H, W,
                                       0, False, False, sign=1.0)
            y = acc.view(B, H, W, e["C1"])[..., :e["cout"]] * e["scale64"] + e["shift64"]
            if e["relu"] is True:
                y = torch.where(y < 0, torch.zeros_like(y), y)   # NaN-propagating, like ATen's
            elif e["relu"]:
                y = torch.where(y < 0, y * e["relu"][1], y)
            cur = y.permute(0, 3, 1, 2).float().contiguous()
            outs.append(cur)
        return outs

    def verify_head(self, model, module, x):
        """(ok, error): the head's activation against ``forward_partial(x, to_module=module)``,
        relative to the largest activation — with the same arbitration by a float64 copy of the
        model as ``verify`` when the difference is within what the library forward itself is off
        by (its fp32 convolutions were measured up to ~1e-5 from the exact result per layer)."""
        with torch.no_grad():
            ref = model.forward_partial(x, to_module=module)
        mine = self.activation(x)
        if ref.shape != mine.shape:
            return False, float("inf")
        scale = float(ref.abs().max().clamp_min(1e-30))
        worst = float((ref - mine).abs().max()) / scale
        if worst <= VERIFY_RTOL:
            return True, worst
        if worst > VERIFY_ARBITRATE_BELOW:
            return False, worst
        rows = _arbitration_rows(x.shape[0], x.shape[0], x.device)
        ref64 = _float64_forward(model, module, x[rows], head=True)
        if ref64 is None or ref64.shape != ref[rows].shape:
            return False, worst
        lib_err = float((ref[rows].double() - ref64).abs().max()) / scale
        mine_err = float((mine[rows].double() - ref64).abs().max()) / scale
        return mine_err <= max(VERIFY_RTOL / 10, lib_err), mine_err

    def unmasked_logits(self, z, inv_perm, buf_hi, buf_lo):
        """[B, n_out] logits with no channel removed (prefix index -1), through the direct form
        of the first layer."""
        saved = self.incremental
        self.incremental = False
        try:
            return self.logits(z, inv_perm, -1, 1, buf_hi, buf_lo).clone()
        finally:
            self.incremental = saved

    def verify(self, model, module, z, perm, inv_perm, k0, P, buf_hi, buf_lo):
        """True iff the plan reproduces ``forward_partial`` on P real masks of this run (both the
        incremental and the direct form of the first layer are checked).

        First against the model's own fp32 forward, to VERIFY_RTOL of the largest logit.  That
        forward is library code with rounding of its own (cuDNN's fp32 convolution algorithms
        differ from the exact result by up to ~2e-5 here, forty times the engine's own error),
        so a difference slightly above the tolerance does not say which side is off: it is then
        arbitrated by a float64 copy of the model, and the plan is accepted when it is at least as
        close to that as the fp32 library forward is.  Structural mistakes are O(1) and fail both."""
        stacked = _lib.prefix_mask_stack(z, inv_perm, k0, P)
        with torch.no_grad():
            ref = model.forward_partial(stacked, to_module=None, from_module=module)
        scale = float(ref.abs().max().clamp_min(1e-30))
        worst = 0.0
        was_incremental = self.incremental
        outs = []
        for mode in ([True, False] if was_incremental else [False]):
            self.incremental = mode
            self.begin(perm, k0)
            mine = self.logits(z, inv_perm, k0, P, buf_hi, buf_lo)
            if ref.shape != mine.shape:
                self.incremental = was_incremental
                return False, float("inf")
            outs.append(mine.clone())
            worst = max(worst, float((ref - mine).abs().max()) / scale)
        self.incremental = was_incremental
        if worst <= VERIFY_RTOL:
            return True, worst
        if worst > VERIFY_ARBITRATE_BELOW:
            return False, worst
        rows = _arbitration_rows(stacked.shape[0], z.shape[0], z.device)
        ref64 = _float64_forward(model, module, stacked[rows])
        if ref64 is None or ref64.shape != ref[rows].shape:
            return False, worst
        lib_err = float((ref[rows].double() - ref64).abs().max()) / scale
        mine_err = max(float((o[rows].double() - ref64).abs().max()) for o in outs) / scale
        return mine_err <= max(VERIFY_RTOL / 10, lib_err), mine_err


def head_plan(model, module, x, sig=None):
    """(plan or None, info): a verified plan for the layers from the model input ``x`` to the
    evaluation module ``module`` (a chain of 3x3 convolutions / BatchNorm / activations /
    MaxPool2d(2)), from the cache or built and verified now.  See ``NativeTail.activation`` for
    why the library computes the cached activation itself when it can."""
    if not (x.dim() == 4 and x.dtype == torch.float32 and x.is_cuda):
        return None, {}
    found, plan, info = cached_plan(model, module, x, kind="head", sig=sig)
    if found:
        return plan, dict(info or {}, cached=True)
    plan = NativeTail.build(model, module, x, head=True)
    info = {"verified": False}
    if plan is not None:
        ok, err = plan.verify_head(model, module, x)
        info = {"verified": ok, "verify_rel_err": err}
        if not ok:
            plan = None
    remember_plan(model, module, x, plan, info, kind="head", sig=sig)
    return plan, info


ARBITRATION_EXAMPLES = 32   # per mask: float64 library convolutions are ~25x slower than fp32


def _arbitration_rows(n_rows, batch, device):
    """Rows of a stacked batch ([masks * batch, ...]) the float64 arbiter looks at: the first
    ARBITRATION_EXAMPLES examples of every mask.  Which side of a rounding-sized disagreement is
    right does not depend on how many examples are asked."""
    per = min(batch, ARBITRATION_EXAMPLES)
    starts = torch.arange(0, n_rows, batch, device=device)
    return (starts[:, None] + torch.arange(per, device=device)[None, :]).reshape(-1)


# float64 copies of models, kept while the model's signature stands (one deep copy per model
# state instead of one per arbitration)
_TWINS = weakref.WeakKeyDictionary()


def _float64_forward(model, module, stacked, head=False):
    """``forward_partial`` of a float64 deep copy of the model (None when the model cannot be
    copied or run that way)."""
    import copy
    try:
        name = next(n for n, m in model.named_modules() if m is module)
        sig = _signature(model)
        cached = _TWINS.get(model)
        if cached is None or cached[0] != sig:
            cached = (sig, copy.deepcopy(model).double())
            _TWINS[model] = cached
        twin = cached[1]
        twin_module = twin.get_submodule(name) if name else twin
        with torch.no_grad():
            if head:
                return twin.forward_partial(stacked.double(), to_module=twin_module)
            return twin.forward_partial(stacked.double(), to_module=None, from_module=twin_module)
    except Exception:
        return None
