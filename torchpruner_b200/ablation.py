"""Ablation / robustness curve: remove the units of one layer one by one, in a given order, and
record the loss and accuracy after each removal.

This is the step that follows scoring in the reference's workflow (experiments notebook
"CIFAR-10 - VGG16 - Layerwise robustness", cell 8, lines 1261-1282): there, every removal is an
``index_fill_``, a ``forward_partial``, a mean loss, an argmax and two ``.item()`` host syncs.
It is the Shapley prefix evaluation with ONE fixed order, so the same machinery applies: masks are
built P at a time (``tp_prefix_mask_stack`` / ``tp_mask_stack_nhwc``), the layers after the
evaluation module run once per stack (tcgen05 engine when the tail is recognised, the model's
``forward_partial`` otherwise), and ``tp_ce_mean_acc`` reduces each masked batch to its mean
cross entropy and top-1 count on the device.  One device->host copy at the end.
"""
import numpy as np
import torch
import torch.nn.functional as F

from . import _lib
from .attributions.native_tail import (NativeTail, cached_activation, cached_plan, head_plan,
                                       model_signature, remember_plan)
from .utils import find_best_module_for_attributions


def ablation_curve(model, x, y, module, ranking, criterion=None, find_best_evaluation_module=True,
                   max_stack_bytes=512 << 20, max_stack=512, native_tail=True,
                   incremental_first=True):
    """Losses and accuracies after removing ``ranking[0]``, ``ranking[:2]``, ... of ``module``.

    x, y: one batch on the model's CUDA device (the notebook uses 1000 test images).
    ``criterion``: the reference protocol ``criterion(output, target, reduction=...)``; None or
    cross entropy selects the fused reduction kernel.  Returns ``(loss, acc)`` float32 arrays of
    ``len(ranking)`` entries — the values the notebook appends per removed unit.
    """
    module_nl = find_best_module_for_attributions(model, module) if find_best_evaluation_module \
        else module
    ranking = np.asarray(ranking, dtype=np.int64)
    with torch.no_grad():
        stack_safe = not any(m.training and isinstance(
            m, (torch.nn.modules.batchnorm._BatchNorm, torch.nn.modules.dropout._DropoutNd))
            for m in model.modules())
        head = head_plan(model, module_nl, x)[0] if (native_tail and stack_safe) else None
        if head is not None:   # the activation itself, exactly (head_plan), cached per batch
            z = cached_activation(model, model_signature(model), head, x, head.run_steps)
        else:
            z = model.forward_partial(x, to_module=module_nl).detach().float().contiguous()
        B, C = z.shape[0], z.shape[1]
        if ranking.size and (ranking.min() < 0 or ranking.max() >= C):
            raise IndexError(f"ranking refers to units outside [0, {C})")
        # position of each channel in the removal order; channels not listed are never removed
        inv = np.full(C, np.iinfo(np.int32).max, dtype=np.int32)
        inv[ranking[::-1]] = np.arange(ranking.size, dtype=np.int32)[::-1]  # first occurrence wins
        inv_dev = torch.from_numpy(inv).to(z.device)
        n = int(ranking.size)
        fused = criterion is None or criterion is F.cross_entropy
        if not fused:
            probe = model.forward_partial(z, from_module=module_nl)
            fused = (probe.dim() == 2 and y.dim() == 1 and y.dtype == torch.int64 and
                     torch.allclose(criterion(probe, y, reduction="none"),
                                    F.cross_entropy(probe, y, reduction="none"), rtol=2e-6, atol=2e-6))
        native = None
        if native_tail and stack_safe and n:
            found, native, info = cached_plan(model, module_nl, z)
            if not found:
                native = NativeTail.build(model, module_nl, z)
        ranking_dev = torch.from_numpy(ranking).to(z.device)
        if native is not None:
            native.prepare_batch(z, allow_incremental=incremental_first)
            per_copy = (4 if native.use_f16 else 8) * native.stack_elems(B, 1)
            P_max = int(max(1, min(max_stack, n, max_stack_bytes // per_copy)))
            buf_hi = torch.empty(native.stack_elems(B, max(P_max, min(2, n))),
                                 dtype=native.stack_dtype, device=z.device)
            buf_lo = torch.empty_like(buf_hi)
            if not found:
                ok, err = native.verify(model, module_nl, z, ranking_dev, inv_dev, 0, min(2, n),
                                        buf_hi, buf_lo)
                if not ok and native.use_f16:
                    native.use_f16 = False
                    buf_hi = torch.empty(native.stack_elems(B, max(P_max, min(2, n))),
                                         dtype=torch.float32, device=z.device)
                    buf_lo = torch.empty_like(buf_hi)
                    ok, err = native.verify(model, module_nl, z, ranking_dev, inv_dev, 0, min(2, n),
                                            buf_hi, buf_lo)
                info = {"verified": ok, "verify_rel_err": err}
                if not ok:
                    native = None
        if native_tail and stack_safe and n and not found:
            remember_plan(model, module_nl, z, native, info or {"verified": False})
        if native is None:
            P_max = int(max(1, min(max_stack, max(n, 1), max_stack_bytes // max(1, z.numel() * 4))))
            if not stack_safe:
                P_max = 1
            stack = torch.empty((P_max * B,) + tuple(z.shape[1:]), dtype=torch.float32, device=z.device)

        losses, counts = [], []
        if native is not None:
            native.begin(ranking_dev, 0)
        for k0 in range(0, n, P_max):
            P = min(P_max, n - k0)
            if native is not None:
                out = native.logits(z, inv_dev, k0, P, buf_hi, buf_lo)
            else:
                masked = _lib.prefix_mask_stack(z, inv_dev, k0, P, out=stack)[:P * B]
                out = model.forward_partial(masked, from_module=module_nl)
            if fused:
                ml, ct = _lib.ce_mean_acc(out.float().contiguous(), y, P, B)
            else:
                per = criterion(out, y.repeat((P,) + (1,) * (y.dim() - 1)), reduction="none")
                ml = per.reshape(P, -1).float().mean(1)
                ct = (out.argmax(dim=1).view(P, B) == y.view(1, B)).sum(1).to(torch.int32)
            losses.append(ml)
            counts.append(ct)
        used_f16 = native is not None and native.use_f16
        if native is not None:
            native.release()
        if not losses:
            return np.zeros(0, np.float32), np.zeros(0, np.float32)
        loss = torch.cat(losses).cpu().numpy()
        if used_f16 and not np.isfinite(loss).all():
            # an activation left fp16's range under some mask (NaN reaches the loss, see
            # common.cuh nan_relu): the cached plan drops to its TF32 flavour and the curve is redone
            native.use_f16 = False
            return ablation_curve(model, x, y, module, ranking, criterion,
                                  find_best_evaluation_module, max_stack_bytes, max_stack,
                                  native_tail, incremental_first)
        acc = (torch.cat(counts).float() / len(y)).cpu().numpy()  # `.sum().float() / len(y_test)`
        return loss, acc
