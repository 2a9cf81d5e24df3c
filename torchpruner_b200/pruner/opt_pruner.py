import torch

from .. import _lib


def sgd_momentum_buffer(optimizer, param):
    """The SGD momentum buffer that belongs to ``param``, or None.

    Reference opt_pruner.py:12-19 looks the parameter up in ``param_groups[0]`` only and then
    indexes ``optimizer.state_dict()["state"]`` by ``id(p)`` — which raises KeyError on every
    torch >= 1.6 because state_dict() is keyed by parameter position.  Same scope (SGD, first
    parameter group), same tensor, reached through ``optimizer.state[p]`` instead.
    """
    if not isinstance(optimizer, torch.optim.SGD):
        return None
    for p in optimizer.param_groups[0]["params"]:
        if p is param:
            buf = optimizer.state.get(p, {}).get("momentum_buffer", None)
            return buf if isinstance(buf, torch.Tensor) else None
    return None


def optimizer_state_tensors(optimizer, param):
    """Every per-parameter state tensor that has to be sliced together with ``param``.

    The reference covers ``torch.optim.SGD`` and its first parameter group only
    (opt_pruner.py:8-13).  That case is kept verbatim (``sgd_momentum_buffer``); in addition, for
    any optimizer and any parameter group, state tensors shaped like the parameter (Adam's
    ``exp_avg`` / ``exp_avg_sq``, RMSprop's ``square_avg``, SGD momentum in later groups, ...) are
    returned too, so a pruned model keeps training without a shape error.  Scalar
    state such as ``step`` is left alone.
    """
    if optimizer is None:
        return []
    found = []
    buf = sgd_momentum_buffer(optimizer, param)
    if buf is not None:
        found.append(buf)
    state = optimizer.state.get(param, None)
    if state:
        for value in state.values():
            if isinstance(value, torch.Tensor) and value.shape == param.shape and \
                    value.dim() > 0 and not any(value is f for f in found):
                found.append(value)
    return found


class OptimizerPruner:
    """Slice optimizer state along with a pruned parameter (reference pruner/opt_pruner.py)."""

    @staticmethod
    def prune(optimizer, param, axis, keep_indices, device):
        if isinstance(optimizer, torch.optim.SGD):
            return OptimizerPruner._prune_sgd(optimizer, param, axis, keep_indices, device)

    @staticmethod
    def _prune_sgd(optimizer, param, axis, keep_indices, device):
        from .pruner import _GatherPlan

        momentum = sgd_momentum_buffer(optimizer, param)
        if momentum is None:
            return
        keep = torch.as_tensor(keep_indices).detach().to("cpu", torch.int64).numpy()
        plan = _GatherPlan(device)
        plan.add(momentum, axis, keep)
        plan.execute()
