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
