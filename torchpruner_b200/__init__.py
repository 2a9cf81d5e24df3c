"""torchpruner_b200 — B200 (sm_100a) implementation of TorchPruner's attribution-scoring and
slicing hot path behind the reference's Python API.

    from torchpruner_b200.attributions import TaylorAttributionMetric, ShapleyAttributionMetric
    from torchpruner_b200.pruner import Pruner
    from torchpruner_b200.utils import get_vgg_pruning_graph

or, to run existing TorchPruner code unchanged, ``torchpruner_b200.install_as_torchpruner()``
before it imports ``torchpruner.attributions`` / ``torchpruner.pruner`` / ``torchpruner.utils``.
"""
import sys

__version__ = "0.1.0"


def install_as_torchpruner():
    """Register this package under the reference's import paths."""
    from . import attributions, pruner, utils

    this = sys.modules[__name__]
    sys.modules["torchpruner"] = this
    sys.modules["torchpruner.attributions"] = attributions
    sys.modules["torchpruner.pruner"] = pruner
    sys.modules["torchpruner.utils"] = utils
    return this


def forget_plans(model=None):
    """Drop every cached execution plan, prepared weight copy, float64 twin and activation the
    library keeps for ``model`` (for all models when None).  Plans are re-validated on every use
    (tensor identities, version counters, a checksum of the weight bits), so this is only needed
    to release memory early."""
    from .attributions import native_tail

    native_tail.forget_plans(model)
