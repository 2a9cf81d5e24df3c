import numpy as np

from .. import attributions as _base


class RandomAttributionMetric(_base._AttributionMetric):
    """Uniform random scores from the *global legacy* numpy RNG (reference
    attributions/methods/random.py).  Stays on the host on purpose: one ``np.random.random``
    call of ``n`` doubles, so a seeded score -> prune -> score loop consumes the RNG stream
    exactly as the reference does."""

    def run(self, module, **kwargs):
        module = super().run(module, **kwargs)
        return np.random.random((module.weight.shape[0],))

    def find_evaluation_module(self, module, find_best_evaluation_module=False):
        return module  # needs a module that owns weights
