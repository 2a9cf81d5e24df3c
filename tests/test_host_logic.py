"""CPU tests of the host-side logic and of the C-ABI surface (no kernel is launched)."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.utils.data import DataLoader, TensorDataset

import torchpruner_b200
from tests import cases
from torchpruner_b200 import _lib, workloads
from torchpruner_b200.attributions import (APoZAttributionMetric, RandomAttributionMetric,
                                           SensitivityAttributionMetric, ShapleyAttributionMetric,
                                           TaylorAttributionMetric, WeightNormAttributionMetric)
from torchpruner_b200.attributions.methods.shapley_values import shard_spans
from torchpruner_b200.pruner import Pruner
from torchpruner_b200.utils import find_best_module_for_attributions, get_vgg_pruning_graph

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


# ------------------------------------------------------------------ C-ABI surface
def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "torchpruner_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(tp_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    declared = _declared_symbols()
    assert len(declared) >= 13
    for name in declared:
        assert hasattr(lib, name), f"{name} is declared in the header but not exported"
    assert sorted(_lib.exported_symbols()) == declared, "ctypes table and header disagree"
    assert lib.tp_version() >= 100
    assert ctypes.sizeof(_lib.GatherDesc) == 56


def test_argument_validation_without_a_gpu():
    lib = _lib.load()
    assert lib.tp_set_tuning(b"no_such_key", 1) != 0
    assert b"unknown key" in lib.tp_last_error()
    assert lib.tp_set_tuning(b"reduce_variant", 0) == 0
    assert lib.tp_taylor_reduce(None, None, -1, 1, 1, 0, None, None) != 0
    assert lib.tp_multi_gather(None, -1, None, 0, None) != 0
    # zero-sized problems are no-ops that never touch the device
    assert lib.tp_apoz_count(None, 0, 4, 9, None, None) == 0
    assert lib.tp_prefix_mask_stack(None, None, 0, 0, 4, 4, 4, None, None) == 0
    assert lib.tp_multi_gather(None, 0, None, 0, None) == 0


def test_no_cpu_fallback():
    z = torch.zeros(2, 3, 4, 4)
    with pytest.raises(_lib.TorchPrunerB200Error):
        _lib.apoz_count(z)
    with pytest.raises(_lib.TorchPrunerB200Error):
        _lib.taylor_reduce(z, z)
    model = nn.Sequential(nn.Linear(3, 2), nn.ReLU(), nn.Linear(2, 1))
    with pytest.raises(_lib.TorchPrunerB200Error):
        Pruner(model, (3,), "cpu").prune_parameter(model[0], "weight", [0], axis=0)
    assert list(model[0].weight.shape) == [2, 3], "a refused slice must leave the model untouched"


# ------------------------------------------------------------------ graph utilities
def test_graph_helpers_match_reference(capsys):
    with open(os.path.join(GOLDEN, "graph.txt")) as f:
        gold = f.read().splitlines()
    model, _ = workloads.build("vgg16_cifar")
    names = {m: n for n, m in model.named_modules()}
    lines = []
    for layer in model.conv_layers() + [m for m in model.classifier if isinstance(m, nn.Linear)]:
        lines.append(f"best {names[layer]} -> {names[find_best_module_for_attributions(model, layer)]}")
    graph = get_vgg_pruning_graph(model)
    assert len(graph) == 15
    for module, cascade in graph:
        lines.append(f"graph {names[module]} -> {','.join(names[c] for c in cascade)}")
    seq = nn.Sequential(nn.Linear(3, 2), nn.BatchNorm1d(2), nn.ReLU(), nn.Linear(2, 1))
    lines.append("seq0 -> %d" % list(seq).index(find_best_module_for_attributions(seq, seq[0])))
    tail = nn.Sequential(nn.Linear(3, 2), nn.ReLU())
    lines.append("tail0 -> %d" % list(tail).index(find_best_module_for_attributions(tail, tail[0])))
    assert lines == gold
    printed = capsys.readouterr().out
    assert "BatchNorm detected: shifting evaluation after" in printed
    assert "Activation detected: shifting evaluation after" in printed


def test_find_evaluation_module_per_metric():
    """torchpruner/tests/test_attributions.py:177-201."""
    x, y, _ = cases.max_model(2)
    loader = DataLoader(TensorDataset(x, y), batch_size=1)
    model = nn.Sequential(nn.Linear(3, 2), nn.BatchNorm1d(2), nn.ReLU(), nn.Linear(2, 1))
    for A in (TaylorAttributionMetric, SensitivityAttributionMetric, ShapleyAttributionMetric,
              APoZAttributionMetric):
        a = A(model, loader, F.mse_loss, "cpu")
        assert a.find_evaluation_module(model[0], find_best_evaluation_module=True) is model[2]
        assert a.find_evaluation_module(model[0]) is model[0]
    for A in (WeightNormAttributionMetric, RandomAttributionMetric):
        a = A(model, loader, F.mse_loss, "cpu")
        assert a.find_evaluation_module(model[0], find_best_evaluation_module=True) is model[0]


def test_module_not_in_model_logs_and_returns_module(caplog):
    model = nn.Sequential(nn.Linear(3, 2))
    stray = nn.Linear(2, 2)
    assert find_best_module_for_attributions(model, stray) is stray
    assert "not in model" in caplog.text


# ------------------------------------------------------------------ metric plumbing
def test_constructor_and_aggregation_contract():
    model = nn.Linear(2, 2)
    with pytest.raises(AssertionError):
        TaylorAttributionMetric(model, [], F.mse_loss, "cpu", reduction="median")
    table = np.arange(12, dtype=np.float32).reshape(3, 4)
    mk = lambda r: APoZAttributionMetric(model, [], F.mse_loss, "cpu", reduction=r)
    np.testing.assert_array_equal(mk("mean").aggregate_over_samples(table), table.mean(0))
    np.testing.assert_array_equal(mk("sum").aggregate_over_samples(table), table.sum(0))
    assert mk("none").aggregate_over_samples(table) is table
    np.testing.assert_array_equal(mk(lambda a: a.max(0)).aggregate_over_samples(table), table.max(0))
    with pytest.raises(AssertionError):
        mk("mean").run(nn.ReLU())


def test_random_metric_consumes_the_global_numpy_stream():
    model = nn.Sequential(nn.Linear(3, 5), nn.ReLU())
    np.random.seed(4)
    got = RandomAttributionMetric(model, [], F.mse_loss, "cpu").run(model[0])
    np.random.seed(4)
    np.testing.assert_array_equal(got, np.random.random((5,)))
    assert got.dtype == np.float64


def test_cudnn_flags_are_restored():
    before = (torch.backends.cudnn.deterministic, torch.backends.cudnn.benchmark)
    a = APoZAttributionMetric(nn.Linear(2, 2), [], F.mse_loss, "cpu")
    a.set_deterministic()
    assert torch.backends.cudnn.deterministic is True and torch.backends.cudnn.benchmark is False
    a.restore_deterministic()
    assert (torch.backends.cudnn.deterministic, torch.backends.cudnn.benchmark) == before


# ------------------------------------------------------------------ sharding
@pytest.mark.parametrize("samples,channels,world", [(25, 64, 8), (25, 512, 8), (5, 7, 3), (1, 4, 8),
                                                    (3, 10, 1), (2, 5, 4)])
def test_shard_spans_partition_the_work(samples, channels, world):
    seen = np.zeros((samples, channels), dtype=int)
    sizes, cuts = [], []
    for rank in range(world):
        spans = shard_spans(samples, channels, world, rank)
        count = 0
        for j, a, b in spans:
            assert 0 <= a < b <= channels and 0 <= j < samples
            seen[j, a:b] += 1
            count += b - a
        sizes.append(count)
        # whole permutations first (no halo, no start inside a permutation) ...
        whole = [s for s in spans if s[1] == 0 and s[2] == channels]
        assert len(whole) >= samples // world
        # ... and the spans that are cut from the left-over permutations are contiguous
        part = [s for s in spans if s not in whole[:samples // world]]
        flat = [j * channels + k for j, a, b in part for k in range(a, b)]
        assert flat == list(range(flat[0], flat[0] + len(flat))) if flat else True
        cuts.append(sum(1 for j, a, b in spans if a > 0))
    assert (seen == 1).all()
    assert max(sizes) - min(sizes) <= 1
    assert max(cuts) <= 1 if channels >= world else True   # at most one halo evaluation per rank


def test_shard_spans_of_the_named_job():
    """sv_samples=25 over 8 GPUs: three whole permutations per rank, the 25th cut in eight."""
    for rank in range(8):
        spans = shard_spans(25, 512, 8, rank)
        assert spans[:3] == [(3 * rank + i, 0, 512) for i in range(3)]
        assert spans[3:] == [(24, 64 * rank, 64 * rank + 64)]
    assert shard_spans(25, 64, 1, 0) == [(j, 0, 64) for j in range(25)]


def test_pruner_dropout_and_validation():
    model = nn.Sequential(nn.Linear(3, 5), nn.Dropout(0.5), nn.Linear(5, 1))
    p = Pruner(model, (3,), "cpu")
    p.prune_module(model[1], [0], direction="in", original_len=5.0)
    assert model[1].p == 0.5 * 4 / 5      # torchpruner/tests/test_pruner.py:162-176
    with pytest.raises(RuntimeError):
        p.prune_module(model[1], [0], direction="in")
    with pytest.raises(AssertionError):
        p.prune_module(model[1], [0], direction="out")
    with pytest.raises(AssertionError):
        p.prune_module(model[0], [0], direction="sideways")


def test_install_as_torchpruner():
    import sys

    saved = {k: sys.modules.get(k) for k in ("torchpruner", "torchpruner.attributions",
                                             "torchpruner.pruner", "torchpruner.utils")}
    try:
        torchpruner_b200.install_as_torchpruner()
        from torchpruner.attributions import ShapleyAttributionMetric as S  # noqa
        from torchpruner.pruner import Pruner as P  # noqa
        from torchpruner.utils import get_vgg_pruning_graph as G  # noqa
        assert S is ShapleyAttributionMetric and P is Pruner and G is get_vgg_pruning_graph
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def test_plan_signature_sees_what_pytorch_can_see():
    """The plan caches are keyed by a signature of the model's tensors and modes (native_tail):
    stable while nothing changes, different after an in-place write, a replaced ``.data`` (what
    pruning does), a resized tensor, a BatchNorm/Dropout mode switch or a swapped module; a tensor
    shared by two modules is no obstacle."""
    from torchpruner_b200.attributions.native_tail import model_signature

    torch.manual_seed(0)
    shared = nn.Linear(6, 6)
    model = nn.Sequential(nn.Conv2d(3, 4, 3), nn.BatchNorm2d(4), nn.ReLU(), nn.Flatten(),
                          nn.Dropout(0.5), shared, shared).eval()
    sig = model_signature(model)
    assert sig == model_signature(model)
    with torch.no_grad():
        model[0].weight.mul_(2.0)                      # in-place write: version counter
    sig2 = model_signature(model)
    assert sig2 != sig
    model[0].bias.data = model[0].bias.data.clone()    # same values, new storage
    sig3 = model_signature(model)
    assert sig3 != sig2
    model[0].weight.data = model[0].weight.data[:3].clone()   # pruned
    sig4 = model_signature(model)
    assert sig4 != sig3
    model[1].train()
    sig5 = model_signature(model)
    assert sig5 != sig4
    model[1].eval()
    assert model_signature(model) == sig4
    model[1].running_mean.add_(1.0)                    # buffers count too
    sig6 = model_signature(model)
    assert sig6 != sig4
    model[5] = nn.Linear(6, 6)
    assert model_signature(model) != sig6
