"""Pins the CPU oracle (oracle/tp_oracle.py) to the reference: its known-answer tests and the
golden vectors the unmodified reference produced (tests/golden/make_golden.py).  CPU only."""
import os

import numpy as np
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.utils.data import DataLoader, TensorDataset

from oracle import tp_oracle as oracle
from tests import cases
from torchpruner_b200 import workloads

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def load(name):
    return np.load(os.path.join(GOLDEN, name))


# ------------------------------------------------------------------ tier A: operator vectors
@pytest.mark.parametrize("name", sorted(cases.OP_SHAPES))
def test_ops_match_reference_vectors(name):
    gold = load("ops.npz")
    z, g = cases.op_inputs(name)
    np.testing.assert_allclose(oracle.taylor_rows(z, g), gold[f"{name}/taylor"], rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(oracle.taylor_rows(z, g, signed=True), gold[f"{name}/taylor_signed"],
                               rtol=1e-6, atol=1e-9)
    np.testing.assert_array_equal(oracle.apoz_rows(z), gold[f"{name}/apoz"])
    np.testing.assert_array_equal(oracle.apoz_rows_int(z), gold[f"{name}/apoz"].astype(np.int64))
    np.testing.assert_allclose(oracle.abs_sum_rows(g), gold[f"{name}/sensitivity"], rtol=1e-6, atol=1e-9)
    # float64 arbiters agree with the fp32 reference to fp32 summation accuracy
    scale = np.abs(gold[f"{name}/taylor"]).max() + 1e-30
    assert np.abs(oracle.taylor_rows_f64(z, g) - gold[f"{name}/taylor"]).max() <= 1e-5 * scale


@pytest.mark.parametrize("name", sorted(cases.WEIGHT_SHAPES))
def test_weight_norm_matches_reference_vectors(name):
    gold = load("ops.npz")
    w = cases.weight_input(name)
    np.testing.assert_array_equal(oracle.weight_norm(w), gold[f"{name}/weight_norm"])
    np.testing.assert_allclose(oracle.weight_norm_f64(w), gold[f"{name}/weight_norm"], rtol=1e-5)


# ------------------------------------------------------------------ reference known answers
def _per_example_rows(model, x, y, layer, kind, signed=False):
    """Scores of `layer`'s output per example (batch size 1, mse), built from oracle ops."""
    rows = []
    for i in range(x.shape[0]):
        model.zero_grad()
        z = layer(x[i:i + 1])
        z.retain_grad()
        rest = list(model.children())[1:]
        out = z
        for m in rest:
            out = m(out)
        if kind == "apoz":
            rows.append(oracle.apoz_rows(z.detach()))
            continue
        F.mse_loss(out, y[i:i + 1]).backward()
        if kind == "taylor":
            rows.append(oracle.taylor_rows(z.detach(), z.grad, signed=signed))
        else:
            rows.append(oracle.abs_sum_rows(z.grad))
    return np.concatenate(rows, 0)


@pytest.mark.parametrize("kind,version,signed,expected", [
    ("apoz", 1, False, [0.5, 0.5, 1, 1]),              # test_attributions.py:95-104
    ("sensitivity", 1, False, [0, 0, 0, 0]),          # :106-115
    ("taylor", 1, False, [0, 0, 0, 0]),               # :117-126
    ("sensitivity", 2, False, [0.2, 0.1, 0.2, 0.04]), # :139-151
    ("taylor", 2, False, [0.1, 0.1, 0.5, 0.1]),       # :153-162
    ("taylor", 2, True, [0.1, 0.1, 0.5, -0.1]),       # :164-175
])
def test_reference_known_answers(kind, version, signed, expected):
    x, y, model = cases.max_model(version)
    layer = list(model.children())[0]
    rows = _per_example_rows(model, x, y, layer, kind, signed)
    np.testing.assert_array_almost_equal(rows.mean(0), expected)
    gold = load("max_model.npz")
    key = f"v{version}/best0/" + ("taylor_signed" if signed else kind)
    np.testing.assert_allclose(rows.mean(0), gold[key], rtol=1e-6, atol=1e-8)


def test_reference_known_answer_weight_norm():
    _, _, model = cases.max_model(1)
    np.testing.assert_array_almost_equal(
        oracle.weight_norm(list(model.children())[0].weight.detach()), [1, 2, 2, 2])  # :84-93


@pytest.mark.parametrize("version", [1, 2])
def test_shapley_slow_path_matches_reference(version):
    gold = load("max_model.npz")
    x, y, model = cases.max_model(version)
    loader = DataLoader(TensorDataset(x, y), batch_size=1, shuffle=False)
    np.random.seed(7)
    sv = oracle.shapley_without_partial(model, loader, F.mse_loss, 20, list(model.children())[0])
    np.testing.assert_allclose(sv, gold[f"v{version}/best0/shapley_slow"], rtol=0, atol=1e-12)
    if version == 1:  # test_attributions.py:128-137 pins the estimator to one decimal
        np.random.seed(0)
        many = oracle.shapley_without_partial(model, loader, F.mse_loss, 300,
                                              list(model.children())[0])
        np.testing.assert_array_almost_equal(many.mean(0), [0.37, 0.37, 1.7, 0.], decimal=1)


def test_shapley_fast_path_matches_reference_small_cnn():
    gold = load("small_cnn.npz")
    model, shape = workloads.build("small_cnn")
    loader = workloads.synthetic_batches(shape, n=128, batch_size=64)
    layer = model.features[0]
    np.random.seed(11)
    sv = oracle.shapley_with_partial(model, loader, workloads.cross_entropy, 5, 128,
                                     module=model.features[1])  # evaluation after the ReLU
    ref = gold["conv1/shapley"]
    assert sv.shape == ref.shape
    scale = np.abs(ref).max()
    assert np.abs(sv - ref).max() <= 1e-6 * scale
    assert layer is model.features[0]


@pytest.mark.parametrize("tag,layer_idx", [("fc1", 1), ("fc2", 3)])
def test_oracle_matches_reference_on_the_mlp(tag, layer_idx):
    """Third model family of the reference's experiments (mnist.py:9-35: Linear layers scored
    after LeakyReLU): the oracle's whole-metric restatements against the unmodified reference."""
    gold = load("mlp.npz")
    model, shape = workloads.build("mlp")
    loader = workloads.synthetic_batches(shape, n=16, batch_size=16, seed=6)
    layer, evaluation = model.fc[layer_idx], model.fc[layer_idx + 1]
    for kind in ("taylor", "sensitivity", "apoz"):
        model.zero_grad()
        got = oracle.hook_scores(model, loader, workloads.cross_entropy, evaluation, kind)
        ref = gold[f"{tag}/{kind}"]
        assert got.shape == ref.shape == (16, 2024)
        if kind == "apoz":
            np.testing.assert_array_equal(got, ref)
        else:
            assert np.abs(got - ref).max() <= 1e-6 * np.abs(ref).max()
    np.testing.assert_allclose(oracle.weight_norm(layer.weight.detach()), gold[f"{tag}/weight_norm"],
                               rtol=1e-6)
    np.random.seed(13)
    sv = oracle.shapley_with_partial(model, loader, workloads.cross_entropy, 1, 16, module=evaluation)
    ref = gold[f"{tag}/shapley16"]
    assert sv.shape == ref.shape == (16, 2024) and sv.dtype == np.float64
    assert np.abs(sv - ref).max() <= 1e-6 * np.abs(ref).max()


def test_mask_stack_and_scatter_restate_the_loop():
    """prefix_mask_stack + loss_delta_scatter compose to shapley_with_partial's inner loops."""
    model, shape = workloads.build("small_cnn")
    loader = workloads.synthetic_batches(shape, n=8, batch_size=8, seed=9)
    module = model.features[4]
    np.random.seed(5)
    perms = [np.random.permutation(64) for _ in range(2)]
    full = oracle.shapley_with_partial(model, loader, workloads.cross_entropy, 2, 8, module,
                                       permutations=perms)
    x, y = next(iter(loader))
    with torch.no_grad():
        z = model.forward_partial(x, to_module=module)
        loss0 = workloads.cross_entropy(model.forward_partial(z, from_module=module), y,
                                        reduction="none").numpy()
        sv = np.zeros((8, 64))
        for j in range(2):
            prev = loss0
            for k0 in range(0, 64, 24):
                P = min(24, 64 - k0)
                stack = torch.from_numpy(oracle.prefix_mask_stack(z, perms[j], k0, P))
                logits = model.forward_partial(stack.reshape((P * 8,) + tuple(z.shape[1:])),
                                               from_module=module)
                losses = oracle.cross_entropy_rows(logits, y.repeat(P)).reshape(P, 8)
                prev = oracle.loss_delta_scatter(losses, prev, perms[j], k0, 2, sv, 0)
    # Evaluating P masks as one [P*B] batch changes the blocking of the CPU GEMMs, so each loss
    # moves by ~1 ulp (2.4e-7 at |loss| ~ 2.3); scores are differences of losses, hence the
    # bound is stated on the loss scale and, relative to max|score|, at the 1e-5 of north_star.
    assert np.abs(sv - full).max() <= 4 * np.finfo(np.float32).eps * float(np.abs(loss0).max())
    assert np.abs(sv - full).max() <= 1e-5 * np.abs(full).max()


# ------------------------------------------------------------------ pruner
def test_nan_trick_known_answers():
    """torchpruner/tests/test_pruner.py:72-121 restated on the oracle's primitives."""
    torch.manual_seed(0)
    x = torch.rand(2, 1, 3, 3)
    conv = nn.Conv2d(1, 3, 2)
    y = torch.from_numpy(oracle.nan_fill(conv(x).detach(), [0]))
    np.testing.assert_array_equal(oracle.nan_indices(torch.flatten(torch.relu(y), 1)), [0, 1, 2, 3])
    pooled = nn.MaxPool2d(2)(torch.relu(y))
    np.testing.assert_array_equal(oracle.nan_indices(torch.flatten(pooled, 1)), [0])
    lin = torch.from_numpy(oracle.nan_fill(torch.rand(2, 2), [0]))
    np.testing.assert_array_equal(oracle.nan_indices(lin), [0])
    np.testing.assert_array_equal(oracle.keep_indices(5, [3, 0, 3]), [1, 2, 4])


@pytest.mark.parametrize("cfg", ["small_cnn", "vgg16_cifar"])
def test_oracle_pruner_matches_reference_digests(cfg):
    from torchpruner_b200.utils import get_vgg_pruning_graph

    with open(os.path.join(GOLDEN, "prune.txt")) as f:
        gold = [line[len(cfg) + 1:].rstrip("\n") for line in f if line.startswith(cfg + " ")]
    model, shape, opt = cases.prune_setup(cfg, device="cpu")
    pruner = oracle.OraclePruner(model, shape, optimizer=opt)
    graph = get_vgg_pruning_graph(model) if cfg == "vgg16_cifar" else model.get_pruning_graph()
    np.random.seed(21)
    for module, cascade in graph:
        pruner.prune_model(module, cases.prune_choice(module.weight.shape[0]), cascade)
    assert cases.model_digest(model, opt) == gold


@pytest.mark.parametrize("cfg,tag,n", [("small_cnn", "features.3", 32), ("vgg16_cifar", "features.34", 32)])
def test_ablation_curve_matches_notebook_loop(cfg, tag, n):
    from torchpruner_b200.utils import find_best_module_for_attributions

    gold = load("ablation.npz")
    model, shape = workloads.build(cfg)
    x, y = next(iter(workloads.synthetic_batches(shape, n=n, batch_size=n, seed=8)))
    module = dict(model.named_modules())[tag]
    ranking = np.random.RandomState(4).permutation(module.out_channels)
    loss, acc = oracle.ablation_curve(model, x, y, workloads.cross_entropy,
                                      find_best_module_for_attributions(model, module), ranking)
    np.testing.assert_allclose(loss, gold[f"{cfg}/loss"], rtol=1e-6)
    np.testing.assert_array_equal(acc, gold[f"{cfg}/acc"])
