"""End-to-end parity of the drop-in Python API on a GPU: the reference's own unit tests restated
against torchpruner_b200 (same fixtures, same expected values), golden vectors produced by the
unmodified reference on the BASELINE configs, and the CPU oracle run side by side."""
import copy
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
from torchpruner_b200.attributions import (APoZAttributionMetric, RandomAttributionMetric,
                                           SensitivityAttributionMetric, ShapleyAttributionMetric,
                                           TaylorAttributionMetric, WeightNormAttributionMetric)
from torchpruner_b200.pruner import Pruner
from torchpruner_b200.utils import get_vgg_pruning_graph

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
DEV = torch.device("cuda:0")


@pytest.fixture(autouse=True)
def fp32_math():
    """The reference's oracle is fp32 on the CPU; keep cuDNN/cuBLAS out of TF32."""
    saved = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = saved


def gold(name):
    return np.load(os.path.join(GOLDEN, name))


def norm_err(got, want):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    assert got.shape == want.shape
    return np.abs(got - want).max() / max(np.abs(want).max(), 1e-30)


def max_model_loader(version):
    x, y, model = cases.max_model(version, device=DEV)
    return DataLoader(TensorDataset(x, y), batch_size=1, shuffle=False), model


# ------------------------------------------------------------------ reference unit tests, restated
@pytest.mark.parametrize("cls,kw,version,expected", [
    (WeightNormAttributionMetric, {}, 1, [1, 2, 2, 2]),
    (APoZAttributionMetric, {}, 1, [0.5, 0.5, 1, 1]),
    (SensitivityAttributionMetric, {}, 1, [0, 0, 0, 0]),
    (TaylorAttributionMetric, {}, 1, [0, 0, 0, 0]),
    (SensitivityAttributionMetric, {}, 2, [0.2, 0.1, 0.2, 0.04]),
    (TaylorAttributionMetric, {}, 2, [0.1, 0.1, 0.5, 0.1]),
    (TaylorAttributionMetric, {"signed": True}, 2, [0.1, 0.1, 0.5, -0.1]),
])
def test_known_answers(cls, kw, version, expected):
    """torchpruner/tests/test_attributions.py:84-175."""
    loader, model = max_model_loader(version)
    attr = cls(model, loader, F.mse_loss, DEV, **kw).run(list(model.children())[0])
    assert list(attr.shape) == [4] and attr.dtype == np.float32
    np.testing.assert_array_almost_equal(attr, expected)
    key = {WeightNormAttributionMetric: "weight_norm", APoZAttributionMetric: "apoz",
           SensitivityAttributionMetric: "sensitivity", TaylorAttributionMetric: "taylor"}[cls]
    if kw.get("signed"):
        key = "taylor_signed"
    for best in (False, True):
        got = cls(model, loader, F.mse_loss, DEV, **kw).run(list(model.children())[0],
                                                           find_best_evaluation_module=best)
        np.testing.assert_allclose(got, gold("max_model.npz")[f"v{version}/best{int(best)}/{key}"],
                                   rtol=1e-6, atol=1e-7)


def test_sv_known_answer_and_random():
    """test_attributions.py:74-82 and :128-137 (Shapley pinned to one decimal, unseeded there)."""
    loader, model = max_model_loader(1)
    np.random.seed(0)
    a = ShapleyAttributionMetric(model, loader, F.mse_loss, DEV, sv_samples=300)
    attr = a.run(list(model.children())[0])
    assert list(attr.shape) == [4] and attr.dtype == np.float64
    np.testing.assert_array_almost_equal(attr, [0.37, 0.37, 1.7, 0.], decimal=1)
    r = RandomAttributionMetric(model, loader, F.mse_loss, DEV).run(list(model.children())[0])
    assert list(r.shape) == [4]


@pytest.mark.parametrize("version", [1, 2])
@pytest.mark.parametrize("best", [False, True])
def test_shapley_slow_path_matches_reference_run(version, best):
    """Same seed, same permutations, same per-example table as the reference's run_module()."""
    loader, model = max_model_loader(version)
    np.random.seed(7)
    a = ShapleyAttributionMetric(model, loader, F.mse_loss, DEV, sv_samples=20, reduction="none")
    got = a.run(list(model.children())[0], find_best_evaluation_module=best)
    want = gold("max_model.npz")[f"v{version}/best{int(best)}/shapley_slow"]
    assert got.shape == want.shape
    np.testing.assert_allclose(got, want, rtol=0, atol=2e-7)


def test_all_metrics_run_with_best_module():
    """test_attributions.py:203-229."""
    loader, model = max_model_loader(1)
    layer = list(model.children())[0]
    for A in (TaylorAttributionMetric, SensitivityAttributionMetric, APoZAttributionMetric,
              WeightNormAttributionMetric):
        a = A(model, loader, F.mse_loss, DEV)
        np.testing.assert_array_almost_equal(a.run(layer, find_best_evaluation_module=False),
                                             a.run(layer, find_best_evaluation_module=True))
    for A in (RandomAttributionMetric, ShapleyAttributionMetric):
        assert list(A(model, loader, F.mse_loss, DEV).run(layer, find_best_evaluation_module=True).shape) == [4]


def test_reductions_and_row_order():
    model, shape = workloads.build("small_cnn")
    model.to(DEV)
    loader = workloads.synthetic_batches(shape, n=96, batch_size=40)  # ragged last batch (16)
    layer = model.features[3]
    table = APoZAttributionMetric(model, loader, workloads.cross_entropy, DEV, reduction="none").run(layer)
    assert table.shape == (96, 64)
    for red, fn in (("mean", lambda t: t.mean(0)), ("sum", lambda t: t.sum(0)),
                    (lambda t: t.max(0), lambda t: t.max(0))):
        got = APoZAttributionMetric(model, loader, workloads.cross_entropy, DEV, reduction=red).run(layer)
        np.testing.assert_array_equal(got, fn(table))
    # rows follow the loader order: recompute batch 2 alone on the CPU oracle
    x, _ = list(loader)[2]
    with torch.no_grad():
        z = model.forward_partial(x.to(DEV), to_module=layer)
    np.testing.assert_array_equal(table[80:], oracle.apoz_rows(z.cpu()))


# ------------------------------------------------------------------ config 1: small CNN vs reference
def test_small_cnn_matches_reference():
    g = gold("small_cnn.npz")
    model, shape = workloads.build("small_cnn")
    model.to(DEV)
    loader = workloads.synthetic_batches(shape, n=128, batch_size=64)
    mk = lambda cls, **kw: cls(model, loader, workloads.cross_entropy, DEV, **kw)
    report = {}
    for name, layer in (("conv1", model.features[0]), ("conv2", model.features[3])):
        for best in (False, True):
            tag = f"{name}/best{int(best)}"
            t = mk(TaylorAttributionMetric, reduction="none").run(layer, find_best_evaluation_module=best)
            model.zero_grad()
            s = mk(SensitivityAttributionMetric, reduction="none").run(layer, find_best_evaluation_module=best)
            model.zero_grad()
            a = mk(APoZAttributionMetric, reduction="none").run(layer, find_best_evaluation_module=best)
            report[tag] = (norm_err(t, g[f"{tag}/taylor"]), norm_err(s, g[f"{tag}/sensitivity"]),
                           float((a != g[f"{tag}/apoz"]).mean()))
            # Tier B: the forward/backward itself runs on different hardware (cuDNN vs oneDNN),
            # so activations differ by fp32 rounding before our kernels see them.
            assert report[tag][0] <= 1e-4 and report[tag][1] <= 1e-4, report
            # a rounding difference can flip `z > 0` at the threshold for a handful of pixels
            assert np.abs(a - g[f"{tag}/apoz"]).max() <= 2 and report[tag][2] <= 1e-3, report
        np.testing.assert_allclose(mk(WeightNormAttributionMetric).run(layer), g[f"{name}/weight_norm"], rtol=1e-5)
        np.random.seed(11)
        sv = mk(ShapleyAttributionMetric, sv_samples=5, reduction="none").run(layer, find_best_evaluation_module=True)
        assert sv.dtype == np.float64 and sv.shape == g[f"{name}/shapley"].shape
        # scores are sums of loss differences: compare on the loss scale (|loss| ~ 2.3)
        assert np.abs(sv - g[f"{name}/shapley"]).max() <= 5e-6, np.abs(sv - g[f"{name}/shapley"]).max()
        assert norm_err(sv.mean(0), g[f"{name}/shapley"].mean(0)) <= 2e-3
    np.testing.assert_allclose(mk(WeightNormAttributionMetric).run(model.classifier[0]),
                               g["fc1/weight_norm"], rtol=1e-5)


# ------------------------------------------------------------------ MLP (mnist.py) vs reference
@pytest.mark.parametrize("tag,layer_idx", [("fc1", 1), ("fc2", 3)])
def test_mlp_matches_reference(tag, layer_idx):
    """784-2024-2024-10 LeakyReLU perceptron: every metric against the unmodified reference's
    CPU output (tests/golden/mlp.npz); Shapley runs on the engine (Linear + LeakyReLU epilogue,
    dense incremental first layer)."""
    g = gold("mlp.npz")
    torch.manual_seed(0)
    model, shape = workloads.build("mlp")
    model.to(DEV)
    loader = workloads.synthetic_batches(shape, n=16, batch_size=16, seed=6)
    layer = model.fc[layer_idx]
    mk = lambda cls, **kw: cls(model, loader, workloads.cross_entropy, DEV, **kw)
    t = mk(TaylorAttributionMetric, reduction="none").run(layer, find_best_evaluation_module=True)
    model.zero_grad()
    s = mk(SensitivityAttributionMetric, reduction="none").run(layer, find_best_evaluation_module=True)
    model.zero_grad()
    a = mk(APoZAttributionMetric, reduction="none").run(layer, find_best_evaluation_module=True)
    assert norm_err(t, g[f"{tag}/taylor"]) <= 1e-4 and norm_err(s, g[f"{tag}/sensitivity"]) <= 1e-4
    assert float((a != g[f"{tag}/apoz"]).mean()) <= 1e-3   # sign flips of ~0 pre-activations
    np.testing.assert_allclose(mk(WeightNormAttributionMetric).run(layer), g[f"{tag}/weight_norm"], rtol=1e-5)
    np.random.seed(13)
    m = mk(ShapleyAttributionMetric, sv_samples=1, reduction="none")
    sv = m.run(layer, find_best_evaluation_module=True)
    want = g[f"{tag}/shapley16"]
    assert sv.shape == want.shape == (16, 2024) and sv.dtype == np.float64
    assert np.abs(sv - want).max() <= 5e-6, np.abs(sv - want).max()   # loss scale, |loss| ~ 2.3
    assert m.last_run_info["stacked"] is True


# ------------------------------------------------------------------ configs 2/3: VGG16 vs reference
def test_vgg16_scores_match_reference():
    g = gold("vgg16.npz")
    model, shape = workloads.build("vgg16_cifar")
    model.to(DEV)
    loader = workloads.synthetic_batches(shape, n=256, batch_size=256)
    names = {m: n for n, m in model.named_modules()}
    mk = lambda cls, **kw: cls(model, loader, workloads.cross_entropy, DEV, **kw)
    worst = {"taylor": 0.0, "sensitivity": 0.0, "apoz": 0.0}
    for layer in model.conv_layers():
        tag = names[layer]
        t = mk(TaylorAttributionMetric).run(layer, find_best_evaluation_module=True)
        model.zero_grad()
        s = mk(SensitivityAttributionMetric).run(layer, find_best_evaluation_module=True)
        model.zero_grad()
        a = mk(APoZAttributionMetric).run(layer, find_best_evaluation_module=True)
        worst["taylor"] = max(worst["taylor"], norm_err(t, g[f"{tag}/taylor"]))
        worst["sensitivity"] = max(worst["sensitivity"], norm_err(s, g[f"{tag}/sensitivity"]))
        worst["apoz"] = max(worst["apoz"], norm_err(a, g[f"{tag}/apoz"]))
        np.testing.assert_allclose(mk(WeightNormAttributionMetric).run(layer), g[f"{tag}/weight_norm"], rtol=1e-5)
    print("VGG16 tier-B worst normalised errors vs reference-CPU:", worst)
    # 13 conv+BN layers of fp32 forward/backward on different hardware sit between the input and
    # these scores; the kernels themselves are held to 1e-5 in test_gpu_kernels.py.
    assert worst["taylor"] <= 2e-3 and worst["sensitivity"] <= 2e-3 and worst["apoz"] <= 2e-3, worst


VGG_CONV_TAGS = ["features.%d" % i for i in (0, 3, 7, 10, 14, 17, 20, 24, 27, 30, 34, 37, 40)]


def spearman(a, b):
    """Rank correlation with average ranks for ties (scipy.stats.spearmanr on flattened input)."""
    from scipy.stats import spearmanr
    return float(spearmanr(np.ravel(a), np.ravel(b)).statistic)


@pytest.mark.parametrize("tag", VGG_CONV_TAGS)
def test_vgg16_shapley_matches_reference(tag):
    """Config 3 at a size the CPU reference finishes in minutes — sv_samples=2, 32 examples — on
    ALL 13 layers, against tables the unmodified reference produced (make_golden.py gen_vgg).
    The workload's classifier is at He scale, so scores are O(0.1-1) numbers: the bar is
    max|difference| <= 1e-4 of the largest score of the layer, and rank correlation >= 0.999 of
    both the [32, C] table and the per-channel means a user prunes by.  (The reference's own fp32
    arithmetic is 0.7-2.2e-5 of max away from its float64 twin on these tables; the engine's
    distance to the same twin is asserted to stay within 5e-5.)"""
    g = gold("vgg16.npz")
    model, shape = workloads.build("vgg16_cifar")
    model.to(DEV)
    small = workloads.synthetic_batches(shape, n=32, batch_size=32, seed=5)
    layer = dict(model.named_modules())[tag]
    np.random.seed(3)
    m = ShapleyAttributionMetric(model, small, workloads.cross_entropy, DEV, sv_samples=2, reduction="none")
    sv = m.run(layer, find_best_evaluation_module=True)
    want, twin = g[f"{tag}/shapley32"], g[f"{tag}/shapley32_f64"]
    assert sv.shape == want.shape == (32, layer.out_channels) and sv.dtype == np.float64
    info = m.last_run_info
    assert info["stacked"] is True and info["native_tail"] and info["incremental_first"], info
    assert info["native_head"], info
    scale = np.abs(want).max()
    assert scale > 0.1     # the comparison can fail: scores are not rounding noise
    err, err_twin, ref_twin = (np.abs(sv - want).max() / scale, np.abs(sv - twin).max() / scale,
                               np.abs(want - twin).max() / scale)
    print(f"C3 parity {tag}: max|score| {scale:.3f}  engine-reference {err:.2e}  "
          f"engine-fp64 {err_twin:.2e}  reference-fp64 {ref_twin:.2e}  (of max|score|)")
    assert err <= 1e-4, err
    assert err_twin <= 5e-5, (err_twin, ref_twin)
    assert spearman(sv, want) >= 0.999
    assert spearman(sv.mean(0), want.mean(0)) >= 0.999
    # the generic path (the user's forward_partial on stacked masks) meets the same bar
    np.random.seed(3)
    m.native_tail = False
    sv2 = m.run(layer, find_best_evaluation_module=True)
    assert m.last_run_info["native_tail"] is False
    # (the model's own fp32 forward on the GPU, cuDNN, measured 0.6-2.4e-4 from the twin)
    assert np.abs(sv2 - want).max() / scale <= 5e-4


def test_run_many_scores_every_layer_from_one_pass():
    """`run_many(modules)`: all 13 VGG16 layers from ONE forward / backward per batch and one
    reduction launch per batch, against 13 separate `run` calls (what the reference does)."""
    from torchpruner_b200 import _lib
    model, shape = workloads.build("vgg16_cifar")
    model.to(DEV)
    loader = workloads.synthetic_batches(shape, n=40, batch_size=16)   # ragged last batch
    convs = model.conv_layers()
    for cls, kw in ((TaylorAttributionMetric, {}), (TaylorAttributionMetric, {"signed": True}),
                    (SensitivityAttributionMetric, {}), (APoZAttributionMetric, {}),
                    (WeightNormAttributionMetric, {})):
        m = cls(model, loader, workloads.cross_entropy, DEV, reduction="none", **kw)
        prof = _lib.profile_begin(timed=False)
        many = m.run_many(convs, find_best_evaluation_module=True)
        launches = {k: v[0] for k, v in prof.records.items()}
        _lib.profile_end()
        model.zero_grad()
        assert len(many) == 13
        if cls is not WeightNormAttributionMetric:
            assert sum(launches.values()) == 3, launches     # one launch per batch
        for conv, got in zip(convs, many):
            want = m.run(conv, find_best_evaluation_module=True)
            model.zero_grad()
            assert got.shape == want.shape and got.dtype == want.dtype
            if cls is APoZAttributionMetric:
                np.testing.assert_array_equal(got, want)
            else:
                assert norm_err(got, want) <= 1e-5
    # the same module asked for twice, and the base class fallback (one run per module)
    a = APoZAttributionMetric(model, loader, workloads.cross_entropy, DEV)
    twice = a.run_many([convs[3], convs[3]], find_best_evaluation_module=True)
    np.testing.assert_array_equal(twice[0], twice[1])
    np.random.seed(0)
    r = RandomAttributionMetric(model, loader, workloads.cross_entropy, DEV).run_many(convs[:2])
    assert [x.shape for x in r] == [(64,), (64,)]


def test_run_many_rescores_layers_whose_activation_is_overwritten():
    """Stock ResNet-50: the evaluation module of conv1 / conv2 is a BatchNorm whose output the
    block's in-place ReLU overwrites before the pass is over, and every conv3's is the shared
    ReLU.  run_many must notice (version counters) and give those layers a pass of their own —
    same numbers as `run`, same error for conv3."""
    model, _ = workloads.build("resnet50_stock")
    model.to(DEV)
    loader = workloads.synthetic_batches((3, 96, 96), n=8, batch_size=4, seed=9)
    mods = dict(model.named_modules())
    picked = [mods[n] for n in ("conv1", "layer1.0.conv1", "layer2.1.conv2", "layer3.0.downsample.0")]
    for cls in (TaylorAttributionMetric, APoZAttributionMetric):
        m = cls(model, loader, workloads.cross_entropy, DEV, reduction="none")
        many = m.run_many(picked, find_best_evaluation_module=True)
        model.zero_grad()
        for mod, got in zip(picked, many):
            want = m.run(mod, find_best_evaluation_module=True)
            model.zero_grad()
            if cls is APoZAttributionMetric:
                np.testing.assert_array_equal(got, want)
            else:
                assert norm_err(got, want) <= 1e-5
        with pytest.raises(RuntimeError, match="different channel counts"):
            m.run_many([picked[0], mods["layer1.0.conv3"]], find_best_evaluation_module=True)
        model.zero_grad()


def test_shapley_schedules_agree():
    """Stacked evaluation, one-mask-at-a-time evaluation (train-mode fallback), any stack size and
    any rank split produce the same table up to fp32 rounding of the batched layers."""
    model, shape = workloads.build("small_cnn")
    model.to(DEV)
    loader = workloads.synthetic_batches(shape, n=24, batch_size=12, seed=4)
    layer = model.features[3]

    def run(**attrs):
        np.random.seed(9)
        m = ShapleyAttributionMetric(model, loader, workloads.cross_entropy, DEV, sv_samples=3, reduction="none")
        for k, v in attrs.items():
            setattr(m, k, v)
        return m.run(layer, find_best_evaluation_module=True), m

    base, m = run()
    assert m.last_run_info["evaluations"] == 2 * 3 * 64
    for attrs in ({"max_stack": 1}, {"max_stack": 7}, {"max_stack_bytes": 1, "max_stack_bytes_engine": 1}):
        other, _ = run(**attrs)
        assert np.abs(other - base).max() <= 2e-6

    # generic-criterion path (not recognised as cross entropy) against the fused one
    def odd_loss(output, target, reduction="mean"):
        return F.cross_entropy(output, target, reduction=reduction) * 1.5
    np.random.seed(9)
    scaled = ShapleyAttributionMetric(model, loader, odd_loss, DEV, sv_samples=3, reduction="none").run(
        layer, find_best_evaluation_module=True)
    assert np.abs(scaled / 1.5 - base).max() <= 2e-6

    # emulate a 4-rank job on one GPU: each "rank" computes its spans, tables are summed
    from torchpruner_b200.attributions.methods import shapley_values as sh
    total = np.zeros_like(base)
    evals = 0
    for rank in range(4):
        np.random.seed(9)
        mm = ShapleyAttributionMetric(model, loader, workloads.cross_entropy, DEV, sv_samples=3, reduction="none")
        mm._dist = lambda rank=rank: (None, 4, rank)
        saved = sh.allreduce_scores
        sh.allreduce_scores = lambda sv, group=None: sv
        try:
            total += mm.run(layer, find_best_evaluation_module=True)
        finally:
            sh.allreduce_scores = saved
        evals += mm.last_run_info["evaluations"]
    assert np.abs(total - base).max() <= 2e-6
    # 192 items / 4 ranks: cuts at 48, 96, 144 fall inside permutations -> 3 halo evaluations
    assert evals == 2 * (3 * 64 + 3)


def test_shapley_train_mode_falls_back_to_sequential():
    net, shape = workloads.build("vgg16_cifar")
    net.to(DEV)
    net.classifier[0].train()  # a Dropout in train mode: stacking masks is no longer legal
    loader = workloads.synthetic_batches(shape, n=4, batch_size=4)
    m = ShapleyAttributionMetric(net, loader, workloads.cross_entropy, DEV, sv_samples=1)
    np.random.seed(0)
    out = m.run(net.features[40], find_best_evaluation_module=True)
    assert out.shape == (512,) and m.last_run_info["stacked"] is False


# ------------------------------------------------------------------ pruner: reference tests restated
def simple_model():
    x, y = torch.ones((10, 3)), torch.randint(0, 10, (10, 1))
    return (x.to(DEV), y.to(DEV)), nn.Sequential(nn.Linear(3, 2), nn.ReLU(), nn.Linear(2, 1)).to(DEV)


def test_prune_parameter_and_module():
    """torchpruner/tests/test_pruner.py:38-70, plus the values (the reference checks shapes only)."""
    (x, y), model = simple_model()
    p = Pruner(model, input_size=(3,), device=DEV)
    module = list(model.children())[0]
    w0, b0 = module.weight.detach().clone(), module.bias.detach().clone()
    weight_id = id(module.weight)
    p.prune_parameter(module, "weight", [0], axis=0)
    p.prune_parameter(module, "bias", [0], axis=0)
    assert list(module.weight.shape) == [1, 3] and list(module.bias.shape) == [1]
    assert id(module.weight) == weight_id
    assert torch.equal(module.weight.data, w0[1:]) and torch.equal(module.bias.data, b0[1:])
    p.prune_parameter(module, "weight", [0], axis=1)
    assert list(module.weight.shape) == [1, 2] and torch.equal(module.weight.data, w0[1:, 1:])
    assert module.weight.is_contiguous()

    (x, y), model = simple_model()
    p = Pruner(model, input_size=(3,), device=DEV)
    module = list(model.children())[0]
    p.prune_module(module, [0], direction="out")
    assert list(module.weight.shape) == [1, 3]
    p.prune_module(module, [0], direction="in")
    assert list(module.weight.shape) == [1, 2]
    # index flavours the reference accepts: unsorted, duplicated, negative, ndarray, tensor
    lin = nn.Linear(4, 6).to(DEV)
    w = lin.weight.detach().clone()
    Pruner(lin, (4,), DEV).prune_parameter(lin, "weight", np.array([5, 2, 2]), axis=0)
    assert torch.equal(lin.weight.data, w[[0, 1, 3, 4]])
    Pruner(lin, (4,), DEV).prune_parameter(lin, "weight", torch.tensor([-1]), axis=1)
    assert torch.equal(lin.weight.data, w[[0, 1, 3, 4]][:, :3])


def test_nan_trick_index_discovery():
    """test_pruner.py:72-121."""
    def probe(model, input_size, poisoned, listeners):
        p = Pruner(model, input_size=input_size, device=DEV)
        kids = list(model.children())
        kids[0].register_forward_hook(p._nanify_hook(poisoned))
        for i in listeners:
            kids[i].register_forward_hook(p._detect_nan_hook())
        p._run_forward()
        return [list(getattr(kids[i], "_nan_indices")) for i in listeners], kids

    got, _ = probe(nn.Sequential(nn.Linear(3, 2), nn.ReLU(), nn.Linear(2, 1)).to(DEV), (3,), [0], [2])
    assert got == [[0]]
    got, _ = probe(nn.Sequential(nn.Conv2d(1, 3, 2), nn.ReLU(), nn.Flatten(), nn.Linear(12, 1)).to(DEV),
                   (1, 3, 3), [0], [3])
    assert got == [[0, 1, 2, 3]]
    got, _ = probe(nn.Sequential(nn.Conv2d(1, 3, 2), nn.ReLU(), nn.MaxPool2d(2), nn.Flatten(),
                                 nn.Linear(3, 1)).to(DEV), (1, 3, 3), [0], [4])
    assert got == [[0]]
    got, kids = probe(nn.Sequential(nn.Linear(3, 2), nn.BatchNorm1d(2), nn.Linear(2, 1)).to(DEV), (3,), [0], [1, 2])
    assert got == [[0], [0]] and kids[1]._activation_len == 2.0
    with pytest.raises(RuntimeError):  # the reference rejects non-int64 index arrays here too
        probe(nn.Sequential(nn.Linear(3, 2), nn.Linear(2, 1)).to(DEV), (3,), np.array([0], np.int32), [1])


def test_prune_model_cascades():
    """test_pruner.py:124-176."""
    (x, y), model = simple_model()
    Pruner(model, (3,), DEV).prune_model(model[0], [0], cascading_modules=[model[2]])
    assert list(model[0].weight.shape) == [1, 3] and list(model[2].weight.shape) == [1, 1]
    assert list(model(x).shape) == list(y.shape)

    model = nn.Sequential(nn.Linear(3, 2), nn.BatchNorm1d(2), nn.Linear(2, 1)).to(DEV)
    Pruner(model, (3,), DEV).prune_model(model[0], [0], [model[1], model[2]])
    assert list(model[0].weight.shape) == [1, 3] and list(model[2].weight.shape) == [1, 1]
    for name in ("weight", "bias", "running_var", "running_mean"):
        assert list(getattr(model[1], name).shape) == [1]
    assert list(model(x).shape) == list(y.shape)

    model = nn.Sequential(nn.Linear(3, 5), nn.Dropout(0.5), nn.Linear(5, 1)).to(DEV)
    Pruner(model, (3,), DEV).prune_model(model[0], [0], [model[1], model[2]])
    assert model[1].p == 0.5 * 4 / 5
    assert not hasattr(model[2], "_nan_indices")


def test_analytic_cascade_falls_back_to_the_probe():
    """Train-mode BatchNorm (the probe forward moves its running statistics, as in the
    reference) and links the trace cannot prove are left to the NaN trick."""
    model = nn.Sequential(nn.Linear(3, 4), nn.BatchNorm1d(4), nn.Linear(4, 1)).to(DEV)
    twin = copy.deepcopy(model)
    p, q = Pruner(model, (3,), DEV), Pruner(twin, (3,), DEV, cascade="probe")
    np.random.seed(5)
    p.prune_model(model[0], [1], [model[1], model[2]])
    np.random.seed(5)
    q.prune_model(twin[0], [1], [twin[1], twin[2]])
    assert p.last_cascade == q.last_cascade == "probe"
    for (k, a), b in zip(model.state_dict().items(), twin.state_dict().values()):
        assert torch.equal(a, b), k
    model.eval()
    p.prune_model(model[0], [0], [model[1], model[2]])
    assert p.last_cascade == "probe"  # the mode changed: the chain is recorded again
    p.prune_model(model[0], [0], [model[1], model[2]])
    assert p.last_cascade == "analytic" and list(model[2].weight.shape) == [1, 1]

    class Scaled(nn.Module):  # a functional op between the modules: no hook can see it
        def __init__(self):
            super().__init__()
            self.a, self.b = nn.Linear(3, 4), nn.Linear(4, 2)

        def forward(self, x):
            return self.b(self.a(x) * 2.0)

    net = Scaled().to(DEV).eval()
    p = Pruner(net, (3,), DEV)
    p.prune_model(net.a, [3], [net.b])
    p.prune_model(net.a, [0], [net.b])
    assert p.last_cascade == "probe" and list(net.b.weight.shape) == [2, 2]


@pytest.mark.parametrize("mode", ["analytic", "probe"])
def test_pruning_one_layer_twice_uses_the_current_widths(mode, capsys):
    """A second prune_model of an already narrowed layer (Dropout in the cascade, negative
    indices) against the oracle: the analytic cascade resolves from a trace recorded at the old
    widths and has to read the current ones from the model (cascade.py)."""
    torch.manual_seed(3)
    cpu = nn.Sequential(nn.Linear(10, 100), nn.ReLU(), nn.Dropout(0.5), nn.Linear(100, 16),
                        nn.ReLU(), nn.Linear(16, 3)).eval()
    gpu = copy.deepcopy(cpu).to(DEV)
    pruner, ref = Pruner(gpu, (10,), DEV, cascade=mode), oracle.OraclePruner(cpu, (10,))
    rounds = [(0, list(range(10)), [2, 3]), (0, [0, 5, -1, 88, 12, 40, 41, 42, 43], [2, 3]),
              (3, [1, -2], [5]), (0, [-81, 7], [2, 3, 5])]
    for r, (layer, units, casc) in enumerate(rounds):
        np.random.seed(40 + r)
        pruner.prune_model(gpu[layer], units, [gpu[i] for i in casc])
        np.random.seed(40 + r)
        ref.prune_model(cpu[layer], units, [cpu[i] for i in casc])
        assert np.random.random() == np.random.RandomState(40 + r).random_sample(21)[-1]
        assert pruner.last_cascade == ("probe" if (mode == "probe" or r == 0) else "analytic")
        assert gpu[2].p == cpu[2].p
        for (k, a), b in zip(gpu.state_dict().items(), cpu.state_dict().values()):
            assert torch.equal(a.cpu(), b), (r, k)
    # the last round listed the layer BEHIND the next Linear as well: everything is dead there, as
    # the NaN trick finds (a NaN input poisons every output of a Linear)
    assert list(gpu[0].weight.shape) == [79, 10] and list(gpu[5].weight.shape) == [3, 0]
    with pytest.raises(IndexError):   # 79 units are left: 79 was a valid index two rounds ago
        pruner.prune_model(gpu[0], [79], [gpu[3]])
    capsys.readouterr()


@pytest.mark.parametrize("mode", ["analytic", "probe"])
def test_resnet_blocks_prune_bit_exact(mode, capsys):
    """Cascading prune of every entry of get_resnet_pruning_graph on a torchvision-style
    bottleneck ResNet (in-place residual adds, shared ReLUs) against the CPU oracle: same tensors,
    same numpy RNG consumption; after the first probe forward the analytic cascade resolves every
    entry from the recorded graph, residual adds included."""
    from tests.test_cascade import _small_resnet
    from torchpruner_b200.utils import get_resnet_pruning_graph
    cpu = _small_resnet("bottleneck")
    gpu = copy.deepcopy(cpu).to(DEV)
    shape = (3, 32, 32)
    pruner, ref = Pruner(gpu, shape, DEV, cascade=mode), oracle.OraclePruner(cpu, shape)
    g_gpu, g_cpu = get_resnet_pruning_graph(gpu), get_resnet_pruning_graph(cpu)
    rng = np.random.RandomState(8)
    how = []
    for r, ((mod, casc), (cmod, ccasc)) in enumerate(zip(g_gpu, g_cpu)):
        n = mod.weight.shape[0]
        units = sorted(rng.choice(n, size=max(1, n // 3), replace=False).tolist())
        np.random.seed(60 + r)
        pruner.prune_model(mod, units, casc)
        how.append(pruner.last_cascade)
        a = np.random.random()
        np.random.seed(60 + r)
        ref.prune_model(cmod, units, ccasc)
        assert a == np.random.random()
    assert how == ["probe"] + [mode] * (len(g_gpu) - 1)
    for (k, a), b in zip(gpu.state_dict().items(), cpu.state_dict().values()):
        assert torch.equal(a.cpu(), b), k
    x = torch.randn((2,) + shape, generator=torch.Generator().manual_seed(1))
    with torch.no_grad():
        assert torch.allclose(gpu(x.to(DEV)).cpu(), cpu(x), atol=1e-4)
    capsys.readouterr()


@pytest.mark.parametrize("momentum", [0.0, 0.01])
def test_training_step_after_pruning(momentum):
    """test_pruner.py:180-228 (the momentum case raises KeyError in the reference on torch>=1.6)."""
    (x, y), _ = simple_model()
    model = nn.Sequential(nn.Linear(3, 2), nn.BatchNorm1d(2), nn.Linear(2, 1)).to(DEV)
    opt = torch.optim.SGD(model.parameters(), lr=0.01, momentum=momentum)
    p = Pruner(model, (3,), DEV, optimizer=opt if momentum else None)

    def step():
        model(x).mean().backward()
        opt.step()

    step()
    p.prune_model(model[0], [0], [model[1], model[2]])
    step()
    if momentum:
        assert list(opt.state[model[0].weight]["momentum_buffer"].shape) == [1, 3]
        assert list(model[0].weight.grad.shape) == [1, 3]


# ------------------------------------------------------------------ config 5 / 1: bit-exact cascade
@pytest.mark.parametrize("mode", ["analytic", "probe"])
@pytest.mark.parametrize("cfg", ["small_cnn", "vgg16_cifar"])
def test_cascading_prune_is_bit_exact(cfg, mode, capsys):
    with open(os.path.join(GOLDEN, "prune.txt")) as f:
        want = [line[len(cfg) + 1:].rstrip("\n") for line in f if line.startswith(cfg + " ")]
    model, shape, opt = cases.prune_setup(cfg, device=DEV)
    cpu_model, _, cpu_opt = cases.prune_setup(cfg, device="cpu")
    pruner = Pruner(model, shape, DEV, optimizer=opt, cascade=mode)
    check = oracle.OraclePruner(cpu_model, shape, optimizer=cpu_opt)
    graph = get_vgg_pruning_graph(model) if cfg == "vgg16_cifar" else model.get_pruning_graph()
    cpu_graph = get_vgg_pruning_graph(cpu_model) if cfg == "vgg16_cifar" else cpu_model.get_pruning_graph()
    # Unit choices and the probe batches both come from the global numpy RNG, interleaved exactly
    # as in the golden run: identical digests therefore also prove identical RNG consumption.
    np.random.seed(21)
    how = []
    for module, cascade in graph:
        pruner.prune_model(module, cases.prune_choice(module.weight.shape[0]), cascading_modules=cascade)
        how.append(pruner.last_cascade)
    # the first probe forward records the chain; every later cascade is resolved without one
    assert how == ["probe"] + [mode] * (len(graph) - 1)
    rng_after_product = np.random.get_state()[1][:8].copy()
    np.random.seed(21)
    for module, cascade in cpu_graph:
        check.prune_model(module, cases.prune_choice(module.weight.shape[0]), cascade)
    assert (np.random.get_state()[1][:8] == rng_after_product).all(), "numpy RNG consumed differently"
    got = cases.model_digest(model, opt)
    assert got == cases.model_digest(cpu_model, cpu_opt), "differs from the CPU oracle"
    assert got == want, "differs from the digests of the unmodified reference"
    printed = capsys.readouterr().out
    assert "Pruning" in printed and "(out)" in printed and "(in)" in printed
    # the pruned model still trains
    x = torch.randn((4,) + tuple(shape), device=DEV)
    model.train()
    model(x).sum().backward()
    opt.step()


# ------------------------------------------------------------------ config 4: ResNet-50
def test_resnet50_all_convs_taylor_apoz_sensitivity():
    """Config 4 on the hook-friendly ResNet-50 (53 convolutions, find_best_evaluation_module=True):
    the API's hook plumbing — including the clone that an in-place ReLU after the BatchNorm evaluation
    module makes necessary — against the same quantities computed with plain torch ops on the
    tensors the hooks see (the reference's arithmetic, taylor.py:40-44 / apoz.py:31-33)."""
    torch.manual_seed(0)
    model, shape = workloads.build("resnet50")
    model.to(DEV)
    loader = workloads.synthetic_batches(shape, n=12, batch_size=6)
    convs = model.conv_layers()
    assert len(convs) == 53
    mk = lambda cls, **kw: cls(model, loader, workloads.cross_entropy, DEV, **kw)
    from torchpruner_b200.utils import find_best_module_for_attributions
    checked = 0
    for conv in convs[::4] + convs[-2:]:
        eval_module = find_best_module_for_attributions(model, conv)
        want_t, want_a, want_s = [], [], []

        def fwd(mod, _inp, out):
            z = out.detach().clone()
            want_a.append((z > 0).float().flatten(2).sum(-1))
            def bwd(g, z=z):
                want_t.append((-1.0 * (g * z)).flatten(2).sum(-1).abs())
                want_s.append(g.abs().flatten(2).sum(-1))
            out.register_hook(bwd)

        h = eval_module.register_forward_hook(fwd)
        for x, y in loader:
            model.zero_grad()
            workloads.cross_entropy(model(x.to(DEV)), y.to(DEV)).backward()
        h.remove()
        model.zero_grad()
        t = mk(TaylorAttributionMetric, reduction="none").run(conv, find_best_evaluation_module=True)
        model.zero_grad()
        s = mk(SensitivityAttributionMetric, reduction="none").run(conv, find_best_evaluation_module=True)
        model.zero_grad()
        a = mk(APoZAttributionMetric, reduction="none").run(conv, find_best_evaluation_module=True)
        wt, ws, wa = (torch.cat(v).cpu().numpy() for v in (want_t, want_s, want_a))
        assert t.shape == wt.shape == (12, eval_module.num_features if hasattr(eval_module, "num_features") else wt.shape[1])
        assert norm_err(t, wt) <= 1e-5 and norm_err(s, ws) <= 1e-5
        np.testing.assert_array_equal(a, wa)
        checked += 1
    assert checked >= 15


def test_run_many_scores_all_53_resnet_layers_in_one_pass():
    """Hook-friendly ResNet-50: every evaluation module's output survives the pass, so run_many
    scores all 53 convolutions from one forward(/backward) per batch — 7x7 = 49-float rows go
    through the single-tensor kernels, everything else through multi-tensor launches — and agrees
    with `run` layer by layer."""
    from torchpruner_b200 import _lib
    torch.manual_seed(0)
    model, _ = workloads.build("resnet50")
    model.to(DEV)
    loader = workloads.synthetic_batches((3, 96, 96), n=8, batch_size=4, seed=3)
    convs = model.conv_layers()
    for cls in (APoZAttributionMetric, TaylorAttributionMetric):
        m = cls(model, loader, workloads.cross_entropy, DEV, reduction="none")
        prof = _lib.profile_begin(timed=False)
        many = m.run_many(convs, find_best_evaluation_module=True)
        launches = sum(v[0] for v in prof.records.values())
        _lib.profile_end()
        model.zero_grad()
        assert len(many) == 53 and launches < 2 * 53     # 2 batches; 53 launches each one by one
        for i in list(range(0, 53, 6)) + [51, 52]:
            want = m.run(convs[i], find_best_evaluation_module=True)
            model.zero_grad()
            if cls is APoZAttributionMetric:
                np.testing.assert_array_equal(many[i], want)
            else:
                assert norm_err(many[i], want) <= 1e-5


def test_stock_resnet50_matches_the_reference_where_it_is_defined():
    """Config 4 on torchvision's ResNet-50 as it ships, find_best_evaluation_module=True, against
    the unmodified reference (tests/golden/resnet50.{npz,txt}, make_golden.py gen_resnet).

    `find_best_module_for_attributions` walks declaration order (utils/graph.py:19-31), so the
    evaluation module is the stem's ReLU for conv1, the following BatchNorm for every
    conv1 / conv2 / downsample convolution — 37 layers on which all three hook metrics are
    defined — and the block's single shared `relu` for every conv3, which fires three times per
    forward with different channel counts: there the reference fails inside numpy / torch
    (apoz.py:37, taylor.py:40) and this library raises a RuntimeError that says why."""
    g = gold("resnet50.npz")
    with open(os.path.join(GOLDEN, "resnet50.txt")) as f:
        rows = [line.split() for line in f if line.strip()]
    model, _ = workloads.build("resnet50_stock")
    model.to(DEV)
    loader = workloads.synthetic_batches((3, 96, 96), n=8, batch_size=4, seed=9)
    modules = dict(model.named_modules())
    classes = {"taylor": TaylorAttributionMetric, "apoz": APoZAttributionMetric,
               "sensitivity": SensitivityAttributionMetric}
    worst = {"taylor": 0.0, "sensitivity": 0.0}
    rank = {"taylor": 1.0, "sensitivity": 1.0}
    apoz_cells = apoz_off = ok = failing = 0
    for tag, kind, status, *rest in rows:
        metric = classes[kind](model, loader, workloads.cross_entropy, DEV, reduction="none")
        model.zero_grad()
        if status == "raises":
            assert tag.endswith("conv3") and rest[0] in ("ValueError", "RuntimeError")
            with pytest.raises(RuntimeError, match="different channel counts|shared"):
                metric.run(modules[tag], find_best_evaluation_module=True)
            failing += 1
            continue
        # tier A: the tensors the hooks see in THIS run, copied out by a hook of the test's own,
        # through the oracle's arithmetic on the CPU
        seen = []

        def spy(_mod, _inp, out, seen=seen):
            z = out.detach().clone()
            if out.requires_grad:
                out.register_hook(lambda grad, z=z: seen.append((z.cpu(), grad.detach().clone().cpu())))
            else:
                seen.append((z.cpu(), None))

        eval_module = modules[rest[0][len("eval="):]]
        handle = eval_module.register_forward_hook(spy)
        try:
            got = metric.run(modules[tag], find_best_evaluation_module=True)
        finally:
            handle.remove()
        assert len(seen) == 2   # two batches
        if kind == "apoz":
            np.testing.assert_array_equal(got, np.concatenate([oracle.apoz_rows(z) for z, _ in seen]))
        elif kind == "taylor":
            assert norm_err(got, np.concatenate([oracle.taylor_rows(z, gr) for z, gr in seen])) <= 1e-5
        else:
            assert norm_err(got, np.concatenate([oracle.abs_sum_rows(gr) for _, gr in seen])) <= 1e-5
        # tier B: against the reference's own run on the CPU
        want = g[f"{tag}/{kind}"]
        assert got.shape == want.shape and got.dtype == want.dtype == np.float32
        if kind == "apoz":
            # counts of `z > 0` behind a BatchNorm: a value within rounding of 0 can land on the
            # other side after a differently rounded forward (tier B)
            apoz_cells += want.size
            apoz_off += int((got != want).sum())
            assert np.abs(got - want).max() <= 3
        else:
            worst[kind] = max(worst[kind], norm_err(got, want))
            rank[kind] = min(rank[kind], spearman(got.mean(0), want.mean(0)))
        ok += 1
    print("stock ResNet-50 vs reference-CPU: worst normalised error", worst, "worst rank correlation",
          rank, f"APoZ cells differing {apoz_off} of {apoz_cells}")
    assert ok == 111 and failing == 48
    # Gradients through 50 layers of ReLU / max-pool gates are discontinuous in the activations:
    # one gate that a differently rounded forward flips re-routes a gradient.  The reference's own
    # fp32 run is 0.4-1.5e-2 (normalised) from its float64 twin on this workload, so end to end
    # only a loose bar is meaningful; the arithmetic itself is held to 1e-5 above.
    assert worst["taylor"] <= 0.3 and worst["sensitivity"] <= 0.3, worst
    assert rank["taylor"] >= 0.95 and rank["sensitivity"] >= 0.95, rank
    assert apoz_off <= apoz_cells // 200


def test_shapley_full_size_telescopes():
    """Size-independent property at BASELINE config 3 size (bs 256, 512 channels): along one
    permutation the marginal contributions telescope, so every example's scores sum to
    loss(all channels removed) - loss(nothing removed)."""
    model, shape = workloads.build("vgg16_cifar")
    model.to(DEV)
    loader = workloads.synthetic_batches(shape, n=256, batch_size=256)
    layer = dict(model.named_modules())["features.24"]
    np.random.seed(2)
    m = ShapleyAttributionMetric(model, loader, workloads.cross_entropy, DEV, sv_samples=1, reduction="none")
    sv = m.run(layer, find_best_evaluation_module=True)
    assert sv.shape == (256, 512) and m.last_run_info["evaluations"] == 512
    assert m.last_run_info["native_tail"] is True
    x, y = next(iter(loader))
    x, y = x.to(DEV), y.to(DEV)
    twin = copy.deepcopy(model).double()   # the exact end points (cuDNN's fp32 is ~1e-4 off here)
    relu = twin.features[26]
    with torch.no_grad():
        z = twin.forward_partial(x.double(), to_module=relu)
        l0 = workloads.cross_entropy(twin.forward_partial(z, from_module=relu), y, reduction="none")
        l1 = workloads.cross_entropy(twin.forward_partial(torch.zeros_like(z), from_module=relu), y,
                                     reduction="none")
    want = (l1 - l0).cpu().numpy()
    # Each fp32 difference of adjacent losses is exact or off by half an ulp of the difference,
    # and the fp64 sum adds nothing: what is left is a random walk of those half-ulps over 512
    # steps plus the distance of the engine's fp32-accurate losses at both ends from their exact
    # values (losses of 2-10: ulp up to 9.5e-7, ten layers deep).
    assert np.abs(sv.sum(1) - want).max() <= 4e-5, np.abs(sv.sum(1) - want).max()
    assert np.abs(want).max() > 1.0   # He-scale classifier: removing a whole layer moves the loss


# ------------------------------------------------------------------ ablation curve
@pytest.mark.parametrize("cfg,tag,n", [("small_cnn", "features.3", 32), ("vgg16_cifar", "features.34", 32)])
@pytest.mark.parametrize("native", [True, False])
def test_ablation_curve_matches_reference(cfg, tag, n, native):
    from torchpruner_b200.ablation import ablation_curve

    g = gold("ablation.npz")
    model, shape = workloads.build(cfg)
    model.to(DEV)
    x, y = next(iter(workloads.synthetic_batches(shape, n=n, batch_size=n, seed=8)))
    module = dict(model.named_modules())[tag]
    ranking = np.random.RandomState(4).permutation(module.out_channels)
    loss, acc = ablation_curve(model, x.to(DEV), y.to(DEV), module, ranking,
                               criterion=workloads.cross_entropy, native_tail=native, max_stack=37)
    assert loss.shape == g[f"{cfg}/loss"].shape and loss.dtype == np.float32
    # mean losses of 2.3-3.5; the generic path's activation and layers are cuDNN's fp32
    rtol = 4e-6 if (native or cfg == "small_cnn") else 2e-5
    np.testing.assert_allclose(loss, g[f"{cfg}/loss"], rtol=rtol)
    flips = int((acc != g[f"{cfg}/acc"]).sum())   # an argmax can flip on a near-tie
    assert flips <= max(1, len(acc) // 100), flips
    # against the oracle on the CPU copy as well, with a partial ranking and a generic criterion
    cpu_model, _ = workloads.build(cfg)
    part = ranking[:9]
    from torchpruner_b200.utils import find_best_module_for_attributions
    cpu_module = dict(cpu_model.named_modules())[tag]
    want_loss, want_acc = oracle.ablation_curve(cpu_model, x, y, workloads.cross_entropy,
                                                find_best_module_for_attributions(cpu_model, cpu_module), part)
    odd = lambda o, t, reduction="mean": F.cross_entropy(o, t, reduction=reduction) * 2.0
    loss2, acc2 = ablation_curve(model, x.to(DEV), y.to(DEV), module, part, criterion=odd,
                                 native_tail=native)
    np.testing.assert_allclose(loss2 / 2.0, want_loss, rtol=rtol)
    assert int((acc2 != want_acc).sum()) <= 1


def test_ablation_then_shapley_on_a_tail_the_engine_does_not_run():
    """The notebook loops Taylor -> ablation -> Shapley over one unchanged model.  When the
    layers after the module are outside the engine's vocabulary (conv1 of the small CNN is
    followed by a 3x3 convolution with padding 2) ablation remembers "no plan"; Shapley must
    read that entry as such and take forward_partial, like ablation did."""
    from torchpruner_b200.ablation import ablation_curve

    model, shape = workloads.build("small_cnn")
    cpu_model = copy.deepcopy(model)
    model.to(DEV)
    loader = workloads.synthetic_batches(shape, n=8, batch_size=8, seed=2)
    x, y = next(iter(loader))
    conv1 = model.features[0]
    ranking = np.random.RandomState(1).permutation(32)
    loss, _ = ablation_curve(model, x.to(DEV), y.to(DEV), conv1, ranking,
                             criterion=workloads.cross_entropy)
    assert np.isfinite(loss).all()
    np.random.seed(6)
    m = ShapleyAttributionMetric(model, loader, workloads.cross_entropy, DEV, sv_samples=2,
                                 reduction="none")
    sv = m.run(conv1, find_best_evaluation_module=True)
    assert m.last_run_info["native_tail"] is False
    np.random.seed(6)
    want = oracle.shapley_with_partial(cpu_model, loader, workloads.cross_entropy, 2, 8,
                                       cpu_model.features[1])
    assert np.abs(sv - want).max() <= 5e-6 * max(1.0, np.abs(want).max())


def test_ablation_curve_redoes_an_fp16_overflow_in_tf32():
    from torchpruner_b200.ablation import ablation_curve
    from torchpruner_b200.attributions import native_tail

    g = gold("ablation.npz")
    model, shape = workloads.build("vgg16_cifar")
    model.to(DEV)
    x, y = next(iter(workloads.synthetic_batches(shape, n=32, batch_size=32, seed=8)))
    module = model.features[34]
    ranking = np.random.RandomState(4).permutation(module.out_channels)
    ablation_curve(model, x.to(DEV), y.to(DEV), module, ranking)          # caches the fp16 plan
    (plan,) = [e[1] for k, e in native_tail._PLANS[model].items() if k[0] == "tail"]
    assert plan.use_f16
    plan.steps[1]["s_in"] *= 2.0 ** 14                                     # force the overflow
    loss, acc = ablation_curve(model, x.to(DEV), y.to(DEV), module, ranking)
    assert plan.use_f16 is False and np.isfinite(loss).all()
    np.testing.assert_allclose(loss, g["vgg16_cifar/loss"], rtol=4e-6)


@pytest.mark.parametrize("seed", range(int(os.environ.get("TP_FUZZ_SEEDS", "6"))))
def test_random_models_prune_bit_exact(seed, capsys):
    """Cascading prune of randomly drawn sequential models (conv/BN/activations/pools/Dropout/
    flatten/linear) with SGD momentum and gradients: every tensor equal to the CPU oracle's, the
    numpy RNG left in the same state, in both cascade modes."""
    from tests.test_cascade import _random_net

    for mode in ("analytic", "probe"):
        rng = np.random.RandomState(2000 + seed)
        torch.manual_seed(seed)
        cpu_model, shape = _random_net(rng)
        model = copy.deepcopy(cpu_model).to(DEV)
        opts = []
        for mdl in (cpu_model, model):
            opt = torch.optim.SGD(mdl.parameters(), lr=0.1, momentum=0.9)
            gen = torch.Generator().manual_seed(seed)
            for p in mdl.parameters():
                p.grad = torch.randn(p.shape, generator=gen).to(p.device)
                opt.state[p]["momentum_buffer"] = torch.randn(p.shape, generator=gen).to(p.device)
            opts.append(opt)
        check = oracle.OraclePruner(cpu_model, shape, optimizer=opts[0])
        pruner = Pruner(model, shape, DEV, optimizer=opts[1], cascade=mode)
        np.random.seed(seed)
        for module, casc in get_vgg_pruning_graph(model):
            n = module.weight.shape[0]
            if n > 2:
                pruner.prune_model(module, sorted(np.random.choice(n, size=n // 3, replace=False).tolist()), casc)
        state = np.random.get_state()[1].copy()
        np.random.seed(seed)
        for module, casc in get_vgg_pruning_graph(cpu_model):
            n = module.weight.shape[0]
            if n > 2:
                check.prune_model(module, sorted(np.random.choice(n, size=n // 3, replace=False).tolist()), casc)
        assert np.array_equal(np.random.get_state()[1], state)
        assert cases.model_digest(model, opts[1]) == cases.model_digest(cpu_model, opts[0]), (seed, mode)
    capsys.readouterr()


def test_adam_state_is_sliced_with_the_parameter():
    """optimizer state beyond SGD / param_groups[0] follows the pruned parameter."""
    (x, y), _ = simple_model()
    model = nn.Sequential(nn.Linear(3, 4), nn.BatchNorm1d(4), nn.Linear(4, 1)).to(DEV)
    opt = torch.optim.Adam([{"params": model[0].parameters()}, {"params": model[1].parameters()},
                            {"params": model[2].parameters()}], lr=1e-3)

    def step():
        opt.zero_grad()
        model(x).mean().backward()
        opt.step()

    step()
    before = {k: v.clone() for k, v in opt.state[model[0].weight].items() if torch.is_tensor(v) and v.dim()}
    Pruner(model, (3,), DEV, optimizer=opt).prune_model(model[0], [1, 3], [model[1], model[2]])
    st = opt.state[model[0].weight]
    assert list(st["exp_avg"].shape) == [2, 3] and list(st["exp_avg_sq"].shape) == [2, 3]
    assert torch.equal(st["exp_avg"], before["exp_avg"][[0, 2]])
    assert list(opt.state[model[2].weight]["exp_avg"].shape) == [1, 2]
    assert list(opt.state[model[1].weight]["exp_avg"].shape) == [2]
    step()   # shapes are consistent: training continues
