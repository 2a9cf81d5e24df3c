"""The fp32-accurate tensor-core layer engine (tp_conv_tf32x3 and companions) against float64
convolutions, through the C-ABI."""
import contextlib
import io
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from torchpruner_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


# Engine-on against engine-off tables.  Engine off means the model's own fp32 forward on the GPU
# (cuDNN / cuBLAS, TF32 off) for the activation and for every masked evaluation, and that is the
# LESS accurate side: on VGG16 it sits 0.6-2.4e-4 of the largest score away from a float64
# evaluation of the same estimator, the engine 0.3-3e-5 (profiles/r02_parity_sources.txt;
# tests/test_gpu_api.py holds the engine to 1e-4 against the reference's own tables).  These
# comparisons catch structural mistakes, which are O(1).
LIB_TOL = 5e-4


def lib_close(got, ref, tol=LIB_TOL):
    return np.abs(got - ref).max() <= tol * max(np.abs(ref).max(), 0.05)


def _ref64(x_nhwc, w, taps, scale, shift, relu):
    """float64 reference: x [N,H,W,Cin], w [Cout, taps*Cin] with k = (r*3+s)*Cin + c."""
    N, H, W, Cin = x_nhwc.shape
    Cout = w.shape[0]
    x = x_nhwc.double().permute(0, 3, 1, 2)
    if taps == 9:
        w4 = w.double().view(Cout, 3, 3, Cin).permute(0, 3, 1, 2)
        y = F.conv2d(x, w4, padding=1)
    else:
        y = F.conv2d(x, w.double().view(Cout, Cin, 1, 1))
    y = y * scale.double().view(1, -1, 1, 1) + shift.double().view(1, -1, 1, 1)
    y = _act(y, relu)
    return y.permute(0, 2, 3, 1).contiguous()


def _act(y, spec):
    if isinstance(spec, tuple):
        return F.leaky_relu(y, spec[1])
    return y.relu() if spec else y


def test_tf32_split():
    x = torch.randn(100003, device=DEV) * 37.0
    hi, lo = _lib.tf32_split(x)
    assert int((hi.view(torch.int32) & 0x1FFF).abs().max()) == 0      # both terms are TF32-exact
    assert int((lo.view(torch.int32) & 0x1FFF).abs().max()) == 0
    rel = ((hi.double() + lo.double() - x.double()).abs() / x.double().abs().clamp_min(1e-30))
    assert float(rel.max()) <= 2.0 ** -22
    assert float((lo.abs() / x.abs().clamp_min(1e-30)).max()) <= 2.0 ** -11 * 1.001
    assert abs(float(((hi + lo).double() - x.double()).mean())) < 37.0 * 2e-9   # unbiased


@pytest.mark.parametrize("N,H,W,Cin,Cout,taps,relu", [
    (4, 32, 32, 64, 64, 9, True),      # VGG features.3-style layer
    (3, 16, 16, 128, 256, 9, True),
    (5, 8, 8, 256, 256, 9, True),      # 2 images per 128-pixel tile
    (19, 4, 4, 512, 512, 9, True),     # 8 images per tile, ragged last tile
    (70, 2, 2, 512, 512, 9, False),    # 32 images per tile, ragged
    (300, 1, 1, 512, 512, 1, True),    # linear layer
    (130, 1, 1, 512, 10, 1, False),    # logits: 10 outputs, unsplit result
    (2, 16, 16, 32, 64, 9, True),      # one k-block per tap
    (40, 32, 32, 64, 64, 9, True),     # 320 tiles: several tiles per persistent CTA
    (150, 8, 8, 256, 512, 9, False),   # position-major tiles (N >= 128), ragged image block
    (300, 4, 4, 512, 512, 9, True),    # position-major: corner / edge / interior tap subsets
    (131, 2, 2, 64, 64, 9, True),      # position-major at 2x2: 4 of 9 taps per tile
    (50, 8, 8, 96, 160, 9, True),      # Cout not a multiple of the 128-wide tile: clipped stores
    (140, 4, 4, 96, 96, 9, True),      # same with position-major tiles
    (33, 1, 1, 64, 96, 1, False),
    (200, 1, 1, 2024, 2024, 1, ("leaky", 0.01)),   # K tail (2024 = 63*32 + 8) + LeakyReLU
    (70, 1, 1, 784, 2024, 1, ("leaky", 0.2)),
])
@pytest.mark.parametrize("variant", [0, 1], ids=["persistent", "tile-per-cta"])
@pytest.mark.parametrize("promote", [1, 2, 8])
def test_conv_tf32x3_matches_float64(N, H, W, Cin, Cout, taps, relu, promote, variant):
    _lib.set_tuning("conv_promote", promote)
    _lib.set_tuning("conv_variant", variant)
    try:
        _conv_case(N, H, W, Cin, Cout, taps, relu, promote)
    finally:
        _lib.set_tuning("conv_promote", 2)
        _lib.set_tuning("conv_variant", 0)


def _conv_case(N, H, W, Cin, Cout, taps, relu, promote):
    gen = torch.Generator(device=DEV).manual_seed(N * 1000 + Cin)
    x = torch.randn((N, H, W, Cin), device=DEV, generator=gen)
    x = x * (torch.rand((N, H, W, Cin), device=DEV, generator=gen) > 0.3)   # ReLU-like zeros
    w = torch.randn((Cout, taps * Cin), device=DEV, generator=gen) / (taps * Cin) ** 0.5
    scale = torch.rand(Cout, device=DEV, generator=gen) + 0.5
    shift = torch.randn(Cout, device=DEV, generator=gen) * 0.1
    x_hi, x_lo = _lib.tf32_split(x)
    w_hi, w_lo = _lib.tf32_split(w)
    split = Cout != 10
    y_hi, y_lo = _lib.conv_tf32x3(x_hi, x_lo, w_hi, w_lo, taps, scale, shift, relu, split=split)
    y = y_hi + y_lo if split else y_hi
    want = _ref64(x, w, taps, scale, shift, relu)
    scale_ref = float(want.abs().max())
    err = float((y.double() - want).abs().max())
    # fp32 library convolution on the same data, for scale: 3xTF32 must be in the same class
    x32 = x.permute(0, 3, 1, 2)
    saved = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        if taps == 9:
            y32 = F.conv2d(x32, w.view(Cout, 3, 3, Cin).permute(0, 3, 1, 2), padding=1)
        else:
            y32 = F.conv2d(x32, w.view(Cout, Cin, 1, 1))
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = saved
    y32 = _act(y32 * scale.view(1, -1, 1, 1) + shift.view(1, -1, 1, 1), relu)
    err32 = float((y32.permute(0, 2, 3, 1).double() - want).abs().max())
    print(f"K={taps * Cin} promote={promote} max abs err: 3xTF32 {err:.3e}  fp32 library {err32:.3e}  "
          f"(scale {scale_ref:.3f})")
    # the tensor core's accumulator truncates: the bound scales with the k-blocks per partial sum
    assert err <= (1e-6 if promote <= 2 else 4e-6) * scale_ref, (err, err32, scale_ref)
    if split:
        assert int((y_hi.view(torch.int32) & 0x1FFF).abs().max()) == 0   # hi is TF32-exact


def test_mask_stack_nhwc_and_pool():
    gen = torch.Generator().manual_seed(5)
    B, C, H, W, P, k0 = 3, 32, 4, 4, 5, 7
    z = torch.randn((B, C, H, W), generator=gen)
    perm = np.random.RandomState(1).permutation(C)
    inv = np.empty(C, np.int32)
    inv[perm] = np.arange(C, dtype=np.int32)
    for pool in (False, True):
        Ho, Wo = (H // 2, W // 2) if pool else (H, W)
        hi = torch.empty(P * B * Ho * Wo * C, device=DEV)
        lo = torch.empty_like(hi)
        ghi, glo = _lib.mask_stack_nhwc(z.to(DEV), torch.from_numpy(inv).to(DEV), k0, P, pool, hi, lo)
        src = F.max_pool2d(z, 2) if pool else z
        for p in range(P):
            want = src.clone()
            want[:, perm[:k0 + p + 1]] = 0.0
            got = (ghi + glo)[p * B:(p + 1) * B].permute(0, 3, 1, 2).cpu()
            assert torch.equal(got == 0, want == 0)            # the mask is exact
            assert float((got - want).abs().max()) <= 2.0 ** -22 * float(want.abs().max())
    x = torch.randn((6, 8, 8, 64), generator=gen)
    phi, plo = _lib.maxpool2_split_nhwc(x.to(DEV))
    want = F.max_pool2d(x.permute(0, 3, 1, 2), 2).permute(0, 2, 3, 1)
    assert float(((phi + plo).cpu() - want).abs().max()) <= 2.0 ** -22 * float(want.abs().max())


@pytest.mark.parametrize("tag", ["features.0", "features.10", "features.27", "features.40", "classifier.1"])
def test_native_tail_agrees_with_forward_partial(tag):
    """Whole-run check of the tcgen05 layer engine: same seed, engine on vs engine off."""
    import numpy as np
    from torchpruner_b200 import workloads
    from torchpruner_b200.attributions import ShapleyAttributionMetric

    saved = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        model, shape = workloads.build("vgg16_cifar")
        model.to(DEV)
        loader = workloads.synthetic_batches(shape, n=16, batch_size=16, seed=5)
        layer = dict(model.named_modules())[tag]
        tables = {}
        for native, incremental in (("auto", True), ("auto", False), (False, False)):
            np.random.seed(3)
            m = ShapleyAttributionMetric(model, loader, workloads.cross_entropy, DEV, sv_samples=1,
                                         reduction="none")
            m.native_tail = native
            m.incremental_first = incremental
            m.max_stack = 48   # several chunks per permutation
            tables[(native, incremental)] = m.run(layer, find_best_evaluation_module=True)
            assert m.last_run_info["native_tail"] is bool(native), m.last_run_info
            # classifier.1's tail is a single ReLU->Dropout->Linear->ReLU->Linear chain: its first
            # layer is chained to another one, so every tag here has an incremental form
            assert m.last_run_info["incremental_first"] is bool(native and incremental), m.last_run_info
            if native:
                print(tag, "incremental" if incremental else "direct", m.last_run_info["native"])
        # second run on the unchanged model: the cached, verified plan also supplies the unmasked
        # loss (no library forward of the tail at all) — same table up to 1 ulp of a loss
        np.random.seed(3)
        m = ShapleyAttributionMetric(model, loader, workloads.cross_entropy, DEV, sv_samples=1,
                                     reduction="none")
        m.max_stack = 48
        again = m.run(layer, find_best_evaluation_module=True)
        assert m.last_run_info["native"].get("cached") is True
        assert np.abs(again - tables[("auto", True)]).max() <= 2e-6   # losses of 2-10: 1-2 ulps
        ref = tables[(False, False)]
        for key in (("auto", True), ("auto", False)):
            diff = np.abs(tables[key] - ref).max()
            print(tag, key, "max |native - forward_partial| =", diff, "max |score| =", np.abs(ref).max())
            assert lib_close(tables[key], ref), diff
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = saved


def test_cached_plan_is_dropped_when_the_model_changes():
    """A cached plan holds prepared copies of the weights: an in-place update (optimizer step),
    pruning (``param.data`` replaced) or a BatchNorm mode switch must all invalidate it."""
    from torchpruner_b200 import workloads
    from torchpruner_b200.attributions import ShapleyAttributionMetric
    from torchpruner_b200.pruner import Pruner

    saved = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
    try:
        model, shape = workloads.build("vgg16_cifar")
        model.to(DEV)
        loader = workloads.synthetic_batches(shape, n=16, batch_size=16, seed=5)
        layer = model.features[30]

        def run(native):
            np.random.seed(3)
            m = ShapleyAttributionMetric(model, loader, workloads.cross_entropy, DEV, sv_samples=1,
                                         reduction="none")
            m.native_tail = native
            return m.run(layer, find_best_evaluation_module=True), m.last_run_info

        first, info = run("auto")
        assert info["native_tail"] and not info["native"].get("cached")
        _, info = run("auto")
        assert info["native"].get("cached") is True
        with torch.no_grad():                      # what an optimizer step does
            model.classifier[6].weight.mul_(3.0)
            model.features[38].running_mean.add_(0.05)
        got, info = run("auto")
        assert info["native_tail"] and not info["native"].get("cached")
        want, _ = run(False)
        assert lib_close(got, want) and np.abs(got - first).max() > 1e-2
        # a write through .data (older masking / fine-tuning code) bumps no version counter and
        # moves no pointer: only the plan's checksum of the weight bits can see it
        _, info = run("auto")
        assert info["native"].get("cached") is True
        before = got
        model.features[34].weight.data[:40].mul_(0.0)
        model.classifier[1].bias.data.add_(0.25)
        got, info = run("auto")
        assert info["native_tail"] and not info["native"].get("cached")
        want, _ = run(False)
        assert lib_close(got, want) and np.abs(got - before).max() > 1e-2
        with contextlib.redirect_stdout(io.StringIO()):
            pruner = Pruner(model, shape, DEV)
            pruner.prune_model(model.features[37], list(range(64)),
                               [model.features[38], model.features[40]])
        assert model.features[37].out_channels == 512     # stale, as in the reference
        got, info = run("auto")
        assert info["native_tail"] and not info["native"].get("cached")   # 448 channels: tileable
        want, _ = run(False)
        assert lib_close(got, want)
        with contextlib.redirect_stdout(io.StringIO()):
            pruner.prune_model(model.features[37], [0, 5, 9], [model.features[38], model.features[40]])
        got, info = run("auto")                    # 445 channels: zero-padded to 448 inside the engine
        assert info["native_tail"] and info["incremental_first"]
        want, _ = run(False)
        assert lib_close(got, want)
        model.features[38].train()                 # train-mode BatchNorm: no stacking, no engine
        _, info = run("auto")
        assert info["native_tail"] is False and info["stacked"] is False
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = saved


def test_score_then_prune_workflow_engine_on_vs_off():
    """The reference's layerwise loop (notebook "VGG16 - Layerwise robustness"): score a layer,
    prune its weakest units, move to the next entry of the pruning graph — the later layers are
    then scored on an already pruned model.  Engine on and engine off must pick the same units
    and end with the same model."""
    import copy
    from torchpruner_b200 import workloads
    from torchpruner_b200.attributions import ShapleyAttributionMetric
    from torchpruner_b200.pruner import Pruner
    from torchpruner_b200.utils import get_vgg_pruning_graph

    saved = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
    try:
        base, shape = workloads.build("vgg16_cifar")
        models = {True: copy.deepcopy(base).to(DEV), False: copy.deepcopy(base).to(DEV)}
        loader = workloads.synthetic_batches(shape, n=16, batch_size=16, seed=5)
        picked = {True: [], False: []}
        used_engine = 0
        for native, model in models.items():
            pruner = Pruner(model, shape, DEV)
            graph = get_vgg_pruning_graph(model)
            for step, (module, cascade) in enumerate(graph[:5]):   # fc2, fc1, conv13, conv12, conv11
                np.random.seed(100 + step)
                m = ShapleyAttributionMetric(model, loader, workloads.cross_entropy, DEV, sv_samples=1)
                m.native_tail = "auto" if native else False
                with contextlib.redirect_stdout(io.StringIO()):
                    scores = m.run(module, find_best_evaluation_module=True)
                    used_engine += int(m.last_run_info["native_tail"])
                    weakest = np.argsort(scores)[:64]
                    pruner.prune_model(module, weakest.tolist(), cascading_modules=cascade)
                picked[native].append((scores, np.sort(weakest)))
        assert used_engine >= 4
        for (s_on, w_on), (s_off, w_off) in zip(picked[True], picked[False]):
            assert lib_close(s_on, s_off)
            # the weakest-64 sets agree except where two scores tie within that rounding
            assert len(set(w_on) ^ set(w_off)) <= 4
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = saved


@pytest.mark.parametrize("widths", [(48, 80, 24), (64, 100, 12)])
def test_native_tail_pads_odd_channel_counts(widths):
    """Layer widths that are not multiples of the 64-channel k-block (what pruning leaves
    behind): the engine zero-pads its own copies of activations and weights."""
    from torchpruner_b200 import workloads
    from torchpruner_b200.attributions import ShapleyAttributionMetric

    a, b, c = widths

    class Net(workloads._PartialSequential):
        def __init__(self):
            super().__init__()
            self.features = torch.nn.Sequential(
                torch.nn.Conv2d(3, a, 3, padding=1), torch.nn.BatchNorm2d(a), torch.nn.ReLU(),
                torch.nn.Conv2d(a, b, 3, padding=1), torch.nn.BatchNorm2d(b), torch.nn.ReLU(),
                torch.nn.MaxPool2d(2),
                torch.nn.Conv2d(b, c, 3, padding=1), torch.nn.LeakyReLU(0.1), torch.nn.MaxPool2d(2))
            self.classifier = torch.nn.Sequential(torch.nn.Linear(c * 4, 36), torch.nn.ReLU(),
                                                  torch.nn.Linear(36, 10))

    saved = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
    try:
        torch.manual_seed(a)
        model = Net()
        workloads.warm_batchnorm(model, (3, 8, 8), n=32)
        model.to(DEV)
        loader = workloads.synthetic_batches((3, 8, 8), n=24, batch_size=24, seed=2)
        for layer in (model.features[0], model.features[3], model.features[7], model.classifier[0]):
            tables = {}
            for native in ("auto", False):
                np.random.seed(9)
                m = ShapleyAttributionMetric(model, loader, workloads.cross_entropy, DEV, sv_samples=2,
                                             reduction="none")
                m.native_tail = native
                with contextlib.redirect_stdout(io.StringIO()):
                    tables[native] = m.run(layer, find_best_evaluation_module=True)
                assert m.last_run_info["native_tail"] is bool(native), (layer, m.last_run_info)
            assert lib_close(tables["auto"], tables[False])
            assert np.abs(tables[False]).max() > 1e-4
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = saved


@pytest.mark.parametrize("tag", ["features.3", "features.27"])
def test_native_tail_over_several_ragged_batches(tag):
    """Three batches, the last one smaller (and small enough that the stacked batch of a 64-channel
    layer drops below one 128-image tile): same permutations for every batch, rows filled batch
    by batch, per-batch scratch re-sized — engine on vs engine off."""
    from torchpruner_b200 import workloads
    from torchpruner_b200.attributions import ShapleyAttributionMetric

    saved = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
    try:
        model, shape = workloads.build("vgg16_cifar")
        model.to(DEV)
        loader = workloads.synthetic_batches(shape, n=21, batch_size=10, seed=11)   # 10 + 10 + 1
        layer = dict(model.named_modules())[tag]
        tables = {}
        for native in ("auto", False):
            np.random.seed(4)
            m = ShapleyAttributionMetric(model, loader, workloads.cross_entropy, DEV, sv_samples=2,
                                         reduction="none")
            m.native_tail = native
            with contextlib.redirect_stdout(io.StringIO()):
                tables[native] = m.run(layer, find_best_evaluation_module=True)
            assert m.last_run_info["native_tail"] is bool(native)
        assert tables["auto"].shape == (21, layer.out_channels)
        assert lib_close(tables["auto"], tables[False])
        assert np.abs(tables[False][20]).max() > 0
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = saved


@pytest.mark.parametrize("contrib", ["stored", "direct"])
def test_sharded_ranks_start_mid_permutation_on_the_engine(contrib, monkeypatch):
    """3 emulated ranks share ONE permutation of a 512-channel VGG16 layer: ranks 1 and 2 start
    inside it, so the incremental first layer has to advance its running sum without output
    (accumulate-only pass) and one halo evaluation supplies the previous loss."""
    from torchpruner_b200 import workloads
    from torchpruner_b200.attributions import ShapleyAttributionMetric, native_tail
    from torchpruner_b200.attributions.methods import shapley_values as sh

    monkeypatch.setattr(native_tail, "CONTRIB", contrib)
    monkeypatch.setattr(sh, "allreduce_scores", lambda sv, group=None: sv)
    saved = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
    try:
        model, shape = workloads.build("vgg16_cifar")
        model.to(DEV)
        loader = workloads.synthetic_batches(shape, n=16, batch_size=16, seed=5)
        layer = model.features[24]

        def run(world, rank):
            np.random.seed(6)
            m = ShapleyAttributionMetric(model, loader, workloads.cross_entropy, DEV, sv_samples=1,
                                         reduction="none")
            m.max_stack = 100
            if world > 1:
                m._dist = lambda: (None, world, rank)
            with contextlib.redirect_stdout(io.StringIO()):
                sv = m.run(layer, find_best_evaluation_module=True)
            assert m.last_run_info["native_tail"] and m.last_run_info["incremental_first"]
            return sv, m.last_run_info["evaluations"]

        base, n_base = run(1, 0)
        parts = [run(3, r) for r in range(3)]
        assert n_base == 512 and sum(n for _, n in parts) == 512 + 2
        total = sum(sv for sv, _ in parts)
        assert np.abs(total - base).max() <= 5e-6     # the same losses (2-10: ulp <= 1e-6), summed in another order
        # each rank touched only its own channels of the permutation
        touched = [(np.abs(sv).sum(0) > 0) for sv, _ in parts]
        assert not (touched[0] & touched[1]).any() and not (touched[1] & touched[2]).any()
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = saved


def _random_vgg_like(rng):
    from torchpruner_b200 import workloads

    feats, c, hw = [], 3, int(rng.choice([4, 8, 16]))
    in_hw = hw
    for _ in range(int(rng.randint(2, 5))):
        out = int(rng.choice([8, 20, 48, 64, 72]))
        feats.append(torch.nn.Conv2d(c, out, 3, padding=1, bias=bool(rng.rand() < 0.7)))
        c = out
        if rng.rand() < 0.7:
            feats.append(torch.nn.BatchNorm2d(c))
        feats.append(torch.nn.ReLU(inplace=bool(rng.rand() < 0.5)) if rng.rand() < 0.7
                     else torch.nn.LeakyReLU(0.05))
        if hw > 1 and rng.rand() < 0.5:
            feats.append(torch.nn.MaxPool2d(2))
            hw //= 2
        if rng.rand() < 0.2:
            feats.append(torch.nn.Dropout2d(0.1))
    cls, f = [], c * hw * hw
    for _ in range(int(rng.randint(0, 3))):
        out = int(rng.choice([12, 40, 100]))
        if rng.rand() < 0.3:
            cls.append(torch.nn.Dropout(0.2))
        cls += [torch.nn.Linear(f, out), torch.nn.ReLU() if rng.rand() < 0.6 else torch.nn.LeakyReLU()]
        f = out
    cls.append(torch.nn.Linear(f, int(rng.choice([3, 10, 17]))))

    class Net(workloads._PartialSequential):
        def __init__(self):
            super().__init__()
            self.features = torch.nn.Sequential(*feats)
            self.classifier = torch.nn.Sequential(*cls)

    return Net(), (3, in_hw, in_hw)


@pytest.mark.parametrize("seed", range(int(os.environ.get("TP_FUZZ_SEEDS", "8"))))
def test_random_chains_engine_on_vs_off(seed):
    """Randomly drawn conv/BN/ReLU/LeakyReLU/MaxPool/Dropout/Linear chains (odd widths, bias-free
    convs, in-place activations, 3/10/17 classes): every conv and hidden Linear scored with the
    engine and with the model's own forward_partial."""
    from torchpruner_b200 import workloads
    from torchpruner_b200.attributions import ShapleyAttributionMetric

    saved = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
    try:
        rng = np.random.RandomState(500 + seed)
        torch.manual_seed(seed)
        model, shape = _random_vgg_like(rng)
        workloads.warm_batchnorm(model, shape, n=32)
        model.to(DEV)
        n_cls = model.classifier[-1].out_features
        loader = workloads.synthetic_batches(shape, n=12, batch_size=12, num_classes=n_cls, seed=seed)
        layers = [m for m in list(model.features) + list(model.classifier)[:-1]
                  if isinstance(m, (torch.nn.Conv2d, torch.nn.Linear))]
        engine_runs = 0
        for layer in layers:
            tables = {}
            for native in ("auto", False):
                np.random.seed(seed)
                m = ShapleyAttributionMetric(model, loader, workloads.cross_entropy, DEV, sv_samples=1,
                                             reduction="none")
                m.native_tail = native
                with contextlib.redirect_stdout(io.StringIO()):
                    tables[native] = m.run(layer, find_best_evaluation_module=True)
                if native:
                    engine_runs += int(m.last_run_info["native_tail"])
            scale = max(np.abs(tables[False]).max(), 1e-3)
            assert lib_close(tables["auto"], tables[False]), (seed, layer)
        assert engine_runs >= len(layers) - 1, (engine_runs, len(layers))
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = saved


def test_engine_propagates_nan_like_aten():
    """ReLU and max pooling keep a NaN (torch.relu / max_pool2d do; fmaxf would not): an fp16
    overflow upstream (inf in hi, inf - inf = NaN in lo) has to reach the logits to be noticed."""
    N, H, W, C = 128, 4, 4, 64
    x = torch.randn((N, H, W, C), device=DEV)
    w = torch.randn((C, 9 * C), device=DEV) / 24.0
    one, zero = torch.ones(C, device=DEV), torch.zeros(C, device=DEV)
    x_hi, x_lo = _lib.f16_split(x, 1.0)
    w_hi, w_lo = _lib.f16_split(w, 1.0)
    x_hi[5, 1, 2, 7] = float("inf")
    x_lo[5, 1, 2, 7] = float("nan")
    for pool in (False, True):
        y_hi, y_lo = _lib.conv_tf32x3(x_hi, x_lo, w_hi, w_lo, 9, one, zero, True, split=True, pool=pool)
        bad = ~torch.isfinite(y_hi.float() + y_lo.float())
        assert bool(bad[5].any()) and not bool(bad[:5].any()) and not bool(bad[6:].any())
    raw = torch.randn((4, 4, 4, 8), device=DEV)
    raw[2, 1, 1, 3] = float("nan")
    hi, lo = _lib.maxpool2_split_nhwc(raw, half=True, scale=1.0)
    assert bool(torch.isnan(hi[2, 0, 0, 3])) and int(torch.isnan(hi).sum()) == 1
    want = F.max_pool2d(raw.permute(0, 3, 1, 2), 2).permute(0, 2, 3, 1)
    assert torch.equal(torch.isnan(hi), torch.isnan(want))


def test_fp16_range_failure_falls_back_to_tf32_with_the_same_permutations(monkeypatch):
    """An activation that leaves fp16's range under some mask makes scores non-finite; the run is
    then repeated on the TF32 flavour with the numpy RNG rewound, so the caller gets the table —
    and the RNG state — of a run that had used TF32 from the start."""
    import copy
    from torchpruner_b200 import workloads
    from torchpruner_b200.attributions import ShapleyAttributionMetric, native_tail

    saved = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
    try:
        model, shape = workloads.build("vgg16_cifar")
        twin = copy.deepcopy(model)
        model.to(DEV)
        twin.to(DEV)
        loader = workloads.synthetic_batches(shape, n=16, batch_size=16, seed=2)
        mk = lambda mdl: ShapleyAttributionMetric(mdl, loader, workloads.cross_entropy, DEV,
                                                  sv_samples=2, reduction="none")
        layer = model.features[30]
        np.random.seed(1)
        m = mk(model)
        m.run(layer, find_best_evaluation_module=True)           # builds + caches the fp16 plan
        assert m.last_run_info["f16"] and m.last_run_info["native_tail"]
        (plan,) = [e[1] for k, e in native_tail._PLANS[model].items() if k[0] == "tail"]
        plan.steps[1]["s_in"] *= 2.0 ** 14                       # next layer's input overflows
        np.random.seed(7)
        m = mk(model)
        sv = m.run(layer, find_best_evaluation_module=True)
        state = np.random.get_state()[1].copy()
        assert np.isfinite(sv).all() and m.last_run_info["f16"] is False and m._f16_banned
        # the same job on a model whose plan never had the fp16 flavour
        monkeypatch.setenv("TORCHPRUNER_B200_F16", "0")
        np.random.seed(1)
        mk(twin).run(twin.features[30], find_best_evaluation_module=True)   # caches its plan too
        np.random.seed(7)
        m2 = mk(twin)
        want = m2.run(twin.features[30], find_best_evaluation_module=True)
        assert m2.last_run_info["f16"] is False
        np.testing.assert_array_equal(sv, want)
        assert np.array_equal(np.random.get_state()[1], state)
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = saved


def test_native_tail_declines_unsupported_models():
    from torchpruner_b200 import workloads
    from torchpruner_b200.attributions.native_tail import NativeTail

    model, shape = workloads.build("small_cnn")   # conv with padding 2 downstream of conv1
    model.to(DEV)
    x = torch.randn((4,) + shape, device=DEV)
    with torch.no_grad():
        z = model.forward_partial(x, to_module=model.features[1])
    assert NativeTail.build(model, model.features[1], z) is None
    with torch.no_grad():
        z = model.forward_partial(x, to_module=model.features[4])   # pool -> flatten -> fc -> fc
    tail = NativeTail.build(model, model.features[4], z)
    assert tail is not None and tail.pool_first and len(tail.steps) == 2


def test_plan_verification_arbitrates_rounding_with_a_float64_twin(monkeypatch):
    """The model's own fp32 forward (cuDNN) is itself up to ~2e-5 off: a plan that differs from it
    by a little more than the tolerance is judged against a float64 copy of the model instead of
    being thrown away (an 8-GPU run lost 35 % to one rank whose first two masks landed at 2.1e-5).
    A structurally wrong plan still fails."""
    from torchpruner_b200 import workloads
    from torchpruner_b200.attributions import native_tail
    from torchpruner_b200.attributions.native_tail import NativeTail

    saved = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
    try:
        model, shape = workloads.build("vgg16_cifar")
        model.to(DEV)
        x = torch.randn((32,) + shape, device=DEV, generator=torch.Generator(device=DEV).manual_seed(3))
        module = model.features[9]
        with torch.no_grad():
            z = model.forward_partial(x, to_module=module).float().contiguous()
        tail = NativeTail.build(model, module, z)
        assert tail is not None
        n = z.shape[1]
        perm = torch.randperm(n, device=DEV)
        inv = torch.empty(n, dtype=torch.int32, device=DEV)
        inv[perm] = torch.arange(n, dtype=torch.int32, device=DEV)
        tail.prepare_batch(z)
        hi = torch.empty(tail.stack_elems(32, 2), dtype=tail.stack_dtype, device=DEV)
        lo = torch.empty_like(hi)
        ok, err = tail.verify(model, module, z, perm, inv, 0, 2, hi, lo)
        assert ok and err <= native_tail.VERIFY_RTOL
        # force the second stage: nothing passes the fp32 comparison any more
        monkeypatch.setattr(native_tail, "VERIFY_RTOL", 1e-9)
        calls = []
        real = native_tail._float64_forward
        monkeypatch.setattr(native_tail, "_float64_forward",
                            lambda *a: calls.append(1) or real(*a))
        ok, err64 = tail.verify(model, module, z, perm, inv, 0, 2, hi, lo)
        assert calls and ok and err64 <= 1e-5, (ok, err64)   # engine vs float64, 10 layers deep
        # a wrong plan (one layer's shift off by 1 %) fails both stages
        tail.steps[1]["shift"].mul_(1.01)
        ok, _ = tail.verify(model, module, z, perm, inv, 0, 2, hi, lo)
        assert not ok
        tail.steps[1]["shift"].div_(1.01)
        tail.release()
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = saved


@pytest.mark.parametrize("B,C,H,W,C1,pool_first,mode,pool_after", [
    (3, 8, 8, 8, 32, False, 0, False),
    (2, 6, 8, 8, 16, True, 0, True),     # MaxPool before the layer and after it
    (4, 5, 4, 4, 64, False, 0, True),
    (5, 12, 4, 4, 24, True, 1, False),   # pool -> flatten -> Linear (taps = 4 positions)
    (7, 16, 1, 1, 32, False, 1, False),  # plain Linear on a [B, C] activation
])
def test_incremental_first_layer_kernels(B, C, H, W, C1, pool_first, mode, pool_after):
    """tp_channel_contrib / tp_contrib_sum / tp_prefix_contrib_stack against float64 torch."""
    gen = torch.Generator().manual_seed(B * 100 + C)
    z = torch.relu(torch.randn((B, C, H, W), generator=gen))
    zp = F.max_pool2d(z, 2) if pool_first else z
    Hp, Wp = zp.shape[2], zp.shape[3]
    if mode == 0:
        w = torch.randn((C1, C, 3, 3), generator=gen) / 3.0
        wt = w.permute(1, 2, 3, 0).reshape(C, 9, C1).contiguous()
        per_channel = [F.conv2d(zp[:, c:c + 1].double(), w[:, c:c + 1].double(), padding=1) for c in range(C)]
        want_T = torch.stack([t.permute(0, 2, 3, 1).reshape(-1) for t in per_channel])   # NHWC
        oh, ow = Hp, Wp
    else:
        S = Hp * Wp
        w = torch.randn((C1, C * S), generator=gen) / (C * S) ** 0.5
        wt = w.view(C1, C, S).permute(1, 2, 0).contiguous()
        want_T = torch.stack([(zp[:, c].reshape(B, S).double() @ w.view(C1, C, S)[:, c].double().t()).reshape(-1)
                              for c in range(C)])
        oh, ow = 1, 1
    M1 = want_T.shape[1]
    T = torch.empty(C * M1, device=DEV)
    pixels = _lib.channel_contrib(z.to(DEV), wt.to(DEV), pool_first, mode, T)
    assert pixels * C1 == M1
    got_T = T.view(C, M1).cpu().double()
    assert float((got_T - want_T).abs().max()) <= 1e-6 * float(want_T.abs().max())
    s_all = torch.empty(M1, dtype=torch.float64, device=DEV)
    _lib.contrib_sum(T, C, M1, s_all)
    assert float((s_all.cpu() - got_T.sum(0)).abs().max()) <= 1e-12 * float(want_T.abs().max()) * C

    perm = np.random.RandomState(C).permutation(C)
    perm_dev = torch.from_numpy(perm.astype(np.int64)).to(DEV)
    scale = (torch.rand(C1, generator=gen) + 0.5).to(DEV)
    shift = (torch.randn(C1, generator=gen) * 0.1).to(DEV)
    base = s_all.clone()
    skip = 2   # a sharded rank skips a prefix without producing output
    _lib.prefix_contrib_stack(T, perm_dev, 0, skip, base, scale, shift, pixels, C1, oh, ow, pool_after, True)
    P = C - skip
    m_out = (pixels // 4 if pool_after else pixels) * C1
    hi = torch.empty(P * m_out, device=DEV)
    lo = torch.empty_like(hi)
    assert _lib.prefix_contrib_stack(T, perm_dev, skip, P, base, scale, shift, pixels, C1, oh, ow,
                                     pool_after, True, hi, lo) == m_out
    got = (hi + lo).view(P, -1).cpu().double()
    run = got_T.sum(0) - got_T[perm[:skip]].sum(0)
    for p in range(P):
        run = run - got_T[perm[skip + p]]
        y = torch.relu(run.view(-1, C1) * scale.cpu().double() + shift.cpu().double())
        if pool_after:
            y = F.max_pool2d(y.view(B, oh, ow, C1).permute(0, 3, 1, 2), 2).permute(0, 2, 3, 1)
        want = y.reshape(-1)
        assert float((got[p] - want).abs().max()) <= 1e-6 * max(float(want.abs().max()), 1e-3)
    # every channel removed: the running sum is back to (numerically) nothing
    assert float(base.abs().max()) <= 1e-12 * float(want_T.abs().max()) * C


@pytest.mark.parametrize("half", [False, True])
@pytest.mark.parametrize("B,C,H,W,C1,pool_first,mode,pool_after", [
    (3, 8, 8, 8, 32, False, 0, False),
    (2, 6, 8, 8, 16, True, 0, True),
    (4, 5, 4, 4, 64, False, 0, True),
    (3, 7, 3, 5, 8, False, 0, False),    # odd spatial size: one pixel per thread
    (2, 4, 2, 2, 128, False, 0, False),  # 2x2 image: every patch entry but the block is padding
    (5, 12, 4, 4, 24, True, 1, False),
    (7, 16, 1, 1, 32, False, 1, False),
    (256, 9, 8, 8, 64, False, 0, True),  # more items than one wave
])
def test_direct_prefix_ladder_matches_stored_contributions(B, C, H, W, C1, pool_first, mode,
                                                           pool_after, half):
    """tp_channel_major + tp_prefix_contrib_direct (contributions recomputed on the fly) against
    tp_channel_contrib + tp_contrib_sum + tp_prefix_contrib_stack: same FMAs in the same order,
    so S_all, the running sums and every output must be identical bit for bit."""
    gen = torch.Generator().manual_seed(B * 100 + C + 1)
    z = torch.relu(torch.randn((B, C, H, W), generator=gen)).to(DEV)
    Hp, Wp = (H // 2, W // 2) if pool_first else (H, W)
    taps = 9 if mode == 0 else Hp * Wp
    wt = (torch.randn((C, taps, C1), generator=gen) / 3.0).to(DEV)
    scale = (torch.rand(C1, generator=gen) - 0.3).to(DEV)
    shift = (torch.randn(C1, generator=gen) * 0.1).to(DEV)
    oh, ow = (Hp, Wp) if mode == 0 else (1, 1)
    pixels = B * oh * ow
    M1 = pixels * C1
    # stored form
    T = torch.empty(C * M1, device=DEV)
    assert _lib.channel_contrib(z, wt, pool_first, mode, T) == pixels
    s_all = torch.empty(M1, dtype=torch.float64, device=DEV)
    _lib.contrib_sum(T, C, M1, s_all)
    # direct form
    zc = _lib.channel_major(z, pool_first)
    want_zc = (F.max_pool2d(z, 2) if pool_first else z).permute(1, 0, 2, 3).contiguous()
    assert torch.equal(zc.view(want_zc.shape), want_zc)
    s_dir = torch.zeros(M1, dtype=torch.float64, device=DEV)
    _lib.prefix_contrib_direct(zc, wt, None, 0, C, s_dir, scale, shift, B, C, Hp, Wp, mode, False,
                               False, sign=1.0)
    assert torch.equal(s_dir, s_all)

    perm = torch.from_numpy(np.random.RandomState(C).permutation(C).astype(np.int64)).to(DEV)
    skip, relu, out_scale = 2, ("leaky", 0.1), 4.0 if half else 1.0
    P = C - skip
    m_out = (pixels // 4 if pool_after else pixels) * C1
    dt = torch.float16 if half else torch.float32
    outs = []
    for direct in (False, True):
        base = s_all.clone()
        hi = torch.zeros(P * m_out, device=DEV, dtype=dt)
        lo = torch.zeros_like(hi)
        if direct:
            _lib.prefix_contrib_direct(zc, wt, perm, 0, skip, base, scale, shift, B, C, Hp, Wp, mode,
                                       pool_after, relu)
            got = _lib.prefix_contrib_direct(zc, wt, perm, skip, P, base, scale, shift, B, C, Hp, Wp,
                                             mode, pool_after, relu, out_hi=hi, out_lo=lo,
                                             out_scale=out_scale)
        else:
            _lib.prefix_contrib_stack(T, perm, 0, skip, base, scale, shift, pixels, C1, oh, ow,
                                      pool_after, relu)
            got = _lib.prefix_contrib_stack(T, perm, skip, P, base, scale, shift, pixels, C1, oh, ow,
                                            pool_after, relu, hi, lo, out_scale=out_scale)
        assert got == m_out
        outs.append((hi, lo, base))
    for a, b in zip(*outs):
        assert torch.equal(a, b)
    assert float(outs[1][0].float().abs().max()) > 0


@pytest.mark.parametrize("half", [False, True])
@pytest.mark.parametrize("B,Hp,Wp,C,C1,pool_after", [
    (3, 8, 8, 9, 32, False),
    (2, 8, 8, 7, 16, True),
    (5, 2, 2, 6, 128, True),
    (7, 1, 1, 12, 32, False),
    (64, 8, 8, 5, 256, True),     # more items than one wave of any configuration
    (64, 8, 8, 5, 256, False),
    (1, 4, 6, 4, 8, True),        # fewer items than a warp
])
def test_prefix_ring_kernel_matches_register_kernel(B, Hp, Wp, C, C1, pool_after, half):
    """The stored-form sweep has two kernels (register loads; cp.async ring in several depth /
    occupancy configurations, picked by size): same arithmetic in the same order, so outputs and
    running sums must be identical bit for bit, with fewer masks than ring stages too."""
    gen = torch.Generator().manual_seed(B * 31 + C)
    pixels = B * Hp * Wp
    M1 = pixels * C1
    m_out = (pixels // 4 if pool_after else pixels) * C1
    T = torch.randn(C * M1, generator=gen).to(DEV)
    s_all = T.view(C, M1).sum(0, dtype=torch.float64)
    scale = (torch.rand(C1, generator=gen) - 0.3).to(DEV)
    shift = (torch.randn(C1, generator=gen) * 0.1).to(DEV)
    perm = torch.from_numpy(np.random.RandomState(C).permutation(C).astype(np.int64)).to(DEV)
    dt = torch.float16 if half else torch.float32

    def sweep(P):
        base = s_all.clone()
        hi = torch.zeros(max(P, 1) * m_out, device=DEV, dtype=dt)
        lo = torch.zeros_like(hi)
        _lib.prefix_contrib_stack(T, perm, 0, 1, base, scale, shift, pixels, C1, Hp, Wp, pool_after,
                                  ("leaky", 0.1))
        _lib.prefix_contrib_stack(T, perm, 1, P, base, scale, shift, pixels, C1, Hp, Wp, pool_after,
                                  ("leaky", 0.1), hi, lo, out_scale=4.0 if half else 1.0)
        return hi, lo, base

    try:
        for P in (C - 1, 1, 2):
            _lib.set_tuning("prefix_variant", 1)
            want = sweep(P)
            assert float(want[0].float().abs().max()) > 0
            for cfg in (1, 2, 3):
                for threads in (0, 32, 96):
                    _lib.set_tuning("prefix_variant", 2)
                    _lib.set_tuning("prefix_ring_cfg", cfg)
                    _lib.set_tuning("prefix_threads", threads)
                    got = sweep(P)
                    for a, b in zip(got, want):
                        assert torch.equal(a, b), (P, cfg, threads)
    finally:
        _lib.set_tuning("prefix_variant", 0)
        _lib.set_tuning("prefix_ring_cfg", 0)
        _lib.set_tuning("prefix_threads", 0)


def test_prefix_sweep_rejects_misaligned_operands():
    """The sweep reads contributions and per-channel vectors with 16-byte accesses: an offset view
    is refused instead of faulting."""
    B, Hp, Wp, C, C1 = 2, 2, 2, 3, 8
    pixels, M1 = B * Hp * Wp, B * Hp * Wp * C1
    T = torch.zeros(C * M1 + 4, device=DEV)
    base = torch.zeros(M1, dtype=torch.float64, device=DEV)
    vec = torch.ones(C1 + 4, device=DEV)
    perm = torch.arange(C, device=DEV)
    hi = torch.zeros(C * M1, device=DEV)
    lo = torch.zeros_like(hi)
    ok = lambda t, sc: _lib.prefix_contrib_stack(t, perm, 0, C, base, sc, vec[:C1], pixels, C1, Hp,
                                                 Wp, False, True, hi, lo)
    assert ok(T[:C * M1], vec[:C1]) == M1
    for variant in (1, 2):
        _lib.set_tuning("prefix_variant", variant)
        try:
            with pytest.raises(_lib.TorchPrunerB200Error, match="aligned"):
                ok(T[1:C * M1 + 1], vec[:C1])
            with pytest.raises(_lib.TorchPrunerB200Error, match="aligned"):
                ok(T[:C * M1], vec[1:C1 + 1])
        finally:
            _lib.set_tuning("prefix_variant", 0)


def test_native_tail_runs_the_reference_mlp():
    """The 784-2024-2024-10 LeakyReLU perceptron of the reference's untrained-network notebook:
    Linear layers whose K is not a multiple of 32, LeakyReLU epilogue, dense incremental first
    layer — engine on vs engine off."""
    import numpy as np
    from torchpruner_b200 import workloads
    from torchpruner_b200.attributions import ShapleyAttributionMetric

    saved = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        torch.manual_seed(0)
        model, shape = workloads.build("mlp")
        model.to(DEV)
        loader = workloads.synthetic_batches(shape, n=24, batch_size=24, seed=6)
        layer = model.fc[1]      # Linear(784, 2024): 2024 units to score
        tables = {}
        for native in ("auto", False):
            np.random.seed(5)
            m = ShapleyAttributionMetric(model, loader, workloads.cross_entropy, DEV, sv_samples=1,
                                         reduction="none")
            m.native_tail = native
            tables[native] = m.run(layer, find_best_evaluation_module=True)
            assert m.last_run_info["native_tail"] is bool(native), m.last_run_info
            if native:
                assert m.last_run_info["incremental_first"] is True
                print("mlp", m.last_run_info["native"])
        diff = np.abs(tables["auto"] - tables[False]).max()
        print("mlp max |native - forward_partial| =", diff, "max |score| =", np.abs(tables[False]).max())
        assert tables[False].shape == (24, 2024) and diff <= 2e-5
    finally:
        torch.backends.cuda.matmul.allow_tf32 = saved


@pytest.mark.parametrize("N,H,W,Cin,Cout,taps,relu", [
    (4, 32, 32, 64, 64, 9, True),
    (3, 16, 16, 128, 256, 9, True),
    (150, 8, 8, 256, 512, 9, False),   # position-major tiles
    (300, 4, 4, 512, 512, 9, True),
    (300, 1, 1, 2048, 2048, 1, True),  # dense form of a conv on 2x2 images
    (130, 1, 1, 512, 10, 1, False),    # logits: fp32 unsplit output
    (70, 1, 1, 520, 128, 1, ("leaky", 0.1)),   # K tail: 520 = 8*64 + 8
])
def test_conv_f16x3_matches_float64(N, H, W, Cin, Cout, taps, relu):
    """fp16x3 flavour: same 11-bit split on fp16 operands with per-tensor power-of-two scales."""
    gen = torch.Generator(device=DEV).manual_seed(N * 1000 + Cin + 1)
    x = torch.randn((N, H, W, Cin), device=DEV, generator=gen) * 3.0
    x = x * (torch.rand((N, H, W, Cin), device=DEV, generator=gen) > 0.3)
    w = torch.randn((Cout, taps * Cin), device=DEV, generator=gen) / (taps * Cin) ** 0.5
    scale = torch.rand(Cout, device=DEV, generator=gen) + 0.5
    shift = torch.randn(Cout, device=DEV, generator=gen) * 0.1
    s_x = 2.0 ** round(np.log2(64.0 / float(x.abs().max())))
    s_w = 2.0 ** round(np.log2(64.0 / float(w.abs().max())))
    x_hi, x_lo = _lib.f16_split(x, s_x)
    w_hi, w_lo = _lib.f16_split(w, s_w)
    assert x_hi.dtype == torch.float16
    rel = ((x_hi.double() + x_lo.double()) / s_x - x.double()).abs().max() / x.abs().max()
    assert float(rel) <= 2.0 ** -21
    split = Cout != 10
    out_scale = 8.0
    y_hi, y_lo = _lib.conv_tf32x3(x_hi, x_lo, w_hi, w_lo, taps, (scale / (s_x * s_w)).contiguous(),
                                  shift, relu, split=split, out_scale=out_scale)
    if split:
        assert y_hi.dtype == torch.float16
        y = (y_hi.double() + y_lo.double()) / out_scale
    else:
        assert y_hi.dtype == torch.float32
        y = y_hi.double()
    want = _ref64(x, w, taps, scale, shift, relu)
    scale_ref = float(want.abs().max())
    err = float((y - want).abs().max())
    print(f"fp16x3 K={taps * Cin} max abs err {err:.3e} (scale {scale_ref:.3f})")
    assert err <= 1e-6 * scale_ref, (err, scale_ref)


@pytest.mark.parametrize("N,H,W,Cin,Cout,relu", [
    (256, 8, 8, 128, 256, True),
    (384, 4, 4, 256, 512, True),
    (200, 4, 4, 64, 128, ("leaky", 0.2)),   # image block cut by the end of the stack
    (128, 2, 2, 64, 64, True),               # BLOCK_N = 64, one window per image
    (1300, 8, 8, 64, 192, False),            # more tiles than SMs; negative outputs under the max
])
def test_conv_f16x3_fused_pool_is_bit_identical(N, H, W, Cin, Cout, relu):
    """tp_conv_f16x3_pool2 (window maximum kept in TMEM) against tp_conv_f16x3 followed by
    tp_maxpool2_split_nhwc_f16: the same partial sums, the same rounding, so the same bits."""
    assert _lib.conv_pool2_fusable(N, H, W, Cin, Cout)
    gen = torch.Generator(device=DEV).manual_seed(N + 7 * Cin)
    x = torch.randn((N, H, W, Cin), device=DEV, generator=gen) * 2.0
    w = torch.randn((Cout, 9 * Cin), device=DEV, generator=gen) / (9 * Cin) ** 0.5
    scale = torch.rand(Cout, device=DEV, generator=gen) + 0.5
    scale[::3] *= -1.0                        # BatchNorm weights may be negative
    shift = torch.randn(Cout, device=DEV, generator=gen) * 0.3
    s_x = 2.0 ** round(np.log2(64.0 / float(x.abs().max())))
    s_w = 2.0 ** round(np.log2(64.0 / float(w.abs().max())))
    x_hi, x_lo = _lib.f16_split(x, s_x)
    w_hi, w_lo = _lib.f16_split(w, s_w)
    sc = (scale / (s_x * s_w)).contiguous()
    out_scale = 4.0
    raw, none = _lib.conv_tf32x3(x_hi, x_lo, w_hi, w_lo, 9, sc, shift, relu, split=False)
    assert none is None and raw.dtype == torch.float32
    want_hi, want_lo = _lib.maxpool2_split_nhwc(raw, half=True, scale=out_scale)
    got_hi, got_lo = _lib.conv_tf32x3(x_hi, x_lo, w_hi, w_lo, 9, sc, shift, relu, split=True,
                                      out_scale=out_scale, pool=True)
    assert got_hi.shape == (N, H // 2, W // 2, Cout) and got_hi.dtype == torch.float16
    assert torch.equal(got_hi, want_hi) and torch.equal(got_lo, want_lo)
    # and against the float64 reference of conv -> affine -> activation -> max pool
    ref = _ref64(x, w, 9, scale, shift, relu).view(N, H // 2, 2, W // 2, 2, Cout).amax(dim=(2, 4))
    y = (got_hi.double() + got_lo.double()) / out_scale
    assert float((y - ref).abs().max()) <= 1e-6 * float(ref.abs().max())


@pytest.mark.parametrize("N,Cin,Cout,relu", [
    (256, 2048, 512, True),                  # VGG16's conv13 at bs 256: CTA pairs, 4 N tiles
    (130, 256, 64, ("leaky", 0.2)),          # BLOCK_N = 64, ragged second image block
    (100, 64, 128, False),                   # one image block (no pair), negative outputs
    (1500, 512, 192, True),                  # more tiles than SMs, Cout not a multiple of 128
])
def test_linear_pool4_is_bit_identical(N, Cin, Cout, relu):
    """tp_linear_f16x3_pool4 (four output groups back to back, maximum kept in TMEM) against
    tp_conv_f16x3 on the [4*Cout, Cin] weight followed by tp_maxpool2_split_nhwc_f16 on the
    [N, 2, 2, Cout] view: the same partial sums and rounding, so the same bits."""
    gen = torch.Generator(device=DEV).manual_seed(N + 7 * Cin)
    x = torch.randn((N, 1, 1, Cin), device=DEV, generator=gen) * 2.0
    w = torch.randn((4 * Cout, Cin), device=DEV, generator=gen) / Cin ** 0.5
    scale = torch.rand(Cout, device=DEV, generator=gen) + 0.5
    scale[::3] *= -1.0
    shift = torch.randn(Cout, device=DEV, generator=gen) * 0.3
    s_x = 2.0 ** round(np.log2(64.0 / float(x.abs().max())))
    s_w = 2.0 ** round(np.log2(64.0 / float(w.abs().max())))
    x_hi, x_lo = _lib.f16_split(x, s_x)
    w_hi, w_lo = _lib.f16_split(w, s_w)
    sc = (scale / (s_x * s_w)).contiguous()
    out_scale = 4.0
    raw, _ = _lib.conv_tf32x3(x_hi, x_lo, w_hi, w_lo, 1, sc.repeat(4), shift.repeat(4), relu,
                              split=False)
    want_hi, want_lo = _lib.maxpool2_split_nhwc(raw.view(N, 2, 2, Cout), half=True, scale=out_scale)
    got_hi, got_lo = _lib.linear_pool4(x_hi, x_lo, w_hi, w_lo, sc, shift, relu, out_scale=out_scale)
    assert got_hi.shape == (N, 1, 1, Cout) and got_hi.dtype == torch.float16
    assert torch.equal(got_hi, want_hi) and torch.equal(got_lo, want_lo)
    ref = _ref64(x, w, 1, scale.repeat(4), shift.repeat(4), relu).view(N, 4, Cout).amax(dim=1)
    y = (got_hi.double() + got_lo.double()).view(N, Cout) / out_scale
    assert float((y - ref).abs().max()) <= 1e-6 * float(ref.abs().max())
    for pair in (1, 2):      # single-CTA tiles and CTA pairs
        _lib.set_tuning("conv_pair", pair)
        try:
            again = _lib.linear_pool4(x_hi, x_lo, w_hi, w_lo, sc, shift, relu, out_scale=out_scale)
        finally:
            _lib.set_tuning("conv_pair", 0)
        assert torch.equal(again[0], want_hi) and torch.equal(again[1], want_lo)


def test_conv_pool2_fusable_limits():
    assert not _lib.conv_pool2_fusable(64, 4, 4, 64, 64)      # fewer than 128 images
    assert not _lib.conv_pool2_fusable(256, 32, 32, 64, 64)   # row-major tiles
    assert not _lib.conv_pool2_fusable(256, 4, 4, 32, 64)     # Cin % 64
    assert not _lib.conv_pool2_fusable(256, 4, 4, 64, 96)     # Cout % 64
    x = torch.zeros((256, 32, 32, 64), device=DEV, dtype=torch.float16)
    w = torch.zeros((64, 576), device=DEV, dtype=torch.float16)
    one = torch.ones(64, device=DEV)
    with pytest.raises(_lib.TorchPrunerB200Error, match="position-major"):
        _lib.conv_tf32x3(x, x, w, w, 9, one, one, True, split=True, pool=True)


@pytest.mark.parametrize("half", [True, False], ids=["f16x3", "tf32x3"])
@pytest.mark.parametrize("N,H,W,Cin,Cout,taps,pool", [
    (8, 32, 32, 64, 64, 9, False),       # row tiles, 8 image blocks of one image, BLOCK_N = 64
    (5, 16, 16, 64, 128, 9, False),      # odd image count: the last pair's second CTA is outside
    (512, 4, 4, 128, 256, 9, False),     # position-major, 4 image blocks, two N tiles
    (384, 4, 4, 64, 192, 9, False),      # 3 image blocks (odd), Cout not a multiple of 128
    (300, 8, 8, 64, 128, 9, True),       # fused MaxPool (running maximum in TMEM), ragged block
    (640, 2, 2, 128, 128, 9, False),     # 2x2 position-major: 4 of 9 taps
    (700, 1, 1, 512, 512, 1, False),     # linear layer, ragged rows
    (256, 1, 1, 200, 64, 1, False),      # K tail zero-filled by TMA
])
def test_conv_cta_pairs_give_the_same_bits_as_single_cta_tiles(N, H, W, Cin, Cout, taps, pool, half):
    """256-row tiles on CTA pairs (tcgen05 cta_group::2: each SM holds its own 128 rows of A and
    half of the weight tile) against the single-CTA kernel: every output element is the same
    K-ordered sum on the same operands, so the results are compared bit for bit."""
    if pool and not half:
        pytest.skip("pool fusion exists in the fp16 flavour only")
    gen = torch.Generator(device=DEV).manual_seed(N + 7 * Cin)
    x = torch.randn((N, H, W, Cin), device=DEV, generator=gen)
    x = x * (torch.rand((N, H, W, Cin), device=DEV, generator=gen) > 0.3)
    w = torch.randn((Cout, taps * Cin), device=DEV, generator=gen) / (taps * Cin) ** 0.5
    scale = torch.rand(Cout, device=DEV, generator=gen) + 0.5
    shift = torch.randn(Cout, device=DEV, generator=gen) * 0.1
    if half:
        x_hi, x_lo = _lib.f16_split(x, 4.0)
        w_hi, w_lo = _lib.f16_split(w, 8.0)
        scale = scale / 32.0
    else:
        x_hi, x_lo = _lib.tf32_split(x)
        w_hi, w_lo = _lib.tf32_split(w)
    outs = {}
    try:
        for pair_off, knob in ((1, 1), (0, 2)):
            _lib.set_tuning("conv_pair", knob)
            outs[pair_off] = _lib.conv_tf32x3(x_hi, x_lo, w_hi, w_lo, taps, scale, shift, True,
                                              split=True, pool=pool, out_scale=2.0 if half else 1.0)
    finally:
        _lib.set_tuning("conv_pair", 0)
    torch.cuda.synchronize()
    for a, b in zip(outs[0], outs[1]):
        assert a.shape == b.shape and torch.equal(a, b)
    y = (outs[0][0].float() + outs[0][1].float()) / (2.0 if half else 1.0)
    want = _ref64(x, w, taps, scale * (32.0 if half else 1.0), shift, True)
    if pool:
        want = F.max_pool2d(want.permute(0, 3, 1, 2), 2).permute(0, 2, 3, 1)
    assert float((y.double() - want).abs().max()) <= 2e-6 * max(1.0, float(want.abs().max()))


def test_activation_cache_gives_the_same_bits():
    """Scoring layer after layer on the same batch continues the head from the cached activation
    of the previous layer (one block instead of the whole prefix).  The tables must be the bits a
    cold cache gives; another batch, another model state or a write through .data must not be
    served from it."""
    from torchpruner_b200 import workloads
    from torchpruner_b200.attributions import ShapleyAttributionMetric, native_tail

    saved = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
    try:
        model, shape = workloads.build("vgg16_cifar")
        model.to(DEV)
        loaders = [workloads.synthetic_batches(shape, n=16, batch_size=16, seed=s) for s in (5, 6)]
        layers = [model.features[i] for i in (20, 24, 27, 30)]

        def run(layer, loader):
            np.random.seed(3)
            m = ShapleyAttributionMetric(model, loader, workloads.cross_entropy, DEV, sv_samples=1,
                                         reduction="none")
            with contextlib.redirect_stdout(io.StringIO()):
                sv = m.run(layer, find_best_evaluation_module=True)
            assert m.last_run_info["native_head"]
            return sv

        warm = {}
        for layer in layers:          # plans get built; activations accumulate in the cache
            for li, loader in enumerate(loaders):
                warm[(id(layer), li)] = run(layer, loader)
        entries = native_tail._ACTS[model]
        assert len(entries) == 2 and all(len(e[2]) == 10 for e in entries.values())   # 10 blocks
        launches = []
        for layer in layers[1:]:      # a cached prefix: at most one block is computed per run
            prof = _lib.profile_begin(timed=False)
            again = run(layer, loaders[1])
            launches.append(prof.records.get("contrib_sum_direct", [0])[0])
            _lib.profile_end()
            assert np.array_equal(again, warm[(id(layer), 1)])
        assert launches == [0, 0, 0], launches
        for layer in layers:          # cold cache: the same bits
            native_tail._ACTS.pop(model, None)
            assert np.array_equal(run(layer, loaders[0]), warm[(id(layer), 0)])
        # the same examples in another order are another batch
        from torch.utils.data import DataLoader, TensorDataset
        x0, y0 = next(iter(loaders[0]))
        flipped = DataLoader(TensorDataset(x0.flip(0), y0.flip(0)), batch_size=16, shuffle=False)
        run(layers[2], loaders[0])
        got = run(layers[2], flipped)
        native_tail._ACTS.pop(model, None)
        assert np.array_equal(got, run(layers[2], flipped))
        assert np.array_equal(got, warm[(id(layers[2]), 0)][::-1])
        # a write through .data: the plan's checksum notices, plans AND activations are dropped
        model.features[17].weight.data.mul_(1.5)
        changed = run(layers[1], loaders[0])
        assert np.abs(changed - warm[(id(layers[1]), 0)]).max() > 1e-3
        native_tail._ACTS.pop(model, None)
        assert np.array_equal(run(layers[1], loaders[0]), changed)
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = saved
