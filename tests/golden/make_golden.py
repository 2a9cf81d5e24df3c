"""Generate the golden vectors under tests/golden/ by running the UNMODIFIED reference.

Run in the build container only (the reference tree is not present on the GPU box):

    python tests/golden/make_golden.py            # all groups
    python tests/golden/make_golden.py ops vgg    # selected groups

The reference (/root/reference, read-only) is imported, never copied.  Two harness-side repairs
are applied by monkey-patch, as SURVEY.md 8(c) records:
  * OptimizerPruner._prune_sgd looks state up through optimizer.state_dict()["state"][id(p)],
    a KeyError on torch >= 1.6; replaced by the same slicing through optimizer.state[p].
  * the reference's VGG16 builder calls a torchvision method that no longer exists; the model
    comes from torchpruner_b200.workloads (same topology and forward_partial protocol).
Inputs are regenerated from seeds by the tests (same torch version => same CPU random streams);
only outputs are stored.
"""
import copy
import hashlib
import io
import contextlib
import os
import sys

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.utils.data import DataLoader, TensorDataset

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.dont_write_bytecode = True
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

from torchpruner.attributions import (  # noqa: E402  (the reference)
    APoZAttributionMetric, SensitivityAttributionMetric, ShapleyAttributionMetric,
    TaylorAttributionMetric, WeightNormAttributionMetric)
from torchpruner.pruner import Pruner  # noqa: E402
from torchpruner.pruner.opt_pruner import OptimizerPruner  # noqa: E402
from torchpruner.utils import find_best_module_for_attributions, get_vgg_pruning_graph  # noqa: E402

from torchpruner_b200 import workloads  # noqa: E402
from tests import cases  # noqa: E402


def _patched_prune_sgd(optimizer, param, axis, keep_indices, device):
    for p in optimizer.param_groups[0]["params"]:
        if id(p) == id(param):
            state = optimizer.state.get(p, {})
            if state.get("momentum_buffer", None) is not None:
                momentum = state["momentum_buffer"]
                momentum.data = momentum.data.index_select(axis, keep_indices.to(device))


OptimizerPruner._prune_sgd = staticmethod(_patched_prune_sgd)


class _Holder:
    pass


def quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()):
        return fn(*a, **k)


def sha(t):
    return hashlib.sha256(np.ascontiguousarray(t).tobytes()).hexdigest()


# --------------------------------------------------------------------------- groups
def gen_ops():
    """Tier A: the reference's hook arithmetic on seeded tensors of ragged shapes."""
    out = {}
    for name, shape in cases.OP_SHAPES.items():
        z, g = cases.op_inputs(name)
        m = _Holder()
        m._tp_activation = z
        TaylorAttributionMetric._backward_hook(_Signed(False))(m, None, (g,))
        out[f"{name}/taylor"] = m._tp_taylor
        m = _Holder()
        m._tp_activation = z
        TaylorAttributionMetric._backward_hook(_Signed(True))(m, None, (g,))
        out[f"{name}/taylor_signed"] = m._tp_taylor
        m = _Holder()
        APoZAttributionMetric._forward_hook()(m, None, z)
        out[f"{name}/apoz"] = m._tp_nonzero_count
        m = _Holder()
        SensitivityAttributionMetric._backward_hook()(m, None, (g,))
        out[f"{name}/sensitivity"] = m._tp_gradient
    for name in cases.WEIGHT_SHAPES:
        w = cases.weight_input(name)
        lin = _Holder()
        lin.weight = nn.Parameter(w)
        wn = WeightNormAttributionMetric.__new__(WeightNormAttributionMetric)
        wn.model = None
        # run() type-checks the module; call the arithmetic the way run() does
        attr = np.abs(lin.weight.detach().cpu().numpy())
        while len(attr.shape) > 1:
            attr = attr.sum(-1)
        out[f"{name}/weight_norm"] = attr
    np.savez_compressed(os.path.join(HERE, "ops.npz"), **out)
    print("ops:", len(out), "arrays")


class _Signed:
    def __init__(self, signed):
        self.signed = signed


def gen_max_model():
    """The reference's own analytic fixture (tests/test_attributions.py:19-45) through run()."""
    out = {}
    for version in (1, 2):
        x, y, model = cases.max_model(version)
        loader = DataLoader(TensorDataset(x, y), batch_size=1, shuffle=False)
        layer = list(model.children())[0]
        for best in (False, True):
            tag = f"v{version}/best{int(best)}"
            mk = lambda cls, **kw: cls(model, loader, F.mse_loss, "cpu", **kw)
            out[f"{tag}/weight_norm"] = quiet(mk(WeightNormAttributionMetric).run, layer, find_best_evaluation_module=best)
            out[f"{tag}/apoz"] = quiet(mk(APoZAttributionMetric).run, layer, find_best_evaluation_module=best)
            out[f"{tag}/sensitivity"] = quiet(mk(SensitivityAttributionMetric).run, layer, find_best_evaluation_module=best)
            out[f"{tag}/taylor"] = quiet(mk(TaylorAttributionMetric).run, layer, find_best_evaluation_module=best)
            out[f"{tag}/taylor_signed"] = quiet(mk(TaylorAttributionMetric, signed=True).run, layer, find_best_evaluation_module=best)
            np.random.seed(7)
            out[f"{tag}/shapley_slow"] = quiet(mk(ShapleyAttributionMetric, sv_samples=20, reduction="none").run, layer, find_best_evaluation_module=best)
    np.savez_compressed(os.path.join(HERE, "max_model.npz"), **out)
    print("max_model:", len(out), "arrays")


def gen_small_cnn():
    """Config 1: small CNN, bs 64, N 128: Taylor / APoZ / WeightNorm / Shapley(sv=5)."""
    out = {}
    model, shape = workloads.build("small_cnn")
    loader = workloads.synthetic_batches(shape, n=128, batch_size=64)
    convs = {"conv1": model.features[0], "conv2": model.features[3]}
    for name, layer in convs.items():
        for best in (False, True):
            tag = f"{name}/best{int(best)}"
            mk = lambda cls, **kw: cls(model, loader, workloads.cross_entropy, "cpu", **kw)
            out[f"{tag}/taylor"] = quiet(mk(TaylorAttributionMetric, reduction="none").run, layer, find_best_evaluation_module=best)
            model.zero_grad()
            out[f"{tag}/sensitivity"] = quiet(mk(SensitivityAttributionMetric, reduction="none").run, layer, find_best_evaluation_module=best)
            model.zero_grad()
            out[f"{tag}/apoz"] = quiet(mk(APoZAttributionMetric, reduction="none").run, layer, find_best_evaluation_module=best)
        out[f"{name}/weight_norm"] = quiet(mk(WeightNormAttributionMetric).run, layer)
        np.random.seed(11)
        out[f"{name}/shapley"] = quiet(mk(ShapleyAttributionMetric, sv_samples=5, reduction="none").run, layer, find_best_evaluation_module=True)
    out["fc1/weight_norm"] = quiet(mk(WeightNormAttributionMetric).run, model.classifier[0])
    np.savez_compressed(os.path.join(HERE, "small_cnn.npz"), **out)
    print("small_cnn:", len(out), "arrays")


def gen_vgg():
    """Configs 2/3 on VGG16-CIFAR: mean Taylor/APoZ/Sensitivity for the 13 convs at bs 256, and
    Shapley (sv=2, "none") on a 32-example batch for all 13 layers."""
    out = {}
    model, shape = workloads.build("vgg16_cifar")
    loader = workloads.synthetic_batches(shape, n=256, batch_size=256)
    names = {m: n for n, m in model.named_modules()}
    for layer in model.conv_layers():
        tag = names[layer]
        mk = lambda cls, **kw: cls(model, loader, workloads.cross_entropy, "cpu", **kw)
        out[f"{tag}/taylor"] = quiet(mk(TaylorAttributionMetric).run, layer, find_best_evaluation_module=True)
        model.zero_grad()
        out[f"{tag}/sensitivity"] = quiet(mk(SensitivityAttributionMetric).run, layer, find_best_evaluation_module=True)
        model.zero_grad()
        out[f"{tag}/apoz"] = quiet(mk(APoZAttributionMetric).run, layer, find_best_evaluation_module=True)
        out[f"{tag}/weight_norm"] = quiet(mk(WeightNormAttributionMetric).run, layer)
        print("  vgg", tag, flush=True)
    # Config 3 at a size the CPU finishes in minutes: Shapley (sv_samples=2, "none") on a
    # 32-example batch for ALL 13 layers, from the reference in fp32 (the golden) and from the
    # same reference code on a float64 copy of the model and data (the arbiter: how far the
    # reference's own fp32 arithmetic is from the exact value of the same estimator).
    small = workloads.synthetic_batches(shape, n=32, batch_size=32, seed=5)
    x32, y32 = next(iter(small))
    twin = copy.deepcopy(model).double()
    small64 = DataLoader(TensorDataset(x32.double(), y32), batch_size=32, shuffle=False)
    twin_modules = dict(twin.named_modules())
    for layer in model.conv_layers():
        tag = names[layer]
        np.random.seed(3)
        m = ShapleyAttributionMetric(model, small, workloads.cross_entropy, "cpu", sv_samples=2, reduction="none")
        out[f"{tag}/shapley32"] = quiet(m.run, layer, find_best_evaluation_module=True)
        np.random.seed(3)
        m = ShapleyAttributionMetric(twin, small64, workloads.cross_entropy, "cpu", sv_samples=2, reduction="none")
        out[f"{tag}/shapley32_f64"] = quiet(m.run, twin_modules[tag], find_best_evaluation_module=True)
        a, b = out[f"{tag}/shapley32"], out[f"{tag}/shapley32_f64"]
        print("  vgg shapley", tag, "max|score| %.3e  reference fp32 vs fp64 twin: %.2e of max" % (
            np.abs(b).max(), np.abs(a - b).max() / np.abs(b).max()), flush=True)
    np.savez_compressed(os.path.join(HERE, "vgg16.npz"), **out)
    print("vgg16:", len(out), "arrays")


def gen_mlp():
    """The 784-2024-2024-10 LeakyReLU perceptron of the "Pruning Untrained Networks" notebook:
    Taylor / Sensitivity / APoZ / WeightNorm and Shapley (sv=1) of both hidden layers on 16
    examples."""
    out = {}
    model, shape = workloads.build("mlp")
    loader = workloads.synthetic_batches(shape, n=16, batch_size=16, seed=6)
    for tag, layer in (("fc1", model.fc[1]), ("fc2", model.fc[3])):
        mk = lambda cls, **kw: cls(model, loader, workloads.cross_entropy, "cpu", **kw)
        out[f"{tag}/taylor"] = quiet(mk(TaylorAttributionMetric, reduction="none").run, layer, find_best_evaluation_module=True)
        model.zero_grad()
        out[f"{tag}/sensitivity"] = quiet(mk(SensitivityAttributionMetric, reduction="none").run, layer, find_best_evaluation_module=True)
        model.zero_grad()
        out[f"{tag}/apoz"] = quiet(mk(APoZAttributionMetric, reduction="none").run, layer, find_best_evaluation_module=True)
        out[f"{tag}/weight_norm"] = quiet(mk(WeightNormAttributionMetric).run, layer)
        np.random.seed(13)
        out[f"{tag}/shapley16"] = quiet(mk(ShapleyAttributionMetric, sv_samples=1, reduction="none").run, layer, find_best_evaluation_module=True)
        print("  mlp", tag, flush=True)
    np.savez_compressed(os.path.join(HERE, "mlp.npz"), **out)
    print("mlp:", len(out), "arrays")


def gen_graph():
    """Host logic: evaluation-module choice and the VGG cascade graph, as module names."""
    lines = []
    model, _ = workloads.build("vgg16_cifar")
    names = {m: n for n, m in model.named_modules()}
    for layer in model.conv_layers() + [m for m in model.classifier if isinstance(m, nn.Linear)]:
        best = quiet(find_best_module_for_attributions, model, layer)
        lines.append(f"best {names[layer]} -> {names[best]}")
    for module, cascade in get_vgg_pruning_graph(model):
        lines.append(f"graph {names[module]} -> {','.join(names[c] for c in cascade)}")
    seq = nn.Sequential(nn.Linear(3, 2), nn.BatchNorm1d(2), nn.ReLU(), nn.Linear(2, 1))
    lines.append("seq0 -> %d" % list(seq).index(quiet(find_best_module_for_attributions, seq, seq[0])))
    tail = nn.Sequential(nn.Linear(3, 2), nn.ReLU())
    lines.append("tail0 -> %d" % list(tail).index(quiet(find_best_module_for_attributions, tail, tail[0])))
    with open(os.path.join(HERE, "graph.txt"), "w") as f:
        f.write("\n".join(lines) + "\n")
    print("graph:", len(lines), "lines")


def gen_prune():
    """Config 5: cascading prune of VGG16-CIFAR with SGD momentum, 25 % of every layer in graph
    order; SHA-256 of every resulting tensor (the tensors total ~40 MB, the digests 10 KB).
    Also config 1's graph on the small CNN."""
    lines = []
    for cfg in ("vgg16_cifar", "small_cnn"):
        model, shape, opt = cases.prune_setup(cfg, device="cpu")
        pruner = Pruner(model, shape, "cpu", optimizer=opt)
        graph = get_vgg_pruning_graph(model) if cfg == "vgg16_cifar" else model.get_pruning_graph()
        np.random.seed(21)
        for module, cascade in graph:
            idx = cases.prune_choice(module.weight.shape[0])
            quiet(pruner.prune_model, module, idx, cascading_modules=cascade)
        lines += [f"{cfg} {line}" for line in cases.model_digest(model, opt)]
    with open(os.path.join(HERE, "prune.txt"), "w") as f:
        f.write("\n".join(lines) + "\n")
    print("prune:", len(lines), "lines")


def gen_ablation():
    """The notebook's ablation loop (cell 8) run literally with the reference's helpers."""
    out = {}
    for cfg, tag, n in (("small_cnn", "features.3", 32), ("vgg16_cifar", "features.34", 32)):
        model, shape = workloads.build(cfg)
        x, y = next(iter(workloads.synthetic_batches(shape, n=n, batch_size=n, seed=8)))
        module = dict(model.named_modules())[tag]
        ranking = np.random.RandomState(4).permutation(module.out_channels)
        loss = workloads.cross_entropy
        losses, accs = [], []
        with torch.no_grad():
            module_nl = quiet(find_best_module_for_attributions, model, module)
            z = model.forward_partial(x, to_module=module_nl)
            for i in ranking:
                z.index_fill_(1, torch.tensor([i]), 0.)
                y_pred = model.forward_partial(z, from_module=module_nl)
                new_loss = loss(y_pred, y)
                current_pred = y_pred.argmax(dim=1, keepdim=True)
                new_acc = current_pred.eq(y.view_as(current_pred)).sum().float() / len(y)
                losses.append(new_loss.item())
                accs.append(new_acc.item())
        out[f"{cfg}/loss"] = np.asarray(losses, np.float32)
        out[f"{cfg}/acc"] = np.asarray(accs, np.float32)
    np.savez_compressed(os.path.join(HERE, "ablation.npz"), **out)
    print("ablation:", len(out), "arrays")


def gen_resnet():
    """Config 4: the reference on STOCK torchvision ResNet-50 with find_best_evaluation_module=True
    (8 examples of 3x96x96 in two batches; 96 leaves 3x3 = 9-float rows in the last stage, which are
    not 16-byte aligned).  For each of the 53 convolutions: Taylor / APoZ / Sensitivity per-example
    tables where the reference is defined, the exception class where it is not."""
    out, lines = {}, []
    model, _ = workloads.build("resnet50_stock")
    loader = workloads.synthetic_batches((3, 96, 96), n=8, batch_size=4, seed=9)
    names = {m: n for n, m in model.named_modules()}
    convs = [m for m in model.modules() if isinstance(m, nn.Conv2d)]
    assert len(convs) == 53
    kinds = (("taylor", TaylorAttributionMetric), ("apoz", APoZAttributionMetric),
             ("sensitivity", SensitivityAttributionMetric))
    for conv in convs:
        tag = names[conv]
        best = quiet(find_best_module_for_attributions, model, conv)
        for kind, cls in kinds:
            for m in model.modules():   # a failed run leaves hooks' attributes behind
                for attr in ("_tp_taylor", "_tp_activation", "_tp_nonzero_count", "_tp_gradient"):
                    if hasattr(m, attr):
                        delattr(m, attr)
                m._forward_hooks.clear()
                m._backward_hooks.clear()
            model.zero_grad()
            try:
                metric = cls(model, loader, workloads.cross_entropy, "cpu", reduction="none")
                out[f"{tag}/{kind}"] = quiet(metric.run, conv, find_best_evaluation_module=True)
                lines.append(f"{tag} {kind} ok eval={names[best]} shape={out[f'{tag}/{kind}'].shape}")
            except Exception as exc:  # noqa: BLE001 - the class of the failure IS the golden
                lines.append(f"{tag} {kind} raises {type(exc).__name__} eval={names[best]}")
        print("  resnet", lines[-1], flush=True)
    np.savez_compressed(os.path.join(HERE, "resnet50.npz"), **out)
    with open(os.path.join(HERE, "resnet50.txt"), "w") as f:
        f.write("\n".join(lines) + "\n")
    print("resnet50:", len(out), "arrays,", sum(" raises " in line for line in lines), "failures")


GROUPS = {"resnet": gen_resnet, "ablation": gen_ablation, "ops": gen_ops, "max_model": gen_max_model, "small_cnn": gen_small_cnn, "vgg": gen_vgg,
          "mlp": gen_mlp, "graph": gen_graph, "prune": gen_prune}

if __name__ == "__main__":
    torch.set_num_threads(8)
    for g in (sys.argv[1:] or list(GROUPS)):
        GROUPS[g]()
