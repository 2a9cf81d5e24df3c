"""Seeded test cases shared by the golden generator (tests/golden/make_golden.py), the oracle
tests and the GPU parity tests.  Everything is regenerated from seeds with torch's CPU generator
(deterministic for a given torch build), so only reference OUTPUTS are stored as fixtures."""
import hashlib

import numpy as np
import torch
import torch.nn as nn

from torchpruner_b200 import workloads

# name -> shape of the activation / gradient pair.  Covers: 2-D (Linear) outputs, HW = 1, rows
# that are not a multiple of 4 floats (7x7 = 49, 3x3 = 9, 5), short aligned rows (2x2, 4x4),
# medium rows (8x8, 16x16), a row longer than one shared-memory tile (72x72 = 5184 > 4096),
# ragged row counts, and a single row.
OP_SHAPES = {
    "linear_2d": (6, 10),
    "hw1": (3, 7, 1, 1),
    "hw4": (5, 12, 2, 2),
    "hw9": (4, 6, 3, 3),
    "hw16": (7, 33, 4, 4),
    "hw49": (3, 19, 7, 7),
    "hw64": (8, 16, 8, 8),
    "hw5_1d": (9, 11, 5),
    "hw256": (2, 24, 16, 16),
    "hw1024": (2, 5, 32, 32),
    "hw5184": (2, 3, 72, 72),
    "hw196": (2, 13, 14, 14),
    "single": (1, 1, 6, 6),
}

WEIGHT_SHAPES = {
    "conv_3x3": (20, 7, 3, 3),
    "conv_first": (64, 3, 3, 3),
    "conv_1x1": (24, 16, 1, 1),
    "linear": (10, 37),
    "linear_wide": (5, 4608),
}


def _gen(name):
    seed = int.from_bytes(hashlib.sha256(name.encode()).digest()[:4], "little")
    return torch.Generator().manual_seed(seed)


def op_inputs(name):
    """(z, g): ReLU-like activation with exact zeros and a gradient of mean-loss scale."""
    shape = OP_SHAPES[name]
    gen = _gen("op/" + name)
    z = torch.relu(torch.randn(shape, generator=gen))
    g = torch.randn(shape, generator=gen) / 256.0
    return z, g


def weight_input(name):
    return torch.randn(WEIGHT_SHAPES[name], generator=_gen("w/" + name))


def max_model(version=1, device="cpu"):
    """The reference's analytic fixture (torchpruner/tests/test_attributions.py:19-45): a
    2 -> 4 -> 1 ReLU network that computes max(x0, x1) on four symmetric inputs; in version 2
    hidden unit D gets an outgoing weight of -0.1."""
    x = torch.tensor([[0., 1.], [1., 0.], [1., 2.], [2., 1.]])
    y = x.max(dim=1, keepdim=True).values
    hidden = torch.tensor([[-0.5, 0.5], [1.0, -1.0], [1.0, 1.0], [1.0, 1.0]])  # [4 units, 2 in]
    readout = torch.tensor([[1.0, 0.5, 0.5, 0.0 if version == 1 else -0.1]])   # [1, 4]
    first = nn.Linear(2, 4, bias=False)
    second = nn.Linear(4, 1, bias=False)
    first.weight.data = hidden.clone()
    second.weight.data = readout.clone()
    model = nn.Sequential(first, nn.ReLU(), second).to(device)
    return x.to(device), y.to(device), model


# ------------------------------------------------------------------ pruning (config 5 / 1)
def prune_setup(cfg, device):
    """Model + SGD(momentum) with gradients, momentum buffers and BatchNorm statistics filled
    from the seeded generator.  (A real training step would make this state depend on the host
    CPU's convolution kernels; synthesising it keeps the bit-exact digests host independent.)"""
    torch.manual_seed(1)
    np.random.seed(1)
    if cfg == "vgg16_cifar":
        model, shape = workloads.VGG16Cifar(), (3, 32, 32)
    elif cfg == "small_cnn":
        model, shape = workloads.SmallCNN(), (1, 28, 28)
    else:
        raise KeyError(cfg)
    model.eval()
    gen = _gen("prune/" + cfg)
    for m in model.modules():
        if isinstance(m, nn.BatchNorm2d):
            m.running_mean.copy_(torch.randn(m.running_mean.shape, generator=gen) * 0.1)
            m.running_var.copy_(torch.rand(m.running_var.shape, generator=gen) + 0.5)
            m.weight.data.copy_(torch.rand(m.weight.shape, generator=gen) + 0.5)
    grads = [torch.randn(p.shape, generator=gen) * 1e-2 for p in model.parameters()]
    moms = [torch.randn(p.shape, generator=gen) * 1e-2 for p in model.parameters()]
    model.to(device)
    opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9, weight_decay=5e-4)
    for p, g, m in zip(model.parameters(), grads, moms):
        p.grad = g.to(device)
        opt.state[p]["momentum_buffer"] = m.to(device)
    return model, shape, opt


def prune_choice(n):
    """A seeded (global numpy RNG) quarter of the units, ascending python ints."""
    k = max(1, n // 4)
    return sorted(int(i) for i in np.random.choice(n, size=k, replace=False))


def _sha(t):
    return hashlib.sha256(np.ascontiguousarray(t.detach().cpu().numpy()).tobytes()).hexdigest()[:32]


def model_digest(model, opt=None):
    """One line per tensor that pruning may touch: name, shape, SHA-256 of its bytes."""
    lines = []
    for name, t in model.state_dict().items():
        lines.append(f"state {name} {tuple(t.shape)} {_sha(t)}")
    for name, p in model.named_parameters():
        if p.grad is not None:
            lines.append(f"grad {name} {tuple(p.grad.shape)} {_sha(p.grad)}")
        if opt is not None and "momentum_buffer" in opt.state.get(p, {}):
            buf = opt.state[p]["momentum_buffer"]
            lines.append(f"momentum {name} {tuple(buf.shape)} {_sha(buf)}")
    for name, m in model.named_modules():
        if isinstance(m, nn.Dropout):
            lines.append(f"dropout {name} {m.p!r}")
    return lines
