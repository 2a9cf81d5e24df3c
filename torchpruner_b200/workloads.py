"""Synthetic workloads of the BASELINE.json configs: model topologies, seeded data, criterion.

These are the shapes the hot path is measured and tested on, not part of the scoring library
itself.  Topologies follow the reference's experiment models (experiments/models/cifar10.py,
fmnist.py) so that layer indices, module order and ``forward_partial`` semantics are the ones
the reference's notebooks use; weights are random (there are no checkpoints offline).
"""
import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.utils.data import DataLoader, TensorDataset

VGG16_CFG = [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"]


def cross_entropy(output, target, reduction="mean"):
    """Criterion protocol of the reference (experiments/models/cifar10.py:90-91)."""
    return F.cross_entropy(output, target, reduction=reduction)


class _PartialSequential(nn.Module):
    """Model made of ``features`` -> flatten -> ``classifier`` with the reference's
    ``forward_partial(x, to_module, from_module)`` protocol (cifar10.py:39-59): run the children
    strictly after ``from_module`` up to and including ``to_module``."""

    def forward_partial(self, x, to_module=None, from_module=None):
        active = from_module is None
        for stage in (self.features, None, self.classifier):
            if stage is None:
                x = torch.flatten(x, 1)
                continue
            for layer in stage.children():
                if active:
                    x = layer(x)
                    if layer is to_module:
                        return x
                elif layer is from_module:
                    active = True
        return x

    def forward(self, x):
        return self.forward_partial(x)


class VGG16Cifar(_PartialSequential):
    """VGG16-BN for 3x32x32 inputs with the 512-512-10 classifier of cifar10.py:62-77.
    ``features`` has torchvision's vgg16_bn layout (features.0 conv, .1 BN, .2 ReLU, ...)."""

    def __init__(self, num_classes=10):
        super().__init__()
        layers, c_in = [], 3
        for v in VGG16_CFG:
            if v == "M":
                layers.append(nn.MaxPool2d(kernel_size=2, stride=2))
            else:
                layers += [nn.Conv2d(c_in, v, kernel_size=3, padding=1), nn.BatchNorm2d(v),
                           nn.ReLU(inplace=True)]
                c_in = v
        self.features = nn.Sequential(*layers)
        self.classifier = nn.Sequential(
            nn.Dropout(), nn.Linear(512, 512), nn.ReLU(True),
            nn.Dropout(), nn.Linear(512, 512), nn.ReLU(True),
            nn.Linear(512, num_classes),
        )
        for m in self.modules():  # torchvision's VGG initialisation
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
                nn.init.constant_(m.bias, 0)
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)
            elif isinstance(m, nn.Linear):
                # torchvision draws the classifier from N(0, 0.01): every logit is then ~1e-4 and
                # every loss ln(10) +- 1e-4, so Shapley scores (differences of losses) are 1e-5
                # numbers that no parity test can tell from zero.  The reference is meant for
                # trained networks (logits of a few units); with no checkpoint offline, He
                # scaling gives the classifier that dynamic range.
                nn.init.kaiming_normal_(m.weight, mode="fan_in", nonlinearity="relu")
                nn.init.constant_(m.bias, 0)

    def conv_layers(self):
        return [m for m in self.features if isinstance(m, nn.Conv2d)]


class SmallCNN(_PartialSequential):
    """2-conv + 2-fc CNN for 1x28x28 (topology of experiments/models/fmnist.py:12-30)."""

    def __init__(self, num_classes=10):
        super().__init__()
        self.features = nn.Sequential(
            nn.Conv2d(1, 32, kernel_size=5, stride=1, padding=2), nn.ReLU(), nn.MaxPool2d(2),
            nn.Conv2d(32, 64, kernel_size=3, stride=1, padding=2), nn.ReLU(), nn.MaxPool2d(2),
        )
        self.classifier = nn.Sequential(nn.Linear(4096, 4096), nn.ReLU(), nn.Linear(4096, num_classes))

    def get_pruning_graph(self):
        conv1, conv2 = self.features[0], self.features[3]
        fc1, fc2 = self.classifier[0], self.classifier[2]
        return [(fc1, [fc2]), (conv2, [fc1]), (conv1, [conv2])]


class MLP(nn.Module):
    """784-2024-2024-10 LeakyReLU perceptron of the reference's "Pruning Untrained Networks"
    notebook (experiments/models/mnist.py:9-35), with its ``forward_partial`` protocol: children
    of ``fc`` strictly after ``from_module`` up to and including ``to_module``."""

    def __init__(self, in_features=784, width=2024, num_classes=10):
        super().__init__()
        self.fc = nn.Sequential(nn.Flatten(1), nn.Linear(in_features, width), nn.LeakyReLU(),
                                nn.Linear(width, width), nn.LeakyReLU(), nn.Linear(width, num_classes))

    def forward_partial(self, x, from_module=None, to_module=None):
        active = from_module is None
        for layer in self.fc.children():
            if active:
                x = layer(x)
                if layer is to_module:
                    return x
            elif layer is from_module:
                active = True
        return x

    def forward(self, x):
        return self.forward_partial(x)


def _bottleneck(c_in, width, stride):
    c_out = width * 4
    main = nn.Sequential(
        nn.Conv2d(c_in, width, 1, bias=False), nn.BatchNorm2d(width), nn.ReLU(inplace=True),
        nn.Conv2d(width, width, 3, stride=stride, padding=1, bias=False), nn.BatchNorm2d(width),
        nn.ReLU(inplace=True),
        nn.Conv2d(width, c_out, 1, bias=False), nn.BatchNorm2d(c_out),
    )
    skip = None
    if stride != 1 or c_in != c_out:
        skip = nn.Sequential(nn.Conv2d(c_in, c_out, 1, stride=stride, bias=False),
                             nn.BatchNorm2d(c_out))
    return main, skip, c_out


class _Residual(nn.Module):
    def __init__(self, c_in, width, stride):
        super().__init__()
        self.main, self.skip, self.c_out = _bottleneck(c_in, width, stride)
        self.relu = nn.ReLU(inplace=True)

    def forward(self, x):
        identity = x if self.skip is None else self.skip(x)
        return self.relu(self.main(x) + identity)


class ResNet50HookFriendly(nn.Module):
    """ResNet-50 (3x224x224) in which every activation use is its own module declared in call
    order, so hook-based metrics are defined for all 53 convolutions (on stock
    torchvision ResNet-50 the block's single ``relu`` fires three times per forward with
    different channel counts and the reference raises for every ``conv3``)."""

    def __init__(self, num_classes=10, widths=(64, 128, 256, 512), depths=(3, 4, 6, 3)):
        super().__init__()
        self.stem = nn.Sequential(nn.Conv2d(3, 64, 7, stride=2, padding=3, bias=False),
                                  nn.BatchNorm2d(64), nn.ReLU(inplace=True),
                                  nn.MaxPool2d(3, stride=2, padding=1))
        blocks, c_in = [], 64
        for i, (w, d) in enumerate(zip(widths, depths)):
            for b in range(d):
                blk = _Residual(c_in, w, stride=2 if (b == 0 and i > 0) else 1)
                blocks.append(blk)
                c_in = blk.c_out
        self.blocks = nn.Sequential(*blocks)
        self.pool = nn.AdaptiveAvgPool2d(1)
        self.fc = nn.Linear(c_in, num_classes)

    def forward(self, x):
        x = self.blocks(self.stem(x))
        return self.fc(torch.flatten(self.pool(x), 1))

    def conv_layers(self):
        return [m for m in self.modules() if isinstance(m, nn.Conv2d)]


def warm_batchnorm(model, shape, n=64, seed=1234):
    """One train-mode forward on ``n`` random examples with BatchNorm momentum 1, so the running
    statistics ARE that batch's statistics (with the default momentum 0.1 one forward moves them a
    tenth of the way from (0, 1): eval-mode activations then shrink layer by layer, to 0.03 after
    the thirteenth), then eval mode."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn((n,) + tuple(shape), generator=g)
    dev = next(model.parameters()).device
    norms = [m for m in model.modules() if isinstance(m, nn.modules.batchnorm._BatchNorm)]
    momenta = [m.momentum for m in norms]
    for m in norms:
        m.momentum = 1.0
    model.train()
    with torch.no_grad():
        model(x.to(dev))
    model.eval()
    for m, mom in zip(norms, momenta):
        m.momentum = mom
    return model


def synthetic_batches(shape, n, batch_size, num_classes=10, seed=1, pin=False):
    """DataLoader over seeded x ~ N(0,1), y ~ U{0..num_classes-1} host tensors."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn((n,) + tuple(shape), generator=g)
    y = torch.randint(0, num_classes, (n,), generator=g)
    if pin and torch.cuda.is_available():
        x, y = x.pin_memory(), y.pin_memory()
    return DataLoader(TensorDataset(x, y), batch_size=batch_size, shuffle=False)


def build(name, seed=1):
    """(model on CPU in eval mode with warmed BatchNorm, input shape) for a named config."""
    torch.manual_seed(seed)
    np.random.seed(seed)
    if name == "vgg16_cifar":
        return warm_batchnorm(VGG16Cifar(), (3, 32, 32)), (3, 32, 32)
    if name == "small_cnn":
        model = SmallCNN()
        model.eval()
        return model, (1, 28, 28)
    if name == "mlp":
        model = MLP()
        model.eval()
        return model, (1, 28, 28)
    if name == "resnet50_stock":
        # torchvision's ResNet-50 as it ships: ONE ReLU module per bottleneck, fired three times
        # per forward with different channel counts.  The reference's hook metrics are defined
        # on 37 of its 53 convolutions and raise on every conv3; config 4's
        # parity set is exactly that behaviour.
        import torchvision

        model = torchvision.models.resnet50(num_classes=10)
        return warm_batchnorm(model, (3, 224, 224), n=4), (3, 224, 224)
    if name == "resnet50":
        return warm_batchnorm(ResNet50HookFriendly(), (3, 224, 224), n=8), (3, 224, 224)
    raise KeyError(name)
