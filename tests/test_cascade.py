"""Analytic cascade resolution (torchpruner_b200/pruner/cascade.py) against the NaN trick of the
reference (pruner.py:21-54, 129-185) as restated by the oracle — on CPU, no kernel launched.

For every (module, cascade) of the reference's pruning graphs the analytic answer must equal what
the probe forward finds: the same ``_activation_len`` and the same ascending int64 positions.
Where the trace cannot prove the answer it must say so (None) instead of guessing."""
import numpy as np
import pytest
import torch
import torch.nn as nn

from oracle import tp_oracle as oracle
from torchpruner_b200 import workloads
from torchpruner_b200.pruner import cascade
from torchpruner_b200.utils import get_vgg_pruning_graph


def probe(model, module, indices, cascading_modules, input_size, seed=0):
    """The reference's discovery: poison, forward a random batch of 2, look where NaNs arrive."""
    found = {}

    def nanify(_, __, output):
        return output.index_fill_(1, torch.tensor(indices), torch.tensor(np.nan))

    def detect(mod, inp, __):
        found[id(mod)] = (float(inp[0].shape[1]), oracle.nan_indices(inp[0].detach()))

    handles = [module.register_forward_hook(nanify)]
    handles += [m.register_forward_hook(detect) for m in cascading_modules]
    rng = np.random.RandomState(seed)
    try:
        with torch.no_grad():
            model(torch.tensor(rng.random_sample((2,) + tuple(input_size))).float())
    finally:
        for h in handles:
            h.remove()
    return found


def assert_same(analytic, probed):
    assert analytic is not None
    assert set(analytic) == set(probed)
    for key, (length, dead) in probed.items():
        a_len, a_dead = analytic[key]
        assert a_len == length and type(a_len) is float
        assert a_dead.dtype == np.int64
        np.testing.assert_array_equal(a_dead, dead)


def random_units(n, frac, seed):
    rng = np.random.RandomState(seed)
    return sorted(rng.choice(n, size=max(1, int(n * frac)), replace=False).tolist())


def test_vgg16_graph_matches_nan_trick():
    model, shape = workloads.build("vgg16_cifar")
    trace = cascade.trace_chain(model, shape)
    assert trace is not None and len(trace.calls) == 51
    graph = get_vgg_pruning_graph(model)
    assert len(graph) == 15
    for i, (module, cascading) in enumerate(graph):
        units = random_units(module.weight.shape[0], 0.25, seed=i)
        assert_same(cascade.resolve(trace, model, module, units, cascading),
                    probe(model, module, units, cascading, shape))


def test_small_cnn_flatten_expands_channels_to_blocks():
    model, shape = workloads.build("small_cnn")
    trace = cascade.trace_chain(model, shape)
    for i, (module, cascading) in enumerate(model.get_pruning_graph()):
        units = random_units(module.weight.shape[0], 0.1, seed=10 + i)
        got = cascade.resolve(trace, model, module, units, cascading)
        assert_same(got, probe(model, module, units, cascading, shape))
    conv2, fc1 = model.features[3], model.classifier[0]
    length, dead = cascade.resolve(trace, model, conv2, [1, 5], [fc1])[id(fc1)]
    hw = 4096 // 64  # reference test_pruner.py:83-92: channel c -> [c*HW, (c+1)*HW)
    assert length == 4096.0
    np.testing.assert_array_equal(dead, np.r_[hw:2 * hw, 5 * hw:6 * hw])


def test_mlp_with_flatten_module_and_leaky_relu():
    model, shape = workloads.build("mlp")
    trace = cascade.trace_chain(model, shape)
    fc1, fc2, fc3 = model.fc[1], model.fc[3], model.fc[5]
    for module, cascading in ((fc1, [fc2]), (fc2, [fc3]), (fc1, [fc3, fc2])):
        units = random_units(module.weight.shape[0], 0.3, seed=3)
        assert_same(cascade.resolve(trace, model, module, units, cascading),
                    probe(model, module, units, cascading, shape))


def test_second_consumer_sees_everything_dead():
    """Behind the first Linear/conv that reads a dead input every unit is NaN in the probe."""
    model, shape = workloads.build("vgg16_cifar")
    trace = cascade.trace_chain(model, shape)
    convs = model.conv_layers()
    cascading = [convs[1], convs[2], model.features[4], model.features[8]]
    assert_same(cascade.resolve(trace, model, convs[0], [3], cascading),
                probe(model, convs[0], [3], cascading, shape))
    length, dead = cascade.resolve(trace, model, convs[0], [3], cascading)[id(convs[2])]
    assert dead.size == length == 64.0


def test_indices_negative_duplicate_unsorted_and_errors():
    model, shape = workloads.build("vgg16_cifar")
    trace = cascade.trace_chain(model, shape)
    conv, cascading = model.conv_layers()[4], [model.conv_layers()[5], model.features[15]]
    for units in ([5, 1, 5, -1], np.array([7, 3, 3], dtype=np.int64), torch.tensor([0, 255, -256])):
        listed = units.tolist() if not isinstance(units, list) else units
        assert_same(cascade.resolve(trace, model, conv, units, cascading),
                    probe(model, conv, listed, cascading, shape))
    with pytest.raises(IndexError):
        cascade.resolve(trace, model, conv, [256], cascading)
    with pytest.raises(RuntimeError, match="int64"):
        cascade.resolve(trace, model, conv, np.array([1, 2], dtype=np.int32), cascading)
    got = cascade.resolve(trace, model, conv, [2], [])
    assert got == {}


def test_module_before_the_pruned_one_is_reached_but_clean():
    model, shape = workloads.build("vgg16_cifar")
    trace = cascade.trace_chain(model, shape)
    convs = model.conv_layers()
    cascading = [convs[0], convs[3]]
    assert_same(cascade.resolve(trace, model, convs[2], [0, 9], cascading),
                probe(model, convs[2], [0, 9], cascading, shape))


class _Residual(nn.Module):
    def __init__(self):
        super().__init__()
        self.a = nn.Conv2d(3, 8, 3, padding=1)
        self.r = nn.ReLU()
        self.b = nn.Conv2d(8, 8, 3, padding=1)
        self.c = nn.Conv2d(8, 4, 1)

    def forward(self, x):
        h = self.r(self.a(x))
        return self.c(self.b(h) + h)


class _Hidden(nn.Module):
    """A functional op between two modules: scaling (new tensor) or an in-place update."""

    def __init__(self, inplace):
        super().__init__()
        self.a = nn.Conv2d(3, 8, 3, padding=1)
        self.b = nn.Conv2d(8, 4, 1)
        self.inplace = inplace

    def forward(self, x):
        h = self.a(x)
        h = h.mul_(2.0) if self.inplace else h * 2.0
        return self.b(h)


class _Twice(nn.Module):
    def __init__(self):
        super().__init__()
        self.a = nn.Conv2d(3, 8, 3, padding=1)
        self.r = nn.ReLU()
        self.b = nn.Conv2d(8, 8, 3, padding=1)

    def forward(self, x):
        return self.r(self.b(self.r(self.a(x))))


def test_unprovable_links_return_none():
    shape = (3, 8, 8)
    m = _Residual().eval()
    trace = cascade.trace_chain(m, shape)
    assert cascade.resolve(trace, m, m.a, [1], [m.b]) is not None       # a -> relu -> b is a chain
    for p in m.parameters():          # no autograd graph, no way to see the add behind m.b
        p.requires_grad_(False)
    blind = cascade.trace_chain(m, shape)
    assert cascade.resolve(blind, m, m.b, [1], [m.c]) is None
    assert cascade.resolve(blind, m, m.a, [1], [m.b]) is not None
    for p in m.parameters():
        p.requires_grad_(True)
    for inplace in (False, True):
        m = _Hidden(inplace).eval()
        assert cascade.resolve(cascade.trace_chain(m, shape), m, m.a, [1], [m.b]) is None
    m = _Twice().eval()
    assert cascade.resolve(cascade.trace_chain(m, shape), m, m.a, [1], [m.r, m.b]) is None
    grouped = nn.Sequential(nn.Conv2d(4, 8, 3, padding=1), nn.Conv2d(8, 8, 3, padding=1, groups=8),
                            nn.Conv2d(8, 2, 1)).eval()
    trace = cascade.trace_chain(grouped, (4, 8, 8))
    assert cascade.resolve(trace, grouped, grouped[0], [1], [grouped[1]]) is not None
    assert cascade.resolve(trace, grouped, grouped[0], [1], [grouped[1], grouped[2]]) is None
    soft = nn.Sequential(nn.Linear(6, 5), nn.Softmax(dim=1), nn.Linear(5, 2)).eval()
    assert cascade.resolve(cascade.trace_chain(soft, (6,)), soft, soft[0], [1], [soft[2]]) is None


def test_training_mode_with_side_effects_is_left_to_the_probe():
    model, shape = workloads.build("vgg16_cifar")
    trace = cascade.trace_chain(model, shape)
    conv, cascading = model.conv_layers()[0], [model.conv_layers()[1], model.features[1]]
    assert cascade.resolve(trace, model, conv, [1], cascading) is not None
    model.train()
    assert cascade.resolve(cascade.trace_chain(model, shape), model, conv, [1], cascading) is None
    assert cascade.structure_signature(model, shape) != trace.signature


def test_trace_leaves_the_model_untouched():
    model, shape = workloads.build("vgg16_cifar")
    before = {k: v.clone() for k, v in model.state_dict().items()}
    hooks = sum(len(m._forward_hooks) + len(m._forward_pre_hooks) for m in model.modules())
    state = torch.get_rng_state()
    np_state = np.random.get_state()[1].copy()
    cascade.trace_chain(model, shape)
    assert sum(len(m._forward_hooks) + len(m._forward_pre_hooks) for m in model.modules()) == hooks
    assert torch.equal(torch.get_rng_state(), state)
    assert np.array_equal(np.random.get_state()[1], np_state)
    for k, v in model.state_dict().items():
        assert torch.equal(v, before[k]), k


def _random_net(rng):
    """A random feature extractor + classifier in the reference's vocabulary."""
    layers, c, hw = [], int(rng.choice([1, 3])), int(rng.choice([8, 12]))
    in_shape = (c, hw, hw)
    for _ in range(int(rng.randint(1, 4))):
        out = int(rng.choice([4, 6, 9]))
        layers.append(nn.Conv2d(c, out, 3, padding=1))
        c = out
        if rng.rand() < 0.6:
            layers.append(nn.BatchNorm2d(c))
        layers.append([nn.ReLU(), nn.LeakyReLU(0.1), nn.ReLU(inplace=True), nn.Tanh()][int(rng.randint(4))])
        if rng.rand() < 0.5 and hw % 2 == 0:
            layers.append([nn.MaxPool2d(2), nn.AvgPool2d(2)][int(rng.randint(2))])
            hw //= 2
        if rng.rand() < 0.3:
            layers.append(nn.Dropout2d(0.2))
    if rng.rand() < 0.3:
        layers.append(nn.AdaptiveAvgPool2d(1))
        hw = 1
    layers.append(nn.Flatten())
    f = c * hw * hw
    for _ in range(int(rng.randint(1, 3))):
        out = int(rng.choice([5, 7]))
        layers += [nn.Linear(f, out), nn.ReLU()]
        if rng.rand() < 0.4:
            layers.append(nn.Dropout(0.3))
        f = out
    layers.append(nn.Linear(f, 3))
    return nn.Sequential(*layers).eval(), in_shape


@pytest.mark.parametrize("seed", range(25))
def test_random_sequential_models_resolve_like_the_nan_trick(seed):
    rng = np.random.RandomState(1000 + seed)
    torch.manual_seed(seed)
    model, shape = _random_net(rng)
    trace = cascade.trace_chain(model, shape)
    assert trace is not None
    prunable = [m for m in model if isinstance(m, (nn.Conv2d, nn.Linear))][:-1]
    listeners = [m for m in model if isinstance(m, (nn.Conv2d, nn.Linear, nn.BatchNorm2d,
                                                    nn.Dropout, nn.Dropout2d))]
    for module in prunable:
        n = module.weight.shape[0]
        units = sorted(rng.choice(n, size=int(rng.randint(1, n)), replace=False).tolist())
        chosen = [m for m in listeners if rng.rand() < 0.7]
        got = cascade.resolve(trace, model, module, units, chosen)
        assert_same(got, probe(model, module, units, chosen, shape, seed=seed))


def test_stale_trace_resolves_against_the_live_widths():
    """A trace recorded before a prune still describes the chain, but every width in it is the
    old one.  Pruning the same layer a second time must use the sizes the model has NOW: negative
    indices wrap modulo the current width, out-of-range ones are rejected like the reference's
    ``index_fill_`` does, Dropout's ``original_len`` is the current input width, and "everything
    behind a Linear is dead" spans the current width."""
    torch.manual_seed(0)
    model = nn.Sequential(nn.Linear(10, 100), nn.ReLU(), nn.Dropout(0.5), nn.Linear(100, 20),
                          nn.ReLU(), nn.Linear(20, 3)).eval()
    shape = (10,)
    trace = cascade.trace_chain(model, shape)
    fc1, drop, fc2, fc3 = model[0], model[2], model[3], model[5]
    np.random.seed(0)
    oracle.OraclePruner(model, shape).prune_model(fc1, list(range(10)), [drop, fc2])
    assert fc1.weight.shape[0] == 90 and drop.p == 0.5 * 0.9
    # second prune of the same layer, through the trace recorded at width 100
    for units in ([0, 5, -1, 88, 12, 40, 41, 42, 43], [89], [-90]):
        got = cascade.resolve(trace, model, fc1, units, [drop, fc2, fc3])
        assert_same(got, probe(model, fc1, units, [drop, fc2, fc3], shape))
        assert got[id(drop)][0] == 90.0
    for bad in ([90], [95], [-91]):   # valid for the recorded width, not for the model as it is
        with pytest.raises(IndexError):
            cascade.resolve(trace, model, fc1, bad, [drop, fc2])
    # the layer behind was narrowed as well: "everything dead" is the current 17 units, not 20
    np.random.seed(0)
    oracle.OraclePruner(model, shape).prune_model(fc2, [1, 2, 3], [fc3])
    got = cascade.resolve(trace, model, fc1, [4], [fc2, fc3])
    assert_same(got, probe(model, fc1, [4], [fc2, fc3], shape))
    assert got[id(fc3)][0] == 17.0 and got[id(fc3)][1].size == 17
    # a model whose widths no longer fit together is left to the probe forward (which raises)
    fc2.weight.data = fc2.weight.data[:, :50]
    assert cascade.resolve(trace, model, fc1, [4], [fc2]) is None


# ------------------------------------------------------------------ residual networks
def test_residual_add_is_resolved_from_the_autograd_graph():
    """`b(h) + h`: no module sees the add, the input of `c` is a new tensor.  Its grad_fn is an
    AddBackward0 whose operands are the outputs of `b` and of the ReLU, which the trace recorded,
    so a dead channel of either operand is dead in the sum."""
    shape = (3, 8, 8)
    m = _Residual().eval()
    trace = cascade.trace_chain(m, shape)
    for module, units, listeners in ((m.b, [1, 6], [m.c]), (m.a, [0, 3], [m.b, m.c]),
                                     (m.a, [7], [m.c])):
        assert_same(cascade.resolve(trace, m, module, units, listeners),
                    probe(m, module, units, listeners, shape))


def _small_resnet(block="bottleneck"):
    import torchvision
    torch.manual_seed(3)
    if block == "bottleneck":
        net = torchvision.models.resnet.ResNet(torchvision.models.resnet.Bottleneck, [2, 2, 1, 1],
                                               num_classes=7, width_per_group=16)
    else:
        net = torchvision.models.resnet.ResNet(torchvision.models.resnet.BasicBlock, [2, 1, 1, 1],
                                               num_classes=7)
    return net.eval()


@pytest.mark.parametrize("block", ["bottleneck", "basic"])
def test_resnet_graph_resolves_like_the_nan_trick(block):
    """torchvision-style ResNets (`out += identity` in place, one shared ReLU per block, a
    functional flatten before `fc`): every entry of get_resnet_pruning_graph, and prunes INTO the
    residual stream with listeners further down, against the probe forward."""
    from torchpruner_b200.utils import get_resnet_pruning_graph
    net = _small_resnet(block)
    shape = (3, 32, 32)
    trace = cascade.trace_chain(net, shape)
    assert trace is not None
    graph = get_resnet_pruning_graph(net)
    n_blocks = sum(1 for m in net.modules() if hasattr(m, "conv1") and m is not net)
    assert len(graph) == (2 * n_blocks + 1 if block == "bottleneck" else n_blocks)
    rng = np.random.RandomState(5)
    for module, listeners in graph:
        n = module.weight.shape[0]
        units = sorted(rng.choice(n, size=max(1, n // 4), replace=False).tolist())
        got = cascade.resolve(trace, net, module, units, listeners)
        want = probe(net, module, units, listeners, shape)
        assert_same(got, want)
        for consumer in listeners:
            if isinstance(consumer, nn.Conv2d):
                np.testing.assert_array_equal(want[id(consumer)][1], units)
    # into the stream: the block's last conv, heard in this block, the next one and the head
    blocks = [m for m in net.modules() if hasattr(m, "conv1") and m is not net]
    last_conv = blocks[0].conv3 if block == "bottleneck" else blocks[0].conv2
    last_bn = blocks[0].bn3 if block == "bottleneck" else blocks[0].bn2
    listeners = [last_bn, blocks[1].conv1, blocks[1].bn1, blocks[2].conv1, net.fc]
    listeners += [m for m in blocks[2].downsample.modules() if isinstance(m, nn.Conv2d)]
    got = cascade.resolve(trace, net, last_conv, [2, 5], listeners)
    assert_same(got, probe(net, last_conv, [2, 5], listeners, shape))
    assert got[id(blocks[1].conv1)][1].tolist() == [2, 5]        # through the add and the ReLU
    assert got[id(net.fc)][1].size == net.fc.weight.shape[1]     # behind a convolution: all dead
