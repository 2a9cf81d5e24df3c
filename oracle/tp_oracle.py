"""CPU oracle for the attribution-scoring / slicing hot path.

TEST INFRASTRUCTURE ONLY.  This module restates, on the CPU, the algorithm of the reference
(TorchPruner, pure Python over torch + numpy) for every operation the CUDA library replaces.
It may be imported by tests/, by __graft_entry__.smoke() and by bench.py's CPU-baseline /
`--impl reference` legs — as the checker or the timed baseline — and by nothing else.  The
product (torchpruner_b200/) never imports it and has no CPU path.

Parity status: PINNED.  The restatement is checked
  * against the reference's own known-answer tests on the 4-unit "max model"
    (torchpruner/tests/test_attributions.py:84-175) and its NaN-trick / shape tests
    (torchpruner/tests/test_pruner.py:72-176), restated in tests/test_oracle_golden.py, and
  * against outputs of the reference itself, produced in the build container by importing
    /root/reference (tests/golden/make_golden.py, vectors committed under tests/golden/).

Floating-point operations use the same torch CPU / numpy calls as the reference so the oracle
is bit-identical to it on the same machine; `*_f64` variants are the float64 arbiters used to
put a scale on fp32 summation-order differences.  Integer / index / byte results (APoZ counts,
keep indices, NaN indices, sliced tensors) are exact by construction.

Citations are file:line under the reference tree (torchpruner/...).
"""
import copy

import numpy as np
import torch

# --------------------------------------------------------------------------- K1-K4


def taylor_rows(z, g, signed=False):
    """taylor.py:40-44: -1 * (grad * activation), flatten(2).sum(-1) if >2-D, abs unless signed."""
    z, g = torch.as_tensor(z), torch.as_tensor(g)
    t = -1.0 * (g * z)
    if t.dim() > 2:
        t = t.flatten(2).sum(-1)
    if not signed:
        t = t.abs()
    return t.numpy()


def taylor_rows_f64(z, g, signed=False):
    z = np.asarray(z, dtype=np.float32)
    g = np.asarray(g, dtype=np.float32)
    prod = (g * z).astype(np.float64)  # the fp32-rounded product, summed exactly enough
    t = -prod.reshape(prod.shape[0], prod.shape[1], -1).sum(-1)
    return t if signed else np.abs(t)


def apoz_rows(z):
    """apoz.py:31-33: (output > 0).float(), sum(-1) until 2-D.  Integer valued."""
    z = torch.as_tensor(z)
    c = (z > 0).float()
    while c.dim() > 2:
        c = c.sum(-1)
    return c.numpy()


def apoz_rows_int(z):
    z = np.asarray(z)
    return (z > 0).reshape(z.shape[0], z.shape[1], -1).sum(-1).astype(np.int64)


def abs_sum_rows(g):
    """sensitivity.py:27-29: abs, flatten(2).sum(-1) if >2-D."""
    g = torch.as_tensor(g).abs()
    if g.dim() > 2:
        g = g.flatten(2).sum(-1)
    return g.numpy()


def abs_sum_rows_f64(g):
    g = np.abs(np.asarray(g, dtype=np.float64))
    return g.reshape(g.shape[0], g.shape[1], -1).sum(-1) if g.ndim > 2 else g


def weight_norm(w):
    """weight_norm.py:15-18: numpy abs, then sum(-1) until 1-D."""
    a = np.abs(np.asarray(w))
    while a.ndim > 1:
        a = a.sum(-1)
    return a


def weight_norm_f64(w):
    a = np.abs(np.asarray(w, dtype=np.float64))
    return a.reshape(a.shape[0], -1).sum(-1)


def hook_scores(model, batches, criterion, module, kind, signed=False):
    """Whole-metric restatement (taylor.py:18-49, apoz.py:15-39, sensitivity.py:13-34): hooks on
    `module`, one forward(+backward) per batch, per-batch host copy + concatenate.  Returns the
    [N, C] table before aggregation.  Used as the timed CPU baseline of config 2."""
    rows = []
    state = {}

    def fwd(mod, _inp, out):
        if kind == "apoz":
            rows.append(apoz_rows(out.detach()))
            return
        if kind == "taylor":
            state["z"] = out.detach().clone()

        def bwd(g):
            if kind == "taylor":
                rows.append(taylor_rows(state["z"], g.detach(), signed))
            else:
                rows.append(abs_sum_rows(g.detach()))
        out.register_hook(bwd)

    h = module.register_forward_hook(fwd)
    try:
        if kind == "apoz":
            with torch.no_grad():
                for x, y in batches:
                    criterion(model(x), y, reduction="none")
        else:
            for x, y in batches:
                criterion(model(x), y).backward()
    finally:
        h.remove()
    return np.concatenate(rows, 0)


# --------------------------------------------------------------------------- K5 / K6


def prefix_mask_stack(z, perm, k0, P):
    """shapley_values.py:53-56: clone, then cumulative index_fill_(1, perm[k], 0.0).
    Returns the P masked copies for prefixes k0 .. k0+P-1 stacked along a new leading dim."""
    z = torch.as_tensor(z).clone()
    out = []
    for k in range(k0 + P):
        z.index_fill_(1, torch.tensor([int(perm[k])]).long(), 0.0)
        if k >= k0:
            out.append(z.clone())
    if not out:
        return np.zeros((0,) + tuple(z.shape), np.float32)
    return torch.stack(out).numpy()


def cross_entropy_rows(logits, target):
    """criterion(y, y_true, reduction="none") for the reference's cross entropy
    (experiments/models/cifar10.py:90-91) on the CPU."""
    import torch.nn.functional as F

    return F.cross_entropy(torch.as_tensor(logits), torch.as_tensor(target), reduction="none").numpy()


def loss_delta_scatter(losses, prev, perm, k0, sv_samples, sv, row0, scatter_from=0):
    """shapley_values.py:58-61 for P consecutive prefixes: delta = new_loss - loss (fp32),
    sv[rows, perm[k]] += (delta / sv_samples) in float64, loss = new_loss.  `losses` is [P, B]
    fp32.  Returns the updated `prev` (the last row)."""
    losses = np.asarray(losses, dtype=np.float32)
    prev = np.asarray(prev, dtype=np.float32).copy()
    P, B = losses.shape
    for p in range(P):
        delta = (losses[p] - prev).astype(np.float32)
        if p >= scatter_from:
            sv[row0:row0 + B, int(perm[k0 + p])] += (delta / np.float32(sv_samples)).astype(np.float32)
        prev = losses[p]
    return prev


def shapley_with_partial(model, batches, criterion, sv_samples, d, module, permutations=None,
                         record_losses=None):
    """shapley_values.py:28-64, restated: for every batch, permutation and prefix position zero
    one more channel, run the layers after `module`, add the per-example loss increase to
    sv[row, channel].  Permutations come from the global numpy RNG on the first batch unless
    given.  `record_losses`, if a list, receives (batch, j, k, loss[B]) tuples."""
    sv = None
    row = 0
    with torch.no_grad():
        for bi, (x, y) in enumerate(batches):
            z0 = model.forward_partial(x, to_module=module)
            loss0 = criterion(model.forward_partial(z0, from_module=module), y, reduction="none")
            n = z0.shape[1]
            if permutations is None:
                permutations = [np.random.permutation(n) for _ in range(sv_samples)]
            if sv is None:
                sv = np.zeros((d, n))
            B = z0.shape[0]
            for j in range(sv_samples):
                loss = loss0.detach().clone()
                z = z0.clone().detach()
                for k, i in enumerate(permutations[j]):
                    z.index_fill_(1, torch.tensor(np.array([i])).long(), 0.0)
                    new_loss = criterion(model.forward_partial(z, from_module=module), y,
                                         reduction="none")
                    delta = new_loss - loss
                    sv[row:row + B, i] += (delta / sv_samples).squeeze().detach().numpy()
                    if record_losses is not None:
                        record_losses.append((bi, j, k, new_loss.reshape(-1).numpy().copy()))
                    loss = new_loss
            row += B
    return sv


def shapley_eval_losses(model, x, y, criterion, module, perm, k_positions):
    """Per-example losses L[k, b] with channels perm[0..k] of `module`'s output zeroed, for the
    listed prefix positions only (bounded sample of shapley_values.py:55-57)."""
    out = {}
    with torch.no_grad():
        z0 = model.forward_partial(x, to_module=module)
        for k in k_positions:
            z = z0.clone()
            z.index_fill_(1, torch.as_tensor(np.asarray(perm[:k + 1])).long(), 0.0)
            out[k] = criterion(model.forward_partial(z, from_module=module), y,
                               reduction="none").reshape(-1).numpy()
    return out


def shapley_prefix_steps(model, z0, loss0, y, criterion, module, perm, n_steps, sv, sv_samples,
                          row0=0):
    """The body of the reference's hot loop (shapley_values.py:52-61) for the first `n_steps`
    prefix positions of one permutation: clone, then per step index_fill_ one more channel, run
    the layers after `module`, per-example loss, delta, host-side `sv +=`.  This is what the CPU
    baseline times (the cached activation z0 and loss0 are computed once per batch outside)."""
    with torch.no_grad():
        loss = loss0.detach().clone()
        z = z0.clone().detach()
        B = z.shape[0]
        for i in perm[:n_steps]:
            z.index_fill_(1, torch.tensor(np.array([i])).long(), 0.0)
            new_loss = criterion(model.forward_partial(z, from_module=module), y, reduction="none")
            delta = new_loss - loss
            sv[row0:row0 + B, i] += (delta / sv_samples).squeeze().detach().cpu().numpy()
            loss = new_loss
    return sv


def shapley_without_partial(model, batches, criterion, samples, module):
    """shapley_values.py:66-99, restated (whole-model forward per evaluation, masking hook)."""
    mask = []

    def hook(mod, _, output):
        mod._oracle_prune_dim = output.shape[1]
        return output.index_fill_(1, torch.tensor(mask).long(), 0.0)

    def all_losses():
        with torch.no_grad():
            return torch.cat([criterion(model(x), y, reduction="none") for x, y in batches], 0)

    handle = module.register_forward_hook(hook)
    try:
        original = all_losses()
        n = module._oracle_prune_dim
        sv = np.zeros((original.shape[0], n))
        for _ in range(samples):
            mask.clear()
            loss = original.detach().clone()
            for i in np.random.permutation(n):
                mask.append(int(i))
                new_loss = all_losses()
                sv[:, i] += ((new_loss - loss) / samples).squeeze().numpy()
                loss = new_loss
    finally:
        handle.remove()
        if hasattr(module, "_oracle_prune_dim"):
            delattr(module, "_oracle_prune_dim")
    return sv


def ablation_curve(model, x, y, criterion, module_nl, ranking):
    """Experiments notebook "CIFAR-10 - VGG16 - Layerwise robustness", cell 8 (:1272-1277),
    restated: activation at `module_nl`, then for each unit of `ranking` zero it (cumulatively),
    run the remaining layers, record the mean loss and the top-1 accuracy."""
    losses, accs = [], []
    with torch.no_grad():
        z = model.forward_partial(x, to_module=module_nl)
        for i in ranking:
            z.index_fill_(1, torch.tensor([int(i)]), 0.)
            y_pred = model.forward_partial(z, from_module=module_nl)
            new_loss = criterion(y_pred, y)
            current_pred = y_pred.argmax(dim=1, keepdim=True)
            new_acc = current_pred.eq(y.view_as(current_pred)).sum().float() / len(y)
            losses.append(new_loss.item())
            accs.append(new_acc.item())
    return np.asarray(losses, np.float32), np.asarray(accs, np.float32)


# --------------------------------------------------------------------------- K7 / K8


def keep_indices(n, indices):
    """pruner.py:105-108: mask = ones(n); mask[indices] = False; arange(n)[mask]."""
    mask = np.ones(n, dtype=bool)
    mask[indices] = False
    return np.arange(n)[mask]


def index_select(t, axis, keep):
    """pruner.py:109: param.data.index_select(axis, keep) — a contiguous bit-exact copy."""
    return np.ascontiguousarray(np.take(np.asarray(t), np.asarray(keep, dtype=np.int64), axis=axis))


def nan_fill(x, indices):
    """pruner.py:137-142."""
    x = torch.as_tensor(x).clone()
    x.index_fill_(1, torch.as_tensor(np.asarray(indices)).long(), float("nan"))
    return x.numpy()


def nan_indices(x):
    """pruner.py:153-163: sum(-1) until 2-D, sum(0), isnan, nonzero (ascending)."""
    t = torch.as_tensor(x)
    while t.dim() > 2:
        t = t.sum(-1)
    t = t.sum(0).flatten(0)
    return torch.isnan(t).nonzero().flatten(0).numpy()


class OraclePruner:
    """pruner.py:14-185 + opt_pruner.py restated on CPU tensors (index_select slicing, NaN-trick
    cascade discovery, Dropout rescale, SGD momentum slicing through optimizer.state — the
    reference's own lookup raises KeyError on torch >= 1.6, see tests/golden/make_golden.py)."""

    def __init__(self, model, input_size, optimizer=None):
        self.model, self.input_size, self.optimizer = model, input_size, optimizer

    def prune_model(self, module, indices, cascading_modules=None):
        from torch.nn.modules.batchnorm import _BatchNorm
        from torch.nn.modules.conv import _ConvNd
        from torch.nn.modules.dropout import _DropoutNd
        import torch.nn as nn

        cascading_modules = cascading_modules or []
        found = {}

        def nanify(_, __, output):
            return output.index_fill_(1, torch.tensor(indices), torch.tensor(np.nan))

        def detect(mod, inp, __):
            found[mod] = (float(inp[0].shape[1]), nan_indices(inp[0].detach()))

        handles = [module.register_forward_hook(nanify)]
        handles += [m.register_forward_hook(detect) for m in cascading_modules]
        x = torch.tensor(np.random.random((2,) + tuple(self.input_size))).float()
        self.model(x)
        for h in handles:
            h.remove()
        for m in cascading_modules:
            length, idx = found[m]
            if len(idx) == 0:
                continue
            if isinstance(m, (nn.Linear, _ConvNd)):
                self.prune_parameter(m, "weight", idx, 1)
            elif isinstance(m, _BatchNorm):
                for name in ("weight", "bias", "running_mean", "running_var"):
                    self.prune_parameter(m, name, idx, 0)
            elif isinstance(m, _DropoutNd):
                m.p *= (1.0 - len(idx) / length)
        self.prune_parameter(module, "weight", indices, 0)
        self.prune_parameter(module, "bias", indices, 0)

    def prune_parameter(self, module, name, indices, axis):
        param = getattr(module, name)
        if param is None:
            return
        keep = torch.tensor(keep_indices(param.data.shape[axis], indices))
        param.data = param.data.index_select(axis, keep)
        if param.grad is not None:
            param.grad.data = param.grad.data.index_select(axis, keep)
        if isinstance(self.optimizer, torch.optim.SGD):
            for p in self.optimizer.param_groups[0]["params"]:
                if p is param:
                    buf = self.optimizer.state.get(p, {}).get("momentum_buffer", None)
                    if buf is not None:
                        buf.data = buf.data.index_select(axis, keep)


def clone_model(model):
    return copy.deepcopy(model)
