"""Where does the distance between the GPU Shapley tables and the float64 twin come from?
For VGG16 layers (sv_samples=2, 32 examples, the tests' golden case): the table's distance to the
reference's float64-twin table, normalised by the largest score, with the cached activation z
taken from (a) the model's own fp32 forward on the GPU (cuDNN), (b) the library's own head plan ("engine": fp32 products, float64 channel sums),
(c) a float64 forward rounded to fp32 — each for the engine tail and the forward_partial tail.

    python profiles/parity_probe.py [features.N ...]  >  profiles/r02_parity_sources.txt
"""
import contextlib
import copy
import io
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from torchpruner_b200 import workloads  # noqa: E402
from torchpruner_b200.attributions import ShapleyAttributionMetric  # noqa: E402

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
DEV = torch.device("cuda:0")
g = np.load(os.path.join(ROOT, "tests", "golden", "vgg16.npz"))
model, shape = workloads.build("vgg16_cifar")
model.to(DEV)
twin = copy.deepcopy(model).double()
names = {m: n for n, m in model.named_modules()}
small = workloads.synthetic_batches(shape, n=32, batch_size=32, seed=5)
tags = sys.argv[1:] or ["features.%d" % i for i in (0, 3, 7, 10, 14, 17, 20, 24, 27, 30, 34, 37, 40)]
print("# distance to the reference's float64-twin table / max|score|; columns: z source, tail")
print("# layer        max|score|  reference(fp32 CPU)  " + "  ".join(
    f"{z}/{t}" for z in ("cudnn", "engine", "f64") for t in ("engine", "fwd_partial")))
for tag in tags:
    layer = dict(model.named_modules())[tag]
    want, ref = g[f"{tag}/shapley32_f64"], g[f"{tag}/shapley32"]
    scale = np.abs(want).max()
    cells = []
    for z_src in ("cudnn", "engine", "f64"):
        for native in ("auto", False):
            if z_src == "engine" and native is False:
                cells.append("    -   ")
                continue
            np.random.seed(3)
            m = ShapleyAttributionMetric(model, small, workloads.cross_entropy, DEV, sv_samples=2,
                                         reduction="none")
            m.native_tail = native
            m.native_head = z_src == "engine"
            if z_src == "f64":
                orig = m.run_forward_partial

                def patched(xx, y_true=None, to_module=None, from_module=None, _orig=orig):
                    if from_module is None and to_module is not None:
                        with torch.no_grad():
                            z64 = twin.forward_partial(
                                xx.to(DEV).double(),
                                to_module=dict(twin.named_modules())[names[to_module]])
                        return z64.float(), None
                    return _orig(xx, y_true=y_true, to_module=to_module, from_module=from_module)
                m.run_forward_partial = patched
            with contextlib.redirect_stdout(io.StringIO()):
                sv = m.run(layer, find_best_evaluation_module=True)
            assert m.last_run_info["native_head"] == (z_src == "engine"), m.last_run_info
            cells.append("%.2e" % (np.abs(sv - want).max() / scale))
    print(f"{tag:12s}  {scale:9.3f}   {np.abs(ref - want).max() / scale:.2e}            "
          + "  ".join(cells), flush=True)


# Per-evaluation losses (the 1e-5 relative bar of north_star is meaningful on these): one
# permutation, 32 examples; L[k, b] rebuilt from the score table (sv_samples=1: the fp32 loss
# differences are exact in the float64 table, so L_k = L_0 + cumulative sum along the permutation)
# against the same ladder evaluated by the float64 twin.
print("# per-evaluation loss: max over (k, b) of |L - L_f64|, divided by the largest loss of the ladder")
print("# (losses of single examples range from 1e-3 to 10: an element-wise relative error would")
print("#  measure how small the smallest loss is)")
for tag in tags:
    layer = dict(model.named_modules())[tag]
    np.random.seed(11)
    m = ShapleyAttributionMetric(model, small, workloads.cross_entropy, DEV, sv_samples=1,
                                 reduction="none")
    with contextlib.redirect_stdout(io.StringIO()):
        sv = m.run(layer, find_best_evaluation_module=True)
        module = m.find_evaluation_module(layer, find_best_evaluation_module=True)
    np.random.seed(11)
    perm = np.random.permutation(layer.out_channels)
    x, y = next(iter(small))
    x, y = x.to(DEV), y.to(DEV)
    tm = dict(twin.named_modules())[names[module]]
    worst = biggest = 0.0
    with torch.no_grad():
        z = twin.forward_partial(x.double(), to_module=tm)
        l0 = workloads.cross_entropy(twin.forward_partial(z, from_module=tm), y, reduction="none")
        ladder = l0.cpu().numpy() + np.cumsum(sv[:, perm], axis=1).T      # [C, 32], engine
        for k, c in enumerate(perm):
            z[:, c] = 0.0
            lk = workloads.cross_entropy(twin.forward_partial(z, from_module=tm), y,
                                         reduction="none").cpu().numpy()
            worst = max(worst, float(np.max(np.abs(ladder[k] - lk))))
            biggest = max(biggest, float(np.max(np.abs(lk))))
    print(f"{tag:12s}  {worst / biggest:.2e}   (max |L| {biggest:.2f}, max |dL| {worst:.2e})", flush=True)
