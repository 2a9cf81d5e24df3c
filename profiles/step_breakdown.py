"""Per-shape breakdown of the conv launches of one bench step (CUDA events per launch)."""
import os, sys, json
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.argv = ["bench.py"]
import bench
from torchpruner_b200 import _lib, workloads
from torchpruner_b200.attributions import ShapleyAttributionMetric
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
model, shape = bench.build_model(dev)
hx, hy = bench.host_batch(shape, pin=False)
dx, dy = hx.to(dev), hy.to(dev)
def step(s):
    np.random.seed(s)
    m = ShapleyAttributionMetric(model, bench.OneBatch(dx, dy), workloads.cross_entropy, dev, sv_samples=1, reduction="none")
    for conv in model.conv_layers():
        m._scores_on_device(m.find_evaluation_module(conv, find_best_evaluation_module=True), 1)
with bench.quiet():
    step(0); step(1)
    prof = _lib.profile_begin(timed=True); prof.by_shape = True
    step(2)
    _lib.profile_end()
summ = prof.summary()
tot = sum(v["ms"] for v in summ.values())
for k, v in sorted(summ.items(), key=lambda kv: -kv[1]["ms"]):
    tf = v["flops"] / v["ms"] / 1e9 if v["flops"] else 0
    print(f"{v['ms']:8.2f} ms {100*v['ms']/tot:5.1f}%  n={v['launches']:5d}  {3*tf:7.1f} exec TF/s  {k}")
print("total own-kernel ms", tot)
