"""What fills the time between the library's own kernels in a thinly loaded Shapley step
(sv_samples=3 on one GPU ~ one rank's share of the 25-permutation job on 8): torch.profiler's
CUDA timeline, own kernels (tp::*) against everything else (torch elementwise ops of the head,
plan checksums, criterion, copies) and idle time."""
import os, sys
import numpy as np, torch
from torch.profiler import profile, ProfilerActivity
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.argv = ["bench.py"]
import bench
from torchpruner_b200 import workloads
from torchpruner_b200.attributions import ShapleyAttributionMetric
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
model, shape = bench.build_model(dev)
hx, hy = bench.host_batch(shape, pin=False)
dx, dy = hx.to(dev), hy.to(dev)
def step(s, sv=3):
    np.random.seed(s)
    m = ShapleyAttributionMetric(model, bench.OneBatch(dx, dy), workloads.cross_entropy, dev, sv_samples=sv, reduction="none")
    for conv in model.conv_layers():
        m._scores_on_device(m.find_evaluation_module(conv, find_best_evaluation_module=True), sv)
with bench.quiet():
    step(0); step(1)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); step(2); b.record(); torch.cuda.synchronize()
    plain_ms = a.elapsed_time(b)
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        step(3)
        torch.cuda.synchronize()
own = other = 0.0
rows = []
for e in prof.key_averages():
    t = getattr(e, "self_device_time_total", 0) or getattr(e, "self_cuda_time_total", 0)
    if t <= 0:
        continue
    if "tp::" in e.key:
        own += t
    else:
        other += t
    rows.append((t, e.count, e.key[:90]))
print(f"step without profiler: {plain_ms:.1f} ms; own kernels {own/1e3:.1f} ms, other GPU work {other/1e3:.1f} ms, "
      f"idle ~ {plain_ms - (own + other)/1e3:.1f} ms")
for t, n, k in sorted(rows, reverse=True)[:40]:
    print(f"{t/1e3:9.2f} ms  n={n:6d}  {k}")
