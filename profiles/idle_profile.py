"""Where the GPU idles in a thinly loaded Shapley step (sv_samples=3 on one GPU ~ one rank's share
of the 25-permutation job on 8 GPUs).  Busy time is the union of the device intervals of every
kernel and copy in torch.profiler's trace; the gaps are listed with the kernel that ends each.

    python profiles/idle_profile.py [sv_samples]"""
import os, sys
import numpy as np, torch
from torch.profiler import profile, ProfilerActivity
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
SV = int(sys.argv[1]) if len(sys.argv) > 1 else 3
sys.argv = ["bench.py"]
import bench
from torchpruner_b200 import workloads
from torchpruner_b200.attributions import ShapleyAttributionMetric, native_tail
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
model, shape = bench.build_model(dev)
hx, hy = bench.host_batch(shape, pin=False)
dx, dy = hx.to(dev), hy.to(dev)
convs = list(model.conv_layers())


def step(s):
    native_tail._ACTS.clear()       # as bench.py: no activation survives from the previous step
    np.random.seed(s)
    m = ShapleyAttributionMetric(model, bench.OneBatch(dx, dy), workloads.cross_entropy, dev,
                                 sv_samples=SV, reduction="none")
    for conv in convs:
        m.run(conv, find_best_evaluation_module=True)


with bench.quiet():
    step(0); step(1)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); step(2); b.record(); torch.cuda.synchronize()
    plain_ms = a.elapsed_time(b)
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        step(3)
        torch.cuda.synchronize()

ev = [e for e in prof.events() if getattr(e, "device_type", None) is not None
      and str(e.device_type).endswith("CUDA") and e.time_range is not None]
iv = sorted((e.time_range.start, e.time_range.end, e.name) for e in ev)
t0, t1 = iv[0][0], max(e for _, e, _ in iv)
busy, own, cur_end, gaps = 0.0, 0.0, iv[0][0], []
for s, e, name in iv:
    if "tp::" in name:
        own += e - s
    if s > cur_end:
        gaps.append((s - cur_end, name))
        busy += e - s
        cur_end = e
    elif e > cur_end:
        busy += e - cur_end
        cur_end = e
span = t1 - t0
print(f"sv_samples={SV}: step without profiler {plain_ms:.1f} ms; under the profiler span {span/1e3:.1f} ms, "
      f"busy {busy/1e3:.1f} ms (own kernels {own/1e3:.1f}), idle {(span-busy)/1e3:.1f} ms in {len(gaps)} gaps")
by = {}
for g, name in gaps:
    k = name[:70]
    by.setdefault(k, [0.0, 0])
    by[k][0] += g; by[k][1] += 1
print("idle time by the kernel that ends the gap:")
for k, (g, n) in sorted(by.items(), key=lambda kv: -kv[1][0])[:25]:
    print(f"  {g/1e3:8.2f} ms  n={n:5d}  {k}")
print("largest single gaps:")
for g, name in sorted(gaps, reverse=True)[:15]:
    print(f"  {g/1e3:8.3f} ms  before {name[:80]}")
