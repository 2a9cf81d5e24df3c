"""Per scored layer: device time of the whole Shapley run (CUDA events around it) against the sum
of the library's own kernel durations inside it — the difference is torch ops (the forward to
the evaluation module, argmax, copies) plus GPU idle time while the host prepares launches."""
import os, sys, time
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
SV = bench.PERMS_PER_STEP
def run_layer(conv, s, timed):
    np.random.seed(s)
    m = ShapleyAttributionMetric(model, bench.OneBatch(dx, dy), workloads.cross_entropy, dev, sv_samples=SV, reduction="none")
    ev = m.find_evaluation_module(conv, find_best_evaluation_module=True)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter(); a.record()
    m._scores_on_device(ev, SV)
    t1 = time.perf_counter(); b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b), (t1 - t0) * 1e3
with bench.quiet():
    for s in range(2):
        for conv in model.conv_layers(): run_layer(conv, s, False)
    rows = []
    for i, conv in enumerate(model.conv_layers()):
        dev_ms, host_ms = run_layer(conv, 7, True)           # untimed kernels: true device time
        prof = _lib.profile_begin(timed=True)
        run_layer(conv, 7, True)
        _lib.profile_end()
        summ = prof.summary()
        own = sum(v["ms"] for v in summ.values()); n = sum(v["launches"] for v in summ.values())
        rows.append((i, conv.out_channels, dev_ms, host_ms, own, n))
print("layer  C   device_ms  host_enqueue_ms  own_kernel_ms  launches  gap_ms")
for i, c, d, h, o, n in rows:
    print(f"{i:3d} {c:4d} {d:10.2f} {h:12.2f} {o:12.2f} {n:8d} {d-o:8.2f}")
print("total", sum(r[2] for r in rows), sum(r[3] for r in rows), sum(r[4] for r in rows), sum(r[5] for r in rows))
