"""cProfile of one Shapley step (sv_samples=3) on one GPU: where the HOST time of a layer run goes
(per-layer fixed costs are what limits strong scaling once the evaluations are spread thin)."""
import cProfile, pstats, os, sys, io
import numpy as np, torch
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
    pr = cProfile.Profile()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    import time
    t0 = time.perf_counter(); a.record()
    pr.enable(); step(2); pr.disable()
    b.record(); torch.cuda.synchronize(); wall = time.perf_counter() - t0
out = io.StringIO()
pstats.Stats(pr, stream=out).sort_stats("cumulative").print_stats(45)
print("wall %.1f ms, device %.1f ms" % (wall * 1e3, a.elapsed_time(b)))
print(out.getvalue()[:9000])
