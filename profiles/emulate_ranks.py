"""Run layer `L` of the 8-rank bench step on ONE GPU, one emulated rank after the other, to find
which rank's permutations make a layer slow (fp16 range fallback)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.argv = ["bench.py"]
import bench
from torchpruner_b200 import _lib, workloads
from torchpruner_b200.attributions import ShapleyAttributionMetric
from torchpruner_b200.attributions.methods import shapley_values as _sv
_sv.allreduce_scores = lambda sv, group=None: sv
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
model, shape = bench.build_model(dev)
hx, hy = bench.host_batch(shape, pin=False)
dx, dy = hx.to(dev), hy.to(dev)
WORLD = 8
SV = bench.PERMS_PER_STEP * WORLD
layers = [int(a) for a in sys.argv[1:]] or [0]
for L in layers:
    conv = model.conv_layers()[L]
    for seed in (1002, 1003):
        for r in range(WORLD):
            np.random.seed(seed)
            m = ShapleyAttributionMetric(model, bench.OneBatch(dx, dy), workloads.cross_entropy, dev, sv_samples=SV, reduction="none")
            m._dist = lambda r=r: (None, WORLD, r)
            ev = m.find_evaluation_module(conv, find_best_evaluation_module=True)
            with bench.quiet():
                torch.cuda.synchronize(); t0 = time.perf_counter()
                sv = m._scores_on_device(ev, SV)
                torch.cuda.synchronize(); t1 = time.perf_counter()
            info = m.last_run_info
            print(f"layer {L} seed {seed} rank {r}: {1e3*(t1-t0):8.1f} ms f16 {info['f16']} banned {getattr(m, '_f16_banned', False)} finite {bool(torch.isfinite(sv).all())} max|sv| {float(sv.abs().max()):.3e}", flush=True)
