"""bench.py's reduction timings (config-2 sweep, config-4 shapes) with the library named by TP_LIB
(A/B of two builds on the same box: run once per build)."""
import os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torchpruner_b200 import _lib
if os.environ.get("TP_LIB"):
    _lib.LIB_PATH = os.path.abspath(os.environ["TP_LIB"])
sys.argv = ["bench.py"]
import bench
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
r = bench.time_reductions(dev, 6552.0)
for k in ("taylor", "apoz", "sensitivity"):
    v = r[k]
    print(os.environ.get("TP_LIB", "default"), k,
          {q: round(v[q], 3) for q in ("c2_sweep_frac", "c2_sweep_13_launches_frac", "c4_sweep_frac",
                                       "c4_sweep_run_many_frac", "c4_largest_frac")})
