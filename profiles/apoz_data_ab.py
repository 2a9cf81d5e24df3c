"""Does the one-launch config-2 sweep depend on the data?  APoZ and |x|-sum, each on post-ReLU
activations (half zeros) and on dense normal data (same shapes, same kernels)."""
import os, sys, statistics
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torchpruner_b200 import _lib
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
shapes = [(256, 64, 32, 32)] * 2 + [(256, 128, 16, 16)] * 2 + [(256, 256, 8, 8)] * 3 + \
    [(256, 512, 4, 4)] * 3 + [(256, 512, 2, 2)] * 3
dense = [torch.randn(s, device=dev) for s in shapes]
relu = [torch.relu(torch.randn(s, device=dev)) for s in shapes]
flush = torch.empty(128 << 20, dtype=torch.float32, device=dev).normal_()
nbytes = sum(4 * int(np.prod(s)) + 4 * s[0] * s[1] for s in shapes)


def timed(fn, iters=15):
    for _ in range(3):
        fn()
    ms = []
    for _ in range(iters):
        for _ in range(6):
            flush.sum()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize()
        ms.append(a.elapsed_time(b))
    return statistics.median(ms)


for op in ("apoz", "abs_sum"):
    for name, data in (("post-ReLU", relu), ("dense", dense)):
        ms = timed(lambda: _lib.reduce_rows_multi(op, data))
        print(f"{op:8s} on {name:9s}: {ms * 1e3:6.1f} us  {nbytes / ms / 1e6:6.0f} GB/s  {nbytes / ms / 1e6 / 6552:.3f} of the copy peak")
