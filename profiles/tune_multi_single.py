"""One tensor at a time: the single-tensor entry points (auto policy) against tp_reduce_rows_multi
with a one-entry table, on config 2 / config 4 shapes (CUDA events, L2 refilled between)."""
import os, statistics, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torchpruner_b200 import _lib
DEV = torch.device("cuda:0")
flush = torch.empty(128 << 20, dtype=torch.float32, device=DEV).normal_()
stream = torch.cuda.current_stream(DEV)
def timed(fn, iters=7):
    for _ in range(2): fn()
    ms = []
    for _ in range(iters):
        for _ in range(4): flush.sum()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream); fn(); b.record(stream); b.synchronize()
        ms.append(a.elapsed_time(b))
    return statistics.median(ms)
shapes = [(256, 64, 32, 32), (256, 128, 16, 16), (256, 256, 8, 8), (256, 512, 4, 4), (256, 512, 2, 2),
          (256, 256, 14, 14), (256, 1024, 14, 14), (256, 128, 28, 28), (256, 512, 28, 28), (256, 64, 56, 56)]
print("shape | op | single us GB/s | multi(1) us GB/s")
for s in shapes:
    z = torch.relu(torch.randn(s, device=DEV)); g = torch.randn(s, device=DEV)
    n, rows = z.numel(), s[0] * s[1]
    for op, f1, f2, nb in (("taylor", lambda: _lib.taylor_reduce(z, g), lambda: _lib.reduce_rows_multi("taylor", [(z, g)]), 8 * n + 4 * rows),
                           ("apoz", lambda: _lib.apoz_count(z), lambda: _lib.reduce_rows_multi("apoz", [z]), 4 * n + 4 * rows)):
        t1, t2 = timed(f1), timed(f2)
        print(f"{'x'.join(map(str, s)):18s} {op:7s} {t1*1e3:7.1f} {nb/t1/1e6:7.0f} | {t2*1e3:7.1f} {nb/t2/1e6:7.0f}")
    del z, g
