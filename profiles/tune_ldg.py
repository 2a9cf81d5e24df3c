import os, sys, statistics, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torchpruner_b200 import _lib
DEV = torch.device("cuda:0")
flush = torch.empty(128 << 20, dtype=torch.float32, device=DEV).normal_()
stream = torch.cuda.current_stream(DEV)
def timed(fn, iters=9):
    for _ in range(2): fn()
    ms = []
    for _ in range(iters):
        for _ in range(4): flush.sum()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream); fn(); b.record(stream); b.synchronize(); ms.append(a.elapsed_time(b))
    return statistics.median(ms)
_lib.set_tuning("reduce_variant", 1)
for shp in [(256,64,32,32),(256,128,16,16),(256,256,8,8),(256,512,4,4),(256,512,2,2),(256,2048,7,7),(256,1024,14,14)]:
    z = torch.relu(torch.randn(shp, device=DEV)); g = torch.randn(shp, device=DEV)
    n, r = z.numel(), shp[0]*shp[1]
    line = ["x".join(map(str, shp))]
    for waves in (1, 2, 4, 16):
        _lib.set_tuning("reduce_ldg_waves", waves)
        t = timed(lambda: _lib.taylor_reduce(z, g)); a = timed(lambda: _lib.apoz_count(z))
        line.append(f"w{waves}: T {(8*n+4*r)/t/1e6:7.0f} A {(4*n+4*r)/a/1e6:7.0f}")
    print("  ".join(line))
