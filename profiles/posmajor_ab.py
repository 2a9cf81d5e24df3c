"""Row-major against position-major tiles for the larger images (conv_pos_major_max_hw 64 vs 1024)
at Shapley stack sizes; same bits either way."""
import os, sys, statistics
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torchpruner_b200 import _lib
DEV = torch.device("cuda:0")


def timed(fn, iters=5):
    ms = []
    for i in range(iters + 1):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); out = fn(); b.record(); b.synchronize()
        if i:
            ms.append(a.elapsed_time(b))
    return statistics.median(ms), out


for (N, H, W, Cin, Cout) in [(64 * 256, 16, 16, 64, 128), (64 * 256, 16, 16, 128, 128),
                             (16 * 256, 32, 32, 64, 64), (256, 16, 16, 128, 128), (256, 32, 32, 64, 64)]:
    gen = torch.Generator(device=DEV).manual_seed(1)
    x = torch.randn((N, H, W, Cin), device=DEV, generator=gen)
    w = torch.randn((Cout, 9 * Cin), device=DEV, generator=gen) / (9 * Cin) ** 0.5
    scale = torch.rand(Cout, device=DEV, generator=gen) + 0.5
    shift = torch.randn(Cout, device=DEV, generator=gen) * 0.3
    s_x = 2.0 ** round(np.log2(64.0 / float(x.abs().max())))
    s_w = 2.0 ** round(np.log2(64.0 / float(w.abs().max())))
    x_hi, x_lo = _lib.f16_split(x, s_x)
    del x
    w_hi, w_lo = _lib.f16_split(w, s_w)
    sc = (scale / (s_x * s_w)).contiguous()
    fn = lambda: _lib.conv_tf32x3(x_hi, x_lo, w_hi, w_lo, 9, sc, shift, True, split=True, out_scale=4.0)
    _lib.set_tuning("conv_pos_major_max_hw", 64)
    t_row, want = timed(fn)
    _lib.set_tuning("conv_pos_major_max_hw", 1024)
    t_pos, got = timed(fn)
    print(f"N={N} {H}x{W} {Cin}->{Cout}: row-major {t_row:.3f} ms, position-major {t_pos:.3f} ms, same bits "
          f"{torch.equal(got[0], want[0]) and torch.equal(got[1], want[1])}", flush=True)
    del x_hi, x_lo, want, got
    torch.cuda.empty_cache()
