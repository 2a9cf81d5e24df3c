"""conv4 of VGG16 (16x16, 128->128, MaxPool behind it) at its Shapley stack size: row-major tiles +
tp_maxpool2_split_nhwc_f16 against position-major tiles with the pool fused (conv_pos_major_max_hw
raised to 256).  Same bits either way."""
import os, sys, statistics
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torchpruner_b200 import _lib
DEV = torch.device("cuda:0")
N, H, W, Cin, Cout = 64 * 256, 16, 16, 128, 128
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


def timed(fn, iters=5):
    ms = []
    for i in range(iters + 1):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); out = fn(); b.record(); b.synchronize()
        if i:
            ms.append(a.elapsed_time(b))
    return statistics.median(ms), out


def two_kernels():
    raw, _ = _lib.conv_tf32x3(x_hi, x_lo, w_hi, w_lo, 9, sc, shift, True, split=False)
    return _lib.maxpool2_split_nhwc(raw, half=True, scale=4.0)


def fused():
    return _lib.conv_tf32x3(x_hi, x_lo, w_hi, w_lo, 9, sc, shift, True, split=True, out_scale=4.0,
                            pool=True)


def conv_only():
    return _lib.conv_tf32x3(x_hi, x_lo, w_hi, w_lo, 9, sc, shift, True, split=False)


_lib.set_tuning("conv_pos_major_max_hw", 64)     # the limit before this measurement
t2, want = timed(two_kernels)
tc, _ = timed(conv_only)
_lib.set_tuning("conv_pos_major_max_hw", 256)
tpm, _ = timed(conv_only)
tf, got = timed(fused)
print(f"row-major conv {tc:.3f} ms, conv + pool kernel {t2:.3f} ms; position-major conv {tpm:.3f} ms, "
      f"position-major with the pool fused {tf:.3f} ms; same bits: "
      f"{torch.equal(got[0], want[0]) and torch.equal(got[1], want[1])}")
