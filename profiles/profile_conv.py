"""Per-layer throughput of tp_conv_f16x3 / tp_conv_tf32x3 on the VGG16-CIFAR tail shapes at
Shapley batch sizes.

    python profiles/profile_conv.py [ncu] [tf32]

(fp16x3 flavour unless `tf32` is given.)

Without arguments: CUDA-event timing of every layer shape for both kernel variants and several
partial-sum intervals.  With `ncu`: launches two representative shapes a few times (to be wrapped
in `ncu --set full -k regex:conv_tf32x3`)."""
import os
import statistics
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torchpruner_b200 import _lib  # noqa: E402

if os.environ.get("TP_LIB"):   # an alternative build of the library (A/B experiments)
    _lib.LIB_PATH = os.path.abspath(os.environ["TP_LIB"])

DEV = torch.device("cuda:0")
# (H, W, Cin, Cout, taps): the 12 conv layers after features.0 and the three linear layers
LAYERS = [(32, 32, 64, 64, 9), (16, 16, 64, 128, 9), (16, 16, 128, 128, 9), (8, 8, 128, 256, 9),
          (8, 8, 256, 256, 9), (4, 4, 256, 512, 9), (4, 4, 512, 512, 9), (2, 2, 512, 512, 9),
          (1, 1, 512, 512, 1), (1, 1, 512, 10, 1)]


def images_for(H, W, C):
    """Stack size the Shapley driver would use: <= 512 MiB of hi+lo input, <= 512 masks x 256."""
    per_image = H * W * C * 8
    return int(min(512 * 256, max(256, (512 << 20) // per_image // 256 * 256)))


F16 = "tf32" not in sys.argv


def images_for_f16(H, W, C):
    per_image = H * W * C * 4
    return int(min(512 * 256, max(256, (512 << 20) // per_image // 256 * 256)))


def run(H, W, Cin, Cout, taps, N, iters=5):
    x = torch.randn((N, H, W, Cin), device=DEV)
    w = torch.randn((Cout, taps * Cin), device=DEV) / (taps * Cin) ** 0.5
    split = Cout != 10
    if F16:
        x_hi, x_lo = _lib.f16_split(x, 8.0)
        w_hi, w_lo = _lib.f16_split(w, 64.0)
        scale = torch.full((Cout,), 1.0 / 512.0, device=DEV)
        odt = torch.float16 if split else torch.float32
    else:
        x_hi, x_lo = _lib.tf32_split(x)
        w_hi, w_lo = _lib.tf32_split(w)
        scale = torch.ones(Cout, device=DEV)
        odt = torch.float32
    del x, w
    shift = torch.zeros(Cout, device=DEV)
    out = (torch.empty((N, H, W, Cout), device=DEV, dtype=odt),
           torch.empty((N, H, W, Cout), device=DEV, dtype=odt) if split else None)
    stream = torch.cuda.current_stream(DEV)
    ms = []
    for i in range(iters + 2):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        _lib.conv_tf32x3(x_hi, x_lo, w_hi, w_lo, taps, scale, shift, True, split=split, out=out,
                         out_scale=8.0 if F16 else 1.0)
        b.record(stream)
        b.synchronize()
        if i >= 2:
            ms.append(a.elapsed_time(b))
    t = statistics.median(ms)
    flops = 2.0 * N * H * W * Cout * taps * Cin
    return t, flops / t / 1e9


def main():
    if "ncu" in sys.argv:
        for shp in (LAYERS[0], LAYERS[6]):
            run(*shp, (images_for_f16 if F16 else images_for)(*shp[:3]), iters=1)
        return
    if "quick" in sys.argv:   # default tuning, every layer once (A/B of two builds: TP_LIB=path)
        tot_ms = tot_fl = 0.0
        for (H, W, Cin, Cout, taps) in LAYERS:
            N = (images_for_f16 if F16 else images_for)(H, W, Cin)
            ms, _ = run(H, W, Cin, Cout, taps, N, iters=7)
            fl = 3.0 * _lib.load().tp_conv_tf32x3_flops(N, H, W, Cin, Cout, taps)
            tot_ms += ms
            tot_fl += fl
            print(f"{H:2d} {W:2d} {Cin:4d} {Cout:4d} {taps:4d} {N:6d} {ms:7.3f} {fl / ms / 1e9:10.1f}")
        print(f"# {os.environ.get('TP_LIB', 'default build')}: weighted {tot_fl / tot_ms / 1e9:.1f} executed TF/s")
        return
    if "ncu_pair" in sys.argv:   # two shapes, single-CTA tiles then CTA pairs (wrap in ncu)
        for shp in ((4, 4, 256, 512, 9), (4, 4, 512, 512, 9)):
            for knob in (1, 2):
                _lib.set_tuning("conv_pair", knob)
                run(*shp, images_for_f16(*shp[:3]), iters=1)
        _lib.set_tuning("conv_pair", 0)
        return
    if "pair" in sys.argv:   # A/B: single-CTA tiles (conv_pair=1) against CTA pairs (0)
        print("pair  H  W  Cin Cout taps      N     ms  executed_TF/s (3 MMAs per product, skipped taps not counted)")
        for single in (1, 0):
            _lib.set_tuning("conv_pair", 1 if single else 2)
            tot_ms = tot_fl = 0.0
            for (H, W, Cin, Cout, taps) in LAYERS:
                N = (images_for_f16 if F16 else images_for)(H, W, Cin)
                ms, _ = run(H, W, Cin, Cout, taps, N)
                fl = 3.0 * _lib.load().tp_conv_tf32x3_flops(N, H, W, Cin, Cout, taps)
                tot_ms += ms
                tot_fl += fl
                print(f"{1 - single:4d} {H:2d} {W:2d} {Cin:4d} {Cout:4d} {taps:4d} {N:6d} {ms:7.3f} {fl / ms / 1e9:10.1f}")
            print(f"# pair={1 - single}: weighted {tot_fl / tot_ms / 1e9:.1f} executed TF/s")
        _lib.set_tuning("conv_pair", 0)
        return
    print("variant promote  H  W  Cin Cout taps      N     ms  useful_TF/s  executed_TF/s")
    print("# (variant 2 = persistent kernel with position-major tiles disabled; FLOPs are the")
    print("#  nominal 2*pixels*Cout*9*Cin, so skipped padding taps show up as > peak)")
    for variant in ((0, 2) if F16 else (0, 2, 1)):
        for promote in ((2,) if variant == 2 else (1, 2, 4)):
            _lib.set_tuning("conv_variant", 1 if variant == 1 else 0)
            _lib.set_tuning("conv_pos_major", 1 if variant == 2 else 0)
            _lib.set_tuning("conv_promote", promote)
            tot_ms = tot_fl = 0.0
            for (H, W, Cin, Cout, taps) in LAYERS:
                N = (images_for_f16 if F16 else images_for)(H, W, Cin)
                ms, tf = run(H, W, Cin, Cout, taps, N)
                tot_ms += ms
                tot_fl += tf * ms
                print(f"{variant:7d} {promote:7d} {H:2d} {W:2d} {Cin:4d} {Cout:4d} {taps:4d} {N:6d} {ms:7.3f} {tf:10.1f} {3 * tf:10.1f}")
            print(f"# variant {variant} promote {promote}: weighted {tot_fl / tot_ms:.1f} useful TF/s")


if __name__ == "__main__":
    main()
