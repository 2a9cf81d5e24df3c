"""Stored vs direct form of the incremental first layer on the VGG16-CIFAR geometries.

    python profiles/profile_prefix.py [ncu]

Per geometry: one chunk of P masks (the Shapley driver's stack size), CUDA-event timed."""
import os, sys, statistics
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torchpruner_b200 import _lib

if os.environ.get("TP_LIB"):   # an alternative build of the library (A/B experiments)
    _lib.LIB_PATH = os.path.abspath(os.environ["TP_LIB"])

DEV = torch.device("cuda:0")
B = 256
# (C, Hp, Wp, C1, pool_after): first tail layer when scoring conv1, conv2(pool first), conv5, conv8, conv10
GEOMS = [(64, 32, 32, 64, True), (64, 16, 16, 128, False), (256, 8, 8, 256, False),
         (512, 4, 4, 512, False), (512, 4, 4, 512, True), (512, 2, 2, 512, False)]


def masks_for(elems_per_mask):
    return int(max(1, min(512, (512 << 20) // (elems_per_mask * 4))))


def run(C, Hp, Wp, C1, pool, what, iters=3):
    g = torch.Generator(device=DEV).manual_seed(C)
    z = torch.relu(torch.randn((B, C, Hp, Wp), device=DEV, generator=g))
    wt = torch.randn((C, 9, C1), device=DEV, generator=g) / 3.0
    scale = torch.ones(C1, device=DEV); shift = torch.zeros(C1, device=DEV)
    pixels = B * Hp * Wp
    M1 = pixels * C1
    m_out = (pixels // 4 if pool else pixels) * C1
    P = min(C, masks_for(m_out))
    perm = torch.randperm(C, device=DEV, generator=g)
    hi = torch.empty(P * m_out, device=DEV, dtype=torch.float16); lo = torch.empty_like(hi)
    base = torch.zeros(M1, dtype=torch.float64, device=DEV)
    if what == "stored":
        T = torch.empty(C * M1, device=DEV)
        _lib.channel_contrib(z, wt, False, 0, T)
        fn = lambda: _lib.prefix_contrib_stack(T, perm, 0, P, base, scale, shift, pixels, C1, Hp, Wp, pool, True, hi, lo, out_scale=4.0)
    else:
        zc = _lib.channel_major(z, False)
        fn = lambda: _lib.prefix_contrib_direct(zc, wt, perm, 0, P, base, scale, shift, B, C, Hp, Wp, 0, pool, True, out_hi=hi, out_lo=lo, out_scale=4.0)
    ms = []
    for i in range(iters + 1):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize()
        if i: ms.append(a.elapsed_time(b))
    t = statistics.median(ms)
    return P, t, 4.0 * P * m_out / t / 1e6   # GB/s of output written


if "ncu" in sys.argv:
    for gm in (GEOMS[2], GEOMS[4]):
        run(*gm, "stored" if "stored" in sys.argv else "direct", iters=1)
else:
    print("C Hp Wp C1 pool | P | stored ms (GB/s out) | direct ms (GB/s out)")
    for gm in GEOMS:
        P, ts, gs = run(*gm, "stored")
        _, td, gd = run(*gm, "direct")
        print(*gm, "|", P, f"| {ts:7.3f} ({gs:6.0f}) | {td:7.3f} ({gd:6.0f})")
