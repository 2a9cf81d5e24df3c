"""Stored-form prefix sweep: register-load kernel against the cp.async ring kernel, on the thirteen
first-tail-layer geometries of the VGG16 job (bs 256, the chunk size the Shapley driver picks).

    python profiles/profile_prefix_ring.py            # A/B table, outputs compared bit for bit
    python profiles/profile_prefix_ring.py ncu        # geometries 6 and 9, register kernel then auto

GB/s counts the algorithmic bytes: 4 B of contribution read per source element and 4 B (fp16 hi
+ lo) written per output element, per mask."""
import os, sys, statistics
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torchpruner_b200 import _lib

DEV = torch.device("cuda:0")
B = 256
# (C, Hp, Wp, C1, pool_after) of the layer that follows conv1 .. conv13
GEOMS = [(64, 32, 32, 64, True), (64, 16, 16, 128, False), (128, 16, 16, 128, True),
         (128, 8, 8, 256, False), (256, 8, 8, 256, False), (256, 8, 8, 256, True),
         (256, 4, 4, 512, False), (512, 4, 4, 512, False), (512, 4, 4, 512, True),
         (512, 2, 2, 512, False), (512, 2, 2, 512, False), (512, 2, 2, 512, True),
         (512, 1, 1, 512, False)]


def setup(C, Hp, Wp, C1, pool):
    g = torch.Generator(device=DEV).manual_seed(C * 131 + Hp)
    pixels = B * Hp * Wp
    M1 = pixels * C1
    m_out = (pixels // 4 if pool else pixels) * C1
    P = int(max(1, min(512, C, (2 << 30) // (4 * m_out))))
    T = torch.randn(C * M1, device=DEV, generator=g)
    perm = torch.randperm(C, device=DEV, generator=g)
    scale = torch.rand(C1, device=DEV, generator=g) + 0.5
    shift = torch.randn(C1, device=DEV, generator=g)
    base0 = T.view(C, M1).sum(0, dtype=torch.float64)
    return dict(T=T, perm=perm, scale=scale, shift=shift, base0=base0, pixels=pixels, M1=M1,
                m_out=m_out, P=P)


def launch(s, gm, hi, lo, base):
    C, Hp, Wp, C1, pool = gm
    _lib.prefix_contrib_stack(s["T"], s["perm"], 0, s["P"], base, s["scale"], s["shift"],
                              s["pixels"], C1, max(Hp, 1), max(Wp, 1), pool, True, hi, lo,
                              out_scale=4.0)


def timed(s, gm, iters=5):
    hi = torch.empty(s["P"] * s["m_out"], device=DEV, dtype=torch.float16)
    lo = torch.empty_like(hi)
    ms = []
    for i in range(iters + 1):
        base = s["base0"].clone()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); launch(s, gm, hi, lo, base); b.record(); b.synchronize()
        if i:
            ms.append(a.elapsed_time(b))
    return (statistics.median(ms) if ms else 0.0), hi, lo, base


if "ncu" in sys.argv:
    for gm in (GEOMS[5], GEOMS[8]):
        s = setup(*gm)
        for var in (1, 0):
            _lib.set_tuning("prefix_variant", var)
            timed(s, gm, iters=0)
    sys.exit(0)

# (name, prefix_variant, prefix_ring_cfg): pooled cfg 1/2/3 = depth 4 x 3 CTAs, depth 3 x 4 CTAs of 224,
# depth 8 x 1 CTA; unpooled cfg 1/2 = depth 8 x 4 CTAs, depth 16 x 2 CTAs (cfg 3 = cfg 1 there)
variants = [("regs", 1, 0), ("ring1", 2, 1), ("ring2", 2, 2), ("ring3", 2, 3), ("auto", 0, 0)]
print("C Hp Wp C1 pool | P | GB alg | " + " | ".join(f"{n} ms (TB/s)" for n, _, _ in variants) + " | same bits")
tot = {n: 0.0 for n, _, _ in variants}
weights = [25 * gm[0] for gm in GEOMS]     # masks per bench step of that layer
for gm, wgt in zip(GEOMS, weights):
    s = setup(*gm)
    gb = s["P"] * 4.0 * (s["M1"] + s["m_out"]) / 1e9
    ref, cells, same = None, [], True
    for name, var, thr in variants:
        _lib.set_tuning("prefix_variant", var)
        _lib.set_tuning("prefix_ring_cfg", thr)
        t, hi, lo, base = timed(s, gm)
        if ref is None:
            ref = (hi.clone(), lo.clone(), base.clone())
        else:
            same = same and torch.equal(hi, ref[0]) and torch.equal(lo, ref[1]) and torch.equal(base, ref[2])
        cells.append(f"{t:7.3f} ({gb / t:5.2f})")
        tot[name] += t / s["P"] * wgt
        del hi, lo, base
    print(*gm, "|", s["P"], f"| {gb:5.2f} |", " | ".join(cells), "|", same, flush=True)
    del s, ref
    torch.cuda.empty_cache()
print("per bench step (25 permutations, all 13 layers), ms:", {k: round(v, 1) for k, v in tot.items()})
