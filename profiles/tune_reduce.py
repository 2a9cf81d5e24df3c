"""Sweep the row-reduction kernels' variants / ring geometry on the BASELINE shapes.

    python profiles/tune_reduce.py [quick]  > gpurun_out/tune_reduce.txt

Timing: CUDA events around single launches; before every timed launch the L2 is refilled with
clean lines by READING a 512 MB buffer (a write-based flush leaves 126 MB of dirty lines whose
write-back competes with the measured reads)."""
import itertools
import json
import os
import statistics
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torchpruner_b200 import _lib  # noqa: E402

DEV = torch.device("cuda:0")
PEAK = 6548.8
SHAPES = [(256, 64, 32, 32), (256, 128, 16, 16), (256, 256, 8, 8), (256, 512, 4, 4), (256, 512, 2, 2),
          (256, 64, 112, 112), (256, 256, 56, 56), (256, 512, 28, 28), (256, 1024, 14, 14),
          (256, 2048, 7, 7)]


def main():
    quick = "quick" in sys.argv
    flush = torch.empty(128 << 20, dtype=torch.float32, device=DEV).normal_()
    stream = torch.cuda.current_stream(DEV)

    def timed(fn, iters=7):
        for _ in range(2):
            fn()
        ms = []
        for _ in range(iters):
            flush.sum()  # read-only L2 refill
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            fn()
            b.record(stream)
            b.synchronize()
            ms.append(a.elapsed_time(b))
        return statistics.median(ms)

    configs = [("ldg", 1, 0, 0, 0)]
    for stages, tile, ctas in itertools.product((3, 4, 6, 8), (2048, 4096, 8192), (1, 2, 3)):
        configs.append(("tma", 2, stages, tile, ctas))
    if quick:
        configs = [c for c in configs if c[0] == "ldg" or (c[2] in (4, 8) and c[4] in (1, 2))]
    rows = []
    for shp in SHAPES:
        z = torch.relu(torch.randn(shp, device=DEV))
        g = torch.randn(shp, device=DEV)
        n, r = z.numel(), shp[0] * shp[1]
        ops = {"taylor": (lambda: _lib.taylor_reduce(z, g), 8 * n + 4 * r),
               "apoz": (lambda: _lib.apoz_count(z), 4 * n + 4 * r)}
        for name, variant, stages, tile, ctas in configs:
            _lib.set_tuning("reduce_variant", variant)
            if variant == 2:
                _lib.set_tuning("reduce_stages", stages)
                _lib.set_tuning("reduce_tile_floats", tile)
                _lib.set_tuning("reduce_ctas_per_sm", ctas)
            for op, (fn, nbytes) in ops.items():
                try:
                    ms = timed(fn)
                except Exception as e:  # a geometry the launcher refuses
                    print("skip", shp, name, stages, tile, ctas, op, str(e)[:60])
                    continue
                gbs = nbytes / ms / 1e6
                rows.append({"shape": "x".join(map(str, shp)), "op": op, "impl": name, "stages": stages,
                             "tile": tile, "ctas": ctas, "us": round(ms * 1e3, 2), "gbs": round(gbs, 1),
                             "frac": round(gbs / PEAK, 3)})
        del z, g
        torch.cuda.empty_cache()
    # best per (shape, op)
    best = {}
    for r in rows:
        k = (r["shape"], r["op"])
        if k not in best or r["gbs"] > best[k]["gbs"]:
            best[k] = r
    print("# best per shape/op")
    for k, r in best.items():
        print(json.dumps(r))
    print("# ldg per shape/op")
    for r in rows:
        if r["impl"] == "ldg":
            print(json.dumps(r))
    print("# all")
    for r in rows:
        print(json.dumps(r))


if __name__ == "__main__":
    main()
