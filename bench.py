#!/usr/bin/env python
"""Benchmark of the attribution-scoring hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (config 3 of BASELINE.json, the one the headline metric is quoted on): VGG16-BN for
CIFAR-10 shaped inputs, one batch of 256 synthetic examples, layer-wise Monte-Carlo Shapley
scoring of all 13 conv layers (evaluated after BN+ReLU, find_best_evaluation_module=True).

One STEP = the whole job of the config: sv_samples=25 permutations (C prefix masks each) for each
of the 13 layers = 105600 channel-evals, sharded over the N ranks by the product's own
`shard_spans` (three whole permutations per rank at N=8, the 25th cut in eight, one halo
evaluation per cut) and ended by the all-reduce of each layer's score table.  The job is fixed
as N grows: scaling is STRONG.

Printed keys follow the driver's contract; see DESIGN.md "Measurement" for the definitions of
`roofline`, `cpu_baseline`, `e2e` and `reductions`.
"""
import argparse
import contextlib
import io
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from torchpruner_b200 import workloads  # noqa: E402

BATCH = 256
SV_SAMPLES = 25      # the config's sv_samples: one step is the whole fixed job, at any N
EVALS_PER_STEP = SV_SAMPLES * 4224   # 25 * sum of the 13 layers' channel counts
METRIC = "shapley_channel_evals_per_s"
UNIT = "channel-evals/s"


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def tensor_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return float(p["bf16_tflops_sustained"]), "measured (MEASURED_PEAKS.json)"
    return 1400.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel):
    """{"traffic": bytes per launch, "traffic_note": where it comes from} for `kernel`, from
    profiles/ncu_traffic.json (written by profiles/summarize_ncu.py from an `ncu --set full`
    capture of this round), else traffic null."""
    path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        with open(path) as f:
            entry = json.load(f)[kernel]
        return {"traffic": float(entry["dram_bytes_per_launch"]),
                "traffic_note": f"{entry['what']}; algorithmic {entry['algorithmic_bytes']:.4e} B; "
                                f"{entry['source']}"}
    except (OSError, KeyError, ValueError):
        return {"traffic": None, "traffic_note": "no committed ncu capture of this kernel flavour"}


def workload_description(n_gpus):
    return {
        "workload": "C3 VGG16-CIFAR10 layerwise Shapley: bs 256, 13 conv layers scored after "
                    "BN+ReLU, sv_samples=25 = 105600 channel-evals per step, the same job at every "
                    "N (sharded over the ranks: whole permutations first, the remainder cut, one "
                    "halo evaluation per cut, one all-reduce per layer); layers after the scored "
                    "module run fp32-accurate (tcgen05 engine: 3 MMAs per fp32 product on an "
                    "11-bit fp16 split, TF32 split as fallback; verified against forward_partial; "
                    "cuDNN/cuBLAS TF32 off); classifier at He scale so that scores are not "
                    "rounding noise",
        "batch": BATCH, "layers": 13, "channel_evals_per_step": EVALS_PER_STEP,
        "sv_samples": SV_SAMPLES, "parallelism": f"(permutation, prefix)-sharded x{n_gpus}",
        "l2": "working set per step (stacked masked activations, up to 512 MiB per chunk) far "
              "exceeds the 126 MB L2; reduction kernels are timed with an L2 flush between "
              "iterations",
    }


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")
    NAMES = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")

    def __init__(self, indices, enabled=True):
        """One nvidia-smi process for all the job's GPUs, on rank 0 only: every query takes a
        driver lock that kernel launches also need, and eight samplers at 5 Hz were measured to
        stretch an 8-GPU step by 60 %."""
        self.rows, self.proc, self.indices, self.enabled = [], None, indices, enabled

    def __enter__(self):
        if not self.enabled:
            return self
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "500", "-i", ",".join(str(i) for i in self.indices)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *exc):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
            self.thread.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, cell in zip(self.NAMES, r[3:7]):
                if cell.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------- model / data
def build_model(device):
    model, shape = workloads.build("vgg16_cifar", seed=1)
    model.to(device)
    return model, shape


def host_batch(shape, pin):
    g = torch.Generator().manual_seed(1)
    x = torch.randn((BATCH,) + tuple(shape), generator=g)
    y = torch.randint(0, 10, (BATCH,), generator=g)
    if pin:
        x, y = x.pin_memory(), y.pin_memory()
    return x, y


class OneBatch:
    """Minimal loader: an iterable of (x, y) with the `.dataset` length the Shapley metric
    reads (reference shapley_values.py:34)."""

    def __init__(self, x, y):
        self.x, self.y = x, y
        self.dataset = range(x.shape[0])

    def __iter__(self):
        yield self.x, self.y


def quiet():
    return contextlib.redirect_stdout(io.StringIO())


# --------------------------------------------------------------------------- product arm
def time_reductions(device, hbm_peak):
    """Taylor / APoZ / Sensitivity kernels on the activation shapes of config 2 (VGG16, bs 256)
    and the largest config-4 shape, in isolation, timed with CUDA events on the launch stream.

    * per shape: one launch per timed region;
    * config-2 sweep: the 13 layers' launches back to back in one timed region (13 distinct
      tensors, 283-566 MB in total, so nothing is re-read from L2 inside the region).
    Before every timed region the L2 is refilled with clean lines by READING a 512 MB buffer (a
    write-based flush would leave 126 MB of dirty lines whose write-back competes with the reads
    being measured)."""
    from torchpruner_b200 import _lib

    layer_shapes = [(256, 64, 32, 32)] * 2 + [(256, 128, 16, 16)] * 2 + [(256, 256, 8, 8)] * 3 + \
        [(256, 512, 4, 4)] * 3 + [(256, 512, 2, 2)] * 3
    big = (256, 64, 112, 112)  # ResNet-50 stem output, 822 MB
    flush = torch.empty(128 << 20, dtype=torch.float32, device=device).normal_()
    stream = torch.cuda.current_stream(device)

    def timed(fn, iters=9):
        for _ in range(2):
            fn()
        ms = []
        for _ in range(iters):
            # ~0.5 ms of queued read-only work: refills the L2 with clean lines and lets the host
            # enqueue the whole timed region before the GPU reaches it (device-side timing)
            for _ in range(6):
                flush.sum()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            fn()
            b.record(stream)
            b.synchronize()
            ms.append(a.elapsed_time(b))
        return statistics.median(ms)

    zs = [torch.relu(torch.randn(s, device=device)) for s in layer_shapes]
    gs = [torch.randn(s, device=device) for s in layer_shapes]
    ops = {
        "taylor": (lambda i: _lib.taylor_reduce(zs[i], gs[i]), 8),
        "apoz": (lambda i: _lib.apoz_count(zs[i]), 4),
        "sensitivity": (lambda i: _lib.abs_sum(gs[i]), 4),
    }
    # the 13 layers in ONE launch (tp_reduce_rows_multi, what `run_many` issues per batch)
    multi = {
        "taylor": lambda: _lib.reduce_rows_multi("taylor", list(zip(zs, gs))),
        "apoz": lambda: _lib.reduce_rows_multi("apoz", zs),
        "sensitivity": lambda: _lib.reduce_rows_multi("abs_sum", gs),
    }
    out = {}
    for op, (fn, mult) in ops.items():
        nbytes = [mult * int(np.prod(s)) + 4 * s[0] * s[1] for s in layer_shapes]
        multi_ms = timed(multi[op])
        sweep_ms = timed(lambda: [fn(i) for i in range(13)])
        # the same 13 launches captured once in a CUDA graph and replayed: launch-bound inner
        # loops belong in graphs; this removes the host enqueue cost from the device timeline
        graph_ms = None
        try:
            side = torch.cuda.Stream(device)
            side.wait_stream(stream)
            with torch.cuda.stream(side):
                [fn(i) for i in range(13)]
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=side):
                    [fn(i) for i in range(13)]
            stream.wait_stream(side)
            graph_ms = timed(g.replay)
        except Exception as e:  # noqa: BLE001 - capture support is a bonus, not a requirement
            log(f"bench.py: CUDA-graph sweep skipped for {op}: {e}")
        per = {}
        for i in (0, 2, 4, 7, 10):
            per["x".join(map(str, layer_shapes[i]))] = round(nbytes[i] / timed(lambda: fn(i)) / 1e6, 1)
        out[op] = {
            "c2_sweep_gbs": sum(nbytes) / multi_ms / 1e6,
            "c2_sweep_frac": sum(nbytes) / multi_ms / 1e6 / hbm_peak,
            "c2_sweep_us": multi_ms * 1e3, "c2_sweep_bytes": sum(nbytes), "c2_sweep_launches": 1,
            "c2_sweep_note": "all 13 layers' tensors in one tp_reduce_rows_multi launch, eager "
                             "(no CUDA graph); *_13_launches = one launch per layer",
            "c2_sweep_13_launches_gbs": sum(nbytes) / sweep_ms / 1e6,
            "c2_sweep_13_launches_frac": sum(nbytes) / sweep_ms / 1e6 / hbm_peak,
            "c2_sweep_13_launches_us": sweep_ms * 1e3,
            "c2_sweep_graph_gbs": (sum(nbytes) / graph_ms / 1e6) if graph_ms else None,
            "c2_sweep_graph_frac": (sum(nbytes) / graph_ms / 1e6 / hbm_peak) if graph_ms else None,
            "per_shape_gbs": per,
        }
    del zs, gs
    # Config 4: every convolution output of ResNet-50 at bs 256 (53 tensors, 11.4 GB of
    # activations: 25.7 MB ... 822 MB each), one launch per tensor as `run()` issues them.  Shapes
    # are read off torchvision's model; each distinct shape is timed once and weighted by how
    # often it occurs.  Rows of 7x7 = 49 floats are not 16-byte aligned.
    shapes = {}
    try:
        import torchvision
        net = torchvision.models.resnet50(num_classes=10).to(device).eval()
        hooks = [m.register_forward_hook(
            lambda _m, _i, o: shapes.__setitem__(tuple(o.shape[1:]), shapes.get(tuple(o.shape[1:]), 0) + 1))
            for m in net.modules() if isinstance(m, torch.nn.Conv2d)]
        with torch.no_grad():
            net(torch.zeros((1, 3, 224, 224), device=device))
        for h in hooks:
            h.remove()
        del net
    except Exception as e:  # noqa: BLE001
        log(f"bench.py: ResNet-50 shapes unavailable ({e}); timing the largest one only")
        shapes = {(64, 112, 112): 1}
    totals = {op: [0.0, 0.0] for op in ops}   # bytes, ms
    per_shape = {op: {} for op in ops}
    for (C, H, W), count in sorted(shapes.items(), key=lambda kv: -kv[0][0] * kv[0][1] * kv[0][2]):
        z = torch.relu(torch.randn((256, C, H, W), device=device))
        g = torch.randn((256, C, H, W), device=device)
        n, rows = z.numel(), 256 * C
        for op, fn, nb in (("taylor", lambda: _lib.taylor_reduce(z, g), 8 * n + 4 * rows),
                           ("apoz", lambda: _lib.apoz_count(z), 4 * n + 4 * rows),
                           ("sensitivity", lambda: _lib.abs_sum(g), 4 * n + 4 * rows)):
            ms = timed(fn, iters=5)
            totals[op][0] += count * nb
            totals[op][1] += count * ms
            per_shape[op][f"256x{C}x{H}x{W} (x{count})"] = round(nb / ms / 1e6, 1)
            if (C, H, W) == (64, 112, 112):
                out[op]["c4_largest_gbs"] = nb / ms / 1e6
                out[op]["c4_largest_frac"] = nb / ms / 1e6 / hbm_peak
        del z, g
    # the same 53 tensors as `run_many` would reduce them on a model whose activations survive the
    # pass (the hook-friendly ResNet-50): all resident at once, rows of up to 1024 floats (14x14,
    # 28x28) batched into multi-tensor launches, the rest one launch each
    try:
        zs53, gs53 = [], []
        for (C, H, W), count in shapes.items():
            for _ in range(count):
                zs53.append(torch.relu(torch.randn((256, C, H, W), device=device)))
                gs53.append(torch.randn((256, C, H, W), device=device))
        many = {"taylor": lambda: _lib.reduce_rows_multi("taylor", list(zip(zs53, gs53))),
                "apoz": lambda: _lib.reduce_rows_multi("apoz", zs53),
                "sensitivity": lambda: _lib.reduce_rows_multi("abs_sum", gs53)}
        for op in ops:
            ms = timed(many[op], iters=3)
            out[op]["c4_sweep_run_many_gbs"] = totals[op][0] / ms / 1e6
            out[op]["c4_sweep_run_many_frac"] = totals[op][0] / ms / 1e6 / hbm_peak
        del zs53, gs53
    except Exception as e:  # noqa: BLE001
        log(f"bench.py: batched config-4 sweep skipped: {e}")
    for op in ops:
        nbytes, ms = totals[op]
        out[op]["c4_sweep_gbs"] = nbytes / ms / 1e6
        out[op]["c4_sweep_frac"] = nbytes / ms / 1e6 / hbm_peak
        out[op]["c4_sweep_bytes"] = nbytes
        out[op]["c4_sweep_launches"] = sum(shapes.values())
        out[op]["c4_per_shape_gbs"] = per_shape[op]
    del flush
    torch.cuda.empty_cache()
    return out


def run_product(args, rank, world, local_rank):
    import torch.distributed as dist
    from torchpruner_b200 import _lib
    from torchpruner_b200.attributions import ShapleyAttributionMetric

    device = torch.device(f"cuda:{local_rank}")
    torch.cuda.set_device(device)
    torch.backends.cudnn.allow_tf32 = False        # fp32-accurate downstream layers
    torch.backends.cuda.matmul.allow_tf32 = False
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    hbm_peak, peak_src = peaks()

    model, shape = build_model(device)
    convs = model.conv_layers()
    hx, hy = host_batch(shape, pin=True)
    dx, dy = hx.to(device), hy.to(device)
    assert SV_SAMPLES * sum(c.out_channels for c in convs) == EVALS_PER_STEP

    def make_metric(loader):
        m = ShapleyAttributionMetric(model, loader, workloads.cross_entropy, device,
                                     sv_samples=SV_SAMPLES, reduction="none")
        m.shard = world > 1
        return m

    rank_evals = [0]

    from torchpruner_b200.attributions import native_tail

    def device_step(step):
        np.random.seed(1000 + step)  # identical permutations on every rank
        # Every step is the whole job from its input batch: activations cached by the previous
        # step are dropped, so the forward to each evaluation module is computed inside the step
        # (layer l + 1 continues from layer l's activation, as it would in any one job).
        native_tail._ACTS.pop(model, None)
        m = make_metric(OneBatch(dx, dy))
        rank_evals[0] = 0
        for conv in convs:
            module = m.find_evaluation_module(conv, find_best_evaluation_module=True)
            m._scores_on_device(module, SV_SAMPLES)
            rank_evals[0] += m.last_run_info["evaluations"]

    def e2e_step(step):
        np.random.seed(1000 + step)
        native_tail._ACTS.pop(model, None)
        m = make_metric(OneBatch(hx, hy))
        h2d = d2h = 0
        for conv in convs:
            sv = m.run(conv, find_best_evaluation_module=True)
            h2d += hx.numel() * 4 + hy.numel() * 8
            d2h += sv.nbytes
        return h2d, d2h

    def shard_check():
        """Outside every timed region: the table the sharded job all-reduces for one layer equals
        the table one rank computes alone (the same per-evaluation losses; only the order of the
        float64 additions differs)."""
        conv = convs[9]   # features.30: 512 channels at 4x4, engine + incremental first layer
        tables = []
        for shard in (True, False):
            np.random.seed(77)
            m = make_metric(OneBatch(dx, dy))
            m.shard = shard
            module = m.find_evaluation_module(conv, find_best_evaluation_module=True)
            tables.append(m._scores_on_device(module, SV_SAMPLES))
        diff = float((tables[0] - tables[1]).abs().max())
        scale = float(tables[1].abs().max())
        return {"layer": "features.30", "max_abs_diff": diff, "max_abs_score": scale,
                "ok": bool(diff <= 1e-12 * max(scale, 1.0))}

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(device)

    with quiet():
        for s in range(args.warmup):
            device_step(s)
        barrier()
        prof = _lib.profile_begin(timed=True)
        start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with ClockSampler(range(world) if world > 1 else [local_rank], enabled=rank == 0) as clocks:
            start.record(torch.cuda.current_stream(device))
            for s in range(args.steps):
                device_step(args.warmup + s)
            stop.record(torch.cuda.current_stream(device))
            barrier()
        _lib.profile_end()
        total_ms = start.elapsed_time(stop)
        kernel_summary = prof.summary()

        # end to end through the public API with host (pinned) inputs and host results
        e2e_steps = max(1, min(args.steps, 3))
        h2d = d2h = 0
        e2e_s = cold_s = float("nan")
        check = None
        if not args.core_only:
            e2e_step(-1)
            barrier()
            t0 = time.perf_counter()
            for s in range(e2e_steps):
                h2d, d2h = e2e_step(args.warmup + s)
            barrier()
            e2e_s = time.perf_counter() - t0
            # the same call as a user's FIRST one on this model: no cached plan, so every layer
            # pays the trace of forward_partial, the weight preparation (fold, permute, split),
            # the fp16 range calibration and the verification against forward_partial
            native_tail.forget_plans(model)
            barrier()
            t0 = time.perf_counter()
            e2e_step(args.warmup + e2e_steps)
            barrier()
            cold_s = time.perf_counter() - t0
            if world > 1:
                check = shard_check()

    t = torch.tensor([total_ms, e2e_s, cold_s], dtype=torch.float64, device=device)
    mine = t.clone()
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        per_rank = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(per_rank, mine)
        evals = torch.tensor([rank_evals[0]], dtype=torch.int64, device=device)
        per_rank_evals = [torch.zeros_like(evals) for _ in range(world)]
        dist.all_gather(per_rank_evals, evals)
    else:
        per_rank, per_rank_evals = [mine], [torch.tensor([rank_evals[0]])]
    total_ms, e2e_s, cold_s = float(t[0]), float(t[1]), float(t[2])
    ms_per_step = total_ms / args.steps
    value = EVALS_PER_STEP / (ms_per_step / 1e3)
    e2e_value = EVALS_PER_STEP * e2e_steps / e2e_s
    e2e_cold_value = EVALS_PER_STEP / cold_s

    result = None
    if rank == 0:
        own_ms = {k: v["ms"] for k, v in kernel_summary.items()}
        dominant = max(own_ms, key=own_ms.get)
        dom = kernel_summary[dominant]
        if dominant in ("conv_tf32x3", "conv_f16x3"):
            # fp32-accurate layers on tcgen05: every fp32 product is three tensor-core MMAs on an
            # 11-bit split of the operands (hi*hi + hi*lo + lo*hi), so the tensor pipe executes 3x
            # the layer's FLOPs.  fp16 operands run at the bf16/fp16 rate, TF32 ones at half of it.
            bf16 = tensor_peak()
            half = dominant == "conv_f16x3"
            peak = bf16[0] if half else bf16[0] / 2.0
            executed = 3.0 * dom["flops"] / dom["ms"] / 1e9 if dom["ms"] > 0 else 0.0
            roofline = {
                "kernel": dominant, "bound": "tensor", "achieved": executed, "peak": peak,
                "unit": "TFLOP/s", "frac": executed / peak,
                # dram__bytes_read.sum + dram__bytes_write.sum of one launch of this kernel, read
                # from the committed ncu --set full summary (a profiler cannot run inside the
                # timed region); null when no capture of this flavour is committed
                **ncu_traffic(dominant),
                "peak_source": bf16[1] + (" bf16 sustained (fp16 and bf16 MMAs run at the same rate)"
                                         if half else
                                         " bf16 sustained / 2 (TF32 dense runs at half the bf16 rate)"),
                "useful_tflops": executed / 3.0,
                "achieved_definition": "3 MMAs per fp32 product (hi*hi + hi*lo + lo*hi on an 11-bit "
                                       "split) x 2*MACs executed by the layer, summed over launches / "
                                       "summed duration",
                "launches": dom["launches"], "avg_launch_ms": dom["ms"] / max(1, dom["launches"]),
                "share_of_step": dom["ms"] / total_ms,
            }
        else:
            achieved = dom["bytes"] / dom["ms"] / 1e6 if dom["ms"] > 0 else 0.0
            roofline = {
                "kernel": dominant, "bound": "hbm", "achieved": achieved, "peak": hbm_peak,
                "unit": "GB/s", "frac": achieved / hbm_peak, "traffic": None,
                "peak_source": peak_src, "launches": dom["launches"],
                "avg_launch_ms": dom["ms"] / max(1, dom["launches"]),
                "share_of_step": dom["ms"] / total_ms,
                "note": "dominant hand-written kernel of the step; the rest of the step is the "
                        "user model's downstream layers in fp32 cuDNN/cuBLAS (library code)",
            }
        own = {k: {"launches": v["launches"], "ms_per_step": v["ms"] / args.steps,
                   "gbs": (v["bytes"] / v["ms"] / 1e6) if v["ms"] > 0 else None,
                   "tflops": (v["flops"] / v["ms"] / 1e9) if v["ms"] > 0 and v["flops"] else None}
               for k, v in kernel_summary.items()}
        with quiet():
            reductions = time_reductions(device, hbm_peak) if (world == 1 and not args.core_only) else None
            secondary = secondary_configs(device) if (world == 1 and not args.core_only) else None
        cpu = cpu_sample(steps=2, warmup=1) if (world == 1 and not args.core_only) else None
        result = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_description(world),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "steps": e2e_steps,
                    "api": "ShapleyAttributionMetric.run(conv, find_best_evaluation_module=True) "
                           "x13, pinned host batch in, numpy [256,C] float64 out"},
            "e2e_cold": {"value": e2e_cold_value, "unit": UNIT, "seconds": cold_s,
                         "what": "one step of the same API calls with the plan cache emptied "
                                 "first: a user's first run() on each layer (trace, weight "
                                 "preparation, fp16 calibration, verification included)"},
            "per_rank": {"ms_per_step": [float(p[0]) / args.steps for p in per_rank],
                         "evaluations_per_step": [int(e[0]) for e in per_rank_evals],
                         "note": "device time of each rank's own timed loop (the job's time is "
                                 "the max); evaluations include one halo per cut"},
            "shard_check": check,
            "gpu_launches": int(sum(v["launches"] for v in kernel_summary.values())),
            "own_kernels": own,
            "roofline": roofline,
            "reductions": reductions,
            "secondary": secondary,
            "cpu_baseline": cpu,
            "clocks": clocks.summary(),
        }
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return result


def secondary_configs(device):
    """The other BASELINE configs, end to end through the public API with the CPU port beside
    them: config 2 = Taylor / APoZ / Sensitivity `run()` over all 13 conv
    layers at bs 256 (model forward/backward included); config 5 = the 15-module cascading prune
    of VGG16 with SGD momentum.  CPU legs are bounded samples, stated."""
    from oracle import tp_oracle as oracle
    from tests import cases
    from torchpruner_b200 import _lib
    from torchpruner_b200.attributions import (APoZAttributionMetric, SensitivityAttributionMetric,
                                               TaylorAttributionMetric)
    from torchpruner_b200.pruner import Pruner
    from torchpruner_b200.utils import find_best_module_for_attributions, get_vgg_pruning_graph

    out = {}
    model, shape = build_model(device)
    hx, hy = host_batch(shape, pin=True)
    loader = OneBatch(hx, hy)
    convs = model.conv_layers()
    classes = {"taylor": TaylorAttributionMetric, "apoz": APoZAttributionMetric,
               "sensitivity": SensitivityAttributionMetric}
    gpu = {}
    for name, cls in classes.items():
        m = cls(model, loader, workloads.cross_entropy, device)
        for conv in convs[:2]:
            m.run(conv, find_best_evaluation_module=True)  # warm-up
        model.zero_grad()
        torch.cuda.synchronize(device)
        prof = _lib.profile_begin(timed=False)
        t0 = time.perf_counter()
        for conv in convs:
            m.run(conv, find_best_evaluation_module=True)
        torch.cuda.synchronize(device)
        gpu[name] = {"seconds_13_layers": time.perf_counter() - t0,
                     "own_kernel_launches": sum(v[0] for v in prof.records.values())}
        _lib.profile_end()
        model.zero_grad()
    for name, cls in classes.items():   # the 13 layers from ONE pass (run_many; not in the reference)
        m = cls(model, loader, workloads.cross_entropy, device)
        m.run_many(convs[:2], find_best_evaluation_module=True)
        model.zero_grad()
        torch.cuda.synchronize(device)
        t0 = time.perf_counter()
        m.run_many(convs, find_best_evaluation_module=True)
        torch.cuda.synchronize(device)
        gpu[name]["seconds_13_layers_run_many"] = time.perf_counter() - t0
        model.zero_grad()
    # CPU port: first, middle and last conv layer (3 of 13; every layer costs one full
    # forward(+backward) of the same model, so the 13-layer time is 13/3 of the sample)
    cpu_model = _CPU_STATE.setdefault("model", workloads.build("vgg16_cifar", seed=1)[0])
    cx, cy = host_batch(shape, pin=False)
    sample = [cpu_model.conv_layers()[i] for i in (0, 6, 12)]
    for name in classes:
        t0 = time.perf_counter()
        for conv in sample:
            module = find_best_module_for_attributions(cpu_model, conv)
            oracle.hook_scores(cpu_model, [(cx, cy)], workloads.cross_entropy, module, name)
        dt = time.perf_counter() - t0
        cpu_model.zero_grad()
        gpu[name]["cpu_port_seconds_13_layers_extrapolated"] = dt * 13.0 / 3.0
        gpu[name]["speedup_vs_cpu_port"] = gpu[name]["cpu_port_seconds_13_layers_extrapolated"] / \
            gpu[name]["seconds_13_layers"]
    out["c2_scores_e2e"] = dict(gpu, cpu_cores=torch.get_num_threads(),
                                cpu_sample="layers features.0/20/40 of 13, x13/3")

    out["c1_small_cnn"] = config1(device)
    out["c4_scores_e2e"] = config4(device)

    # config 5
    def cascade(pruner_factory, dev):
        mdl, shp, opt = cases.prune_setup("vgg16_cifar", device=dev)
        pr = pruner_factory(mdl, shp, opt)
        graph = get_vgg_pruning_graph(mdl)
        np.random.seed(21)
        if dev != "cpu":
            torch.cuda.synchronize(device)
        t0 = time.perf_counter()
        for module, casc in graph:
            pr.prune_model(module, cases.prune_choice(module.weight.shape[0]), casc)
        if dev != "cpu":
            torch.cuda.synchronize(device)
        return time.perf_counter() - t0

    cascade(lambda m, s_, o: Pruner(m, s_, device, optimizer=o), device)  # warm-up
    prof = _lib.profile_begin(timed=False)
    t_gpu = cascade(lambda m, s_, o: Pruner(m, s_, device, optimizer=o), device)
    launches = {k: v[0] for k, v in prof.records.items()}
    _lib.profile_end()
    cascade(lambda m, s_, o: Pruner(m, s_, device, optimizer=o, cascade="probe"), device)
    t_probe = cascade(lambda m, s_, o: Pruner(m, s_, device, optimizer=o, cascade="probe"), device)
    t_cpu = cascade(lambda m, s_, o: oracle.OraclePruner(m, s_, optimizer=o), "cpu")
    out["c5_prune_cascade"] = {"seconds": t_gpu, "seconds_with_probe_forward": t_probe,
                               "cascade": "analytic after the first probe forward (module chain recorded "
                                          "there); seconds_with_probe_forward = the NaN trick "
                                          "for every module, as the reference does",
                               "cpu_port_seconds": t_cpu, "modules": 15,
                               "own_kernel_launches": launches, "cpu_cores": torch.get_num_threads()}
    return out


def config1(device):
    """Config 1 (small 2-conv + 2-fc CNN, 1x28x28, bs 64, N = 128): Taylor / APoZ / WeightNorm /
    Shapley(sv_samples=5) on both convolutions plus the 3-module cascading prune, through the
    public API, beside the CPU port on the host cores.  conv1's Shapley tail starts with a 3x3
    convolution with padding 2 — outside the engine's vocabulary — so it runs the model's own
    forward_partial (cuDNN / cuBLAS) on stacked masks; conv2's tail runs on the engine."""
    from oracle import tp_oracle as oracle
    from torchpruner_b200.attributions import (APoZAttributionMetric, ShapleyAttributionMetric,
                                               TaylorAttributionMetric, WeightNormAttributionMetric)
    from torchpruner_b200.pruner import Pruner
    from torchpruner_b200.utils import find_best_module_for_attributions

    def gpu_pass(model, shape):
        loader = workloads.synthetic_batches(shape, n=128, batch_size=64, pin=True)
        convs = [model.features[0], model.features[3]]
        engine = []
        for conv in convs:
            for cls in (TaylorAttributionMetric, APoZAttributionMetric, WeightNormAttributionMetric):
                cls(model, loader, workloads.cross_entropy, device).run(
                    conv, find_best_evaluation_module=True)
            np.random.seed(11)
            sh = ShapleyAttributionMetric(model, loader, workloads.cross_entropy, device, sv_samples=5)
            sh.run(conv, find_best_evaluation_module=True)
            engine.append(bool(sh.last_run_info["native_tail"]))
        np.random.seed(21)
        pruner = Pruner(model, shape, device)
        for module, casc in model.get_pruning_graph():
            pruner.prune_model(module, list(range(0, module.weight.shape[0], 4)), casc)
        torch.cuda.synchronize(device)
        return engine

    model, shape = workloads.build("small_cnn")
    gpu_pass(model.to(device), shape)   # warm-up (plans, cuDNN)
    model, shape = workloads.build("small_cnn")
    model.to(device)
    torch.cuda.synchronize(device)
    t0 = time.perf_counter()
    engine = gpu_pass(model, shape)
    t_gpu = time.perf_counter() - t0

    cpu, shape = workloads.build("small_cnn")
    loader = workloads.synthetic_batches(shape, n=128, batch_size=64)
    t0 = time.perf_counter()
    for conv in (cpu.features[0], cpu.features[3]):
        module = find_best_module_for_attributions(cpu, conv)
        for kind in ("taylor", "apoz"):
            oracle.hook_scores(cpu, list(loader), workloads.cross_entropy, module, kind)
        oracle.weight_norm(conv.weight.detach())
        np.random.seed(11)
        oracle.shapley_with_partial(cpu, loader, workloads.cross_entropy, 5, 128, module)
    np.random.seed(21)
    pr = oracle.OraclePruner(cpu, shape)
    for module, casc in cpu.get_pruning_graph():
        pr.prune_model(module, list(range(0, module.weight.shape[0], 4)), casc)
    t_cpu = time.perf_counter() - t0
    return {"seconds": t_gpu, "cpu_port_seconds": t_cpu, "speedup_vs_cpu_port": t_cpu / t_gpu,
            "cpu_cores": torch.get_num_threads(),
            "shapley_tail_on_engine": {"conv1": engine[0], "conv2": engine[1]},
            "what": "Taylor + APoZ + WeightNorm + Shapley(sv_samples=5) on conv1 and conv2, then "
                    "prune_model over the 3-entry graph; 128 examples in batches of 64"}


def config4(device):
    """Config 4 end to end: Taylor and APoZ `run(conv, find_best_evaluation_module=True)` on stock
    torchvision ResNet-50, 3x224x224, bs 256, for a sample of the 37 layers the reference is
    defined on (every run is a full forward(+backward) of the model at bs 256).  CPU port on a
    bs-16 batch, scaled by 16 (the per-example cost of a ResNet-50 pass does not depend on the
    batch size once the cores are busy)."""
    from oracle import tp_oracle as oracle
    from torchpruner_b200.attributions import APoZAttributionMetric, TaylorAttributionMetric
    from torchpruner_b200.utils import find_best_module_for_attributions

    picks = ("conv1", "layer1.0.conv2", "layer2.0.downsample.0", "layer4.2.conv2")
    model, shape = workloads.build("resnet50_stock")
    g = torch.Generator().manual_seed(4)
    hx = torch.randn((BATCH,) + tuple(shape), generator=g).pin_memory()
    hy = torch.randint(0, 10, (BATCH,), generator=g).pin_memory()
    out = {"layers": list(picks), "batch": BATCH, "layers_defined": 37}
    cpu_model = workloads.build("resnet50_stock")[0]
    cx, cy = hx[:16].clone(), hy[:16].clone()
    model.to(device)
    mods, cpu_mods = dict(model.named_modules()), dict(cpu_model.named_modules())
    for name, cls in (("taylor", TaylorAttributionMetric), ("apoz", APoZAttributionMetric)):
        m = cls(model, OneBatch(hx, hy), workloads.cross_entropy, device)
        m.run(mods[picks[0]], find_best_evaluation_module=True)   # warm-up
        model.zero_grad()
        torch.cuda.synchronize(device)
        t0 = time.perf_counter()
        for p in picks:
            m.run(mods[p], find_best_evaluation_module=True)
        torch.cuda.synchronize(device)
        t_gpu = (time.perf_counter() - t0) / len(picks)
        model.zero_grad()
        t0 = time.perf_counter()
        module = find_best_module_for_attributions(cpu_model, cpu_mods[picks[1]])
        oracle.hook_scores(cpu_model, [(cx, cy)], workloads.cross_entropy, module, name)
        t_cpu = (time.perf_counter() - t0) * (BATCH / 16)
        cpu_model.zero_grad()
        out[name] = {"seconds_per_layer": t_gpu, "cpu_port_seconds_per_layer_extrapolated": t_cpu,
                     "speedup_vs_cpu_port": t_cpu / t_gpu,
                     "seconds_37_layers_extrapolated": 37 * t_gpu}
    out["cpu_cores"] = torch.get_num_threads()
    out["cpu_sample"] = "one layer on 16 of the 256 examples, x16"
    del model
    torch.cuda.empty_cache()
    return out


# --------------------------------------------------------------------------- CPU arm
_CPU_STATE = {}


def cpu_sample(steps, warmup, evals_per_layer=6):
    """The reference's algorithm for the path (oracle port: the loop body of
    shapley_values.py:55-61 on torch CPU ops) on this host's cores, on a bounded sample:
    `evals_per_layer` prefix evaluations of one permutation for each of the 13 layers at bs 256.
    Per-evaluation cost does not depend on the mask, so the whole-job rate is extrapolated as
    sum_l n_l / sum_l (n_l * t_l) with n_l = 25*C_l evaluations."""
    from oracle import tp_oracle as oracle

    threads = torch.get_num_threads()
    if "model" not in _CPU_STATE:
        _CPU_STATE["model"] = workloads.build("vgg16_cifar", seed=1)[0]
    if "batch" not in _CPU_STATE:
        _CPU_STATE["batch"] = host_batch((3, 32, 32), pin=False)
    model = _CPU_STATE["model"]
    x, y = _CPU_STATE["batch"]
    convs = model.conv_layers()
    from torchpruner_b200.utils import find_best_module_for_attributions

    with quiet():
        modules = [find_best_module_for_attributions(model, c) for c in convs]
    if "cached" not in _CPU_STATE:  # once per batch in the reference too (shapley_values.py:42-43)
        with torch.no_grad():
            cached = []
            for module in modules:
                z0 = model.forward_partial(x, to_module=module)
                loss0 = workloads.cross_entropy(model.forward_partial(z0, from_module=module), y,
                                                reduction="none")
                cached.append((z0, loss0))
        _CPU_STATE["cached"] = cached
    rates = []
    t_begin = time.perf_counter()
    for it in range(warmup + steps):
        t_layers = []
        for conv, module, (z0, loss0) in zip(convs, modules, _CPU_STATE["cached"]):
            perm = np.random.RandomState(it).permutation(conv.out_channels)
            sv = np.zeros((BATCH, conv.out_channels))
            t0 = time.perf_counter()
            oracle.shapley_prefix_steps(model, z0, loss0, y, workloads.cross_entropy, module, perm,
                                        evals_per_layer, sv, 25)
            t_layers.append((time.perf_counter() - t0) / evals_per_layer)
        if it >= warmup:
            n = [25 * c.out_channels for c in convs]
            rates.append(sum(n) / sum(ni * ti for ni, ti in zip(n, t_layers)))
    wall = time.perf_counter() - t_begin
    value = statistics.median(rates)
    return {
        "value": value, "unit": UNIT, "cores": threads, "host_cpus": os.cpu_count(), "kind": "port",
        "sample": f"{evals_per_layer} prefix evaluations of one permutation per layer x 13 layers "
                  f"at bs 256 per step ({steps} timed step(s), {wall:.1f} s of CPU work); "
                  "whole-job rate extrapolated by per-layer evaluation counts (25*C_l); the "
                  "once-per-batch forward to the evaluation module is outside the timed loop",
        "seconds_per_step": wall / max(1, warmup + steps),
    }


def _host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference(args, rank, world):
    if rank != 0:
        return None
    # All the host threads this arm may use.  torchrun exports OMP_NUM_THREADS=1 and the OpenMP /
    # oneDNN pools are sized when torch is imported (torch.set_num_threads afterwards does not
    # bring them back), which would make the same CPU arm several times slower at N > 1 than at
    # N = 1: run it in a fresh interpreter with the variable set to the core count.
    want = str(_host_threads())
    if os.environ.get("OMP_NUM_THREADS", want) != want and "TP_BENCH_REF_CHILD" not in os.environ:
        env = {k: v for k, v in os.environ.items()
               if k not in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT",
                            "OMP_NUM_THREADS", "MKL_NUM_THREADS")}
        env.update(OMP_NUM_THREADS=want, TP_BENCH_REF_CHILD="1", TP_BENCH_REF_WORLD=str(world))
        out = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference",
                              "--gpus", str(args.gpus), "--steps", str(args.steps), "--warmup",
                              str(args.warmup)], env=env, stdout=subprocess.PIPE, text=True, check=True)
        return json.loads(out.stdout.strip().splitlines()[-1])
    world = int(os.environ.get("TP_BENCH_REF_WORLD", world))
    torch.backends.cudnn.allow_tf32 = False
    cpu = cpu_sample(steps=args.steps, warmup=args.warmup)
    value = cpu["value"]
    return {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": cpu["seconds_per_step"] * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_description(world), "cpu_baseline": cpu,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="product", choices=["product", "reference"])
    ap.add_argument("--core-only", action="store_true",
                    help="timed Shapley steps only (no e2e / reductions / CPU legs); for ncu runs")
    args = ap.parse_args()
    # Everything any library prints to fd 1 (NCCL's version banner, the reference-compatible
    # "... detected: shifting evaluation ..." lines) goes to stderr; the JSON line alone goes to
    # the real stdout.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world == 1 and args.gpus > 1:
        log(f"bench.py: --gpus {args.gpus} needs torchrun (one process per GPU); running 1 GPU")
    if args.impl == "reference":
        result = run_reference(args, rank, world)
    else:
        if args.warmup < 3:
            log("bench.py: timing rules ask for >= 3 warm-up steps")
        result = run_product(args, rank, world, local_rank)
    sys.stdout.flush()
    if rank == 0 and result is not None:
        os.write(real_stdout, (json.dumps(result) + "\n").encode())


if __name__ == "__main__":
    main()
