"""Per-layer device time of one sharded bench step (run under torchrun): where an N-GPU step
spends time beyond the kernels a single GPU runs."""
import os, sys, time
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.argv = ["bench.py"]
import bench
from torchpruner_b200 import _lib, workloads
from torchpruner_b200.attributions import ShapleyAttributionMetric
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
dev = torch.device(f"cuda:{local}"); torch.cuda.set_device(dev)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
model, shape = bench.build_model(dev)
hx, hy = bench.host_batch(shape, pin=False)
dx, dy = hx.to(dev), hy.to(dev)
SV = bench.PERMS_PER_STEP * world
rows = []
ar_log = []
sect = []
if os.environ.get("PROBE_AR"):
    from torchpruner_b200.attributions import native_tail as _nt
    from torchpruner_b200.attributions import attributions as _at
    def wrap(cls, name):
        orig = getattr(cls, name)
        def f(self, *a, **k):
            torch.cuda.synchronize(); t0 = time.perf_counter()
            out = orig(self, *a, **k)
            torch.cuda.synchronize(); sect.append((name, (time.perf_counter() - t0) * 1e3))
            return out
        setattr(cls, name, f)
    wrap(_nt.NativeTail, "prepare_batch"); wrap(_nt.NativeTail, "unmasked_logits")
    wrap(_at._AttributionMetric, "run_forward_partial")
if world > 1 and os.environ.get("PROBE_AR"):
    from torchpruner_b200.attributions.methods import shapley_values as _sv
    def timed_allreduce(sv, group=None):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        dist.barrier(); torch.cuda.synchronize(); t1 = time.perf_counter()
        dist.all_reduce(sv, op=dist.ReduceOp.SUM, group=group)
        torch.cuda.synchronize(); t2 = time.perf_counter()
        ar_log.append((tuple(sv.shape), (t1 - t0) * 1e3, (t2 - t1) * 1e3))
        return sv
    _sv.allreduce_scores = timed_allreduce
def step(s, record):
    np.random.seed(1000 + s)
    m = ShapleyAttributionMetric(model, bench.OneBatch(dx, dy), workloads.cross_entropy, dev, sv_samples=SV, reduction="none")
    m.shard = world > 1
    for i, conv in enumerate(model.conv_layers()):
        ev = m.find_evaluation_module(conv, find_best_evaluation_module=True)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); t0 = time.perf_counter(); a.record()
        m._scores_on_device(ev, SV)
        b.record(); torch.cuda.synchronize(); t1 = time.perf_counter()
        if record:
            rows.append((i, a.elapsed_time(b), (t1 - t0) * 1e3, dict(m.last_run_info)))
with bench.quiet():
    step(0, False); step(1, False)
    step(2, True)
if ar_log:
    os.makedirs("gpurun_out", exist_ok=True)
    with open(f"gpurun_out/probe_rank{rank}.txt", "w") as fh:
        fh.write(repr([(a, round(b, 2), round(c, 2)) for a, b, c in ar_log[-13:]]) + "\n")
        fh.write(repr([(n, round(t, 2)) for n, t in sect[-39:][:9]]) + "\n")
    print(f"rank {rank} allreduce (shape, wait-for-peers ms, collective ms):", [(a, round(b, 2), round(c, 2)) for a, b, c in ar_log[-13:][:4]], flush=True)
print(f"rank {rank}: total dev {sum(r[1] for r in rows):.1f} ms", flush=True)
if rank == 0:
    for i, d, h, info in rows:
        print(f"{i:2d} dev {d:8.2f} ms host {h:8.2f} ms evals {info['evaluations']} cached {info['native'].get('cached')} f16 {info['f16']} incr {info['incremental_first']}")
    print("total", sum(r[1] for r in rows))
if world > 1:
    dist.destroy_process_group()
