"""world_size-2 test of the sharded Shapley path's host logic over gloo on CPU.

Each rank evaluates only the (permutation, prefix) spans `shard_spans` assigns to it — including
the halo evaluation that makes loss differences local — using the CPU oracle as the evaluator,
then the product's `allreduce_scores` combines the partial tables.  The sum must equal the
unsharded oracle run.  (The CUDA evaluator is covered by the -m gpu tests; what is exercised
here is the partition, the halo rule and the collective.)"""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import tp_oracle as oracle
        from torchpruner_b200 import workloads
        from torchpruner_b200.attributions.methods.shapley_values import (allreduce_scores,
                                                                          shard_spans)

        torch.set_num_threads(2)
        model, shape = workloads.build("small_cnn")
        loader = workloads.synthetic_batches(shape, n=6, batch_size=6, seed=2)
        x, y = next(iter(loader))
        module = model.features[4]  # ReLU after conv2: 64 channels
        samples, n = 3, 64
        np.random.seed(13)  # every rank draws the same permutations
        perms = [np.random.permutation(n) for _ in range(samples)]

        sv = np.zeros((6, n))
        with torch.no_grad():
            z = model.forward_partial(x, to_module=module)
            loss0 = workloads.cross_entropy(model.forward_partial(z, from_module=module), y,
                                            reduction="none").numpy()
        for j, k_begin, k_end in shard_spans(samples, n, world, rank):
            ks = list(range(max(k_begin - 1, 0), k_end))
            losses = oracle.shapley_eval_losses(model, x, y, workloads.cross_entropy, module,
                                                perms[j], ks)
            prev = loss0 if k_begin == 0 else losses[k_begin - 1]  # halo
            block = np.stack([losses[k] for k in range(k_begin, k_end)])
            oracle.loss_delta_scatter(block, prev, perms[j], k_begin, samples, sv, 0)

        table = torch.from_numpy(sv)
        allreduce_scores(table)
        if rank == 0:
            full = oracle.shapley_with_partial(model, loader, workloads.cross_entropy, samples, 6,
                                               module, permutations=perms)
            np.save(os.path.join(out_dir, "sharded.npy"), table.numpy())
            np.save(os.path.join(out_dir, "full.npy"), full)
    finally:
        dist.destroy_process_group()


def test_sharded_scores_sum_to_the_unsharded_result(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    sharded = np.load(tmp_path / "sharded.npy")
    full = np.load(tmp_path / "full.npy")
    scale = np.abs(full).max()
    assert scale > 0
    # identical per-evaluation losses; only the fp64 summation order differs across ranks
    assert np.abs(sharded - full).max() <= 1e-12 * max(scale, 1.0)


def _flag_worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from torchpruner_b200.attributions.methods.shapley_values import any_rank

        got = [any_rank(rank == 1, "cpu"), any_rank(False, "cpu"), any_rank(True, "cpu")]
        with open(os.path.join(out_dir, f"flags{rank}.txt"), "w") as f:
            f.write(repr(got))
    finally:
        dist.destroy_process_group()


def test_ranks_agree_on_repeating_a_run(tmp_path):
    """The fp16 -> TF32 rerun contains collectives: a flag raised on ONE rank (its plan ran the
    fp16 flavour) must be seen by all, or the ranks would disagree on whether to run again."""
    world = 2
    mp.spawn(_flag_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    for rank in range(world):
        assert open(tmp_path / f"flags{rank}.txt").read() == "[True, False, True]"


class _FakePlan:
    """Stands in for a NativeTail on the CPU: the host logic under test only reads these."""

    def __init__(self, use_f16=True, incremental=True, head_out=None):
        self.use_f16, self.incremental, self.head_out = use_f16, incremental, head_out
        self.released = 0
        self.exact = [{"pool_before": False, "cout": 1}, {"pool_before": False, "cout": 1}]

    def release(self):
        self.released += 1
        self.incremental = False

    def run_steps(self, z, start, stop=2):   # step 0: 2 * ch0 + ch1; step 1: + 1
        outs = []
        for i in range(start, stop):
            z = z[:, :1] * 2.0 + z[:, 1:2] if i == 0 else z + 1.0
            outs.append(z)
        return outs


def _plan_worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from torchpruner_b200.attributions.methods.shapley_values import (agree_on_plan,
                                                                          sharded_steps)

        lines = []
        # rank 1 lost its plan (verification a hair over the tolerance): both take the library path
        mine = None if rank == 1 else _FakePlan()
        got = agree_on_plan(mine, True, "cpu")
        lines.append(repr((got is None, mine.released if mine else None)))
        # rank 0's plan is on the TF32 flavour, rank 1's cannot run the incremental first layer
        mine = _FakePlan(use_f16=rank == 1, incremental=rank == 0)
        got = agree_on_plan(mine, True, "cpu")
        lines.append(repr((got is mine, got.use_f16, got.incremental)))
        # a rank without work does not veto
        mine = _FakePlan() if rank == 0 else None
        got = agree_on_plan(mine, rank == 0, "cpu")
        lines.append(repr(got is mine))
        # the batch's activation: each rank computes ceil(5 / 2) rows, one all-gather
        x = torch.arange(5 * 2 * 3 * 3, dtype=torch.float32).view(5, 2, 3, 3)
        outs = sharded_steps(_FakePlan(head_out=(1, 3, 3)), x, 0, rank, world)
        want = x[:, :1] * 2.0 + x[:, 1:2]
        ok = len(outs) == 2 and torch.equal(outs[0], want) and torch.equal(outs[1], want + 1.0)
        later = sharded_steps(_FakePlan(head_out=(1, 3, 3)), outs[0], 1, rank, world)
        ok = ok and len(later) == 1 and torch.equal(later[0], want + 1.0)
        lines.append(repr(bool(ok) and tuple(outs[1].shape)))
        with open(os.path.join(out_dir, f"plan{rank}.txt"), "w") as f:
            f.write("\n".join(lines))
    finally:
        dist.destroy_process_group()


def test_ranks_agree_on_the_execution_plan_and_share_the_activation(tmp_path):
    world = 2
    mp.spawn(_plan_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    want0 = ["(True, 1)", "(True, False, False)", "True", "(5, 1, 3, 3)"]
    want1 = ["(True, None)", "(True, False, False)", "True", "(5, 1, 3, 3)"]
    assert open(tmp_path / "plan0.txt").read().splitlines() == want0
    assert open(tmp_path / "plan1.txt").read().splitlines() == want1
