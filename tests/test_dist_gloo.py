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
