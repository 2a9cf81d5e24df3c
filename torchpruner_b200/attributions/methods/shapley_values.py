import logging
import os
import weakref

import numpy as np
import torch
from torch.nn.modules.batchnorm import _BatchNorm
from torch.nn.modules.dropout import _DropoutNd

from .. import attributions as _base
from ..native_tail import (NativeTail, cached_activation, cached_plan, head_plan, model_signature,
                           note_agreement, remember_plan)
from ... import _lib

_ZERO_BITS = 0x00000000


def shard_spans(sv_samples, n_channels, world_size, rank):
    """The half-open prefix ranges ``(j, k_begin, k_end)`` of permutation j that ``rank``
    evaluates; over all ranks every (permutation, prefix position) pair is covered exactly once.

    Whole permutations first: each rank gets ``sv_samples // world_size`` of them, which need no
    halo evaluation and no start inside a permutation.  Only the ``sv_samples % world_size``
    permutations left over (25 = 8 * 3 + 1) are cut: their (j, k) pairs, in j-major order, are
    split into ``world_size`` contiguous spans, so work per rank differs by at most one
    evaluation and a rank pays at most one halo per cut it starts behind.
    """
    spans = []
    if n_channels <= 0 or sv_samples <= 0:
        return spans
    whole = sv_samples // world_size
    spans += [(j, 0, n_channels) for j in range(rank * whole, (rank + 1) * whole)]
    first_cut = whole * world_size
    total = (sv_samples - first_cut) * n_channels
    lo = rank * total // world_size
    hi = (rank + 1) * total // world_size
    if hi > lo:
        for j in range(lo // n_channels, (hi - 1) // n_channels + 1):
            k_begin = max(lo, j * n_channels) - j * n_channels
            k_end = min(hi, (j + 1) * n_channels) - j * n_channels
            spans.append((first_cut + j, k_begin, k_end))
    return spans


def allreduce_scores(sv, group=None):
    """Sum the per-rank partial score tables in place (one collective per ``run``)."""
    import torch.distributed as dist

    dist.all_reduce(sv, op=dist.ReduceOp.SUM, group=group)
    return sv


def any_rank(flag, device, group=None):
    """True on every rank iff ``flag`` is true on at least one (one tiny MAX all-reduce).  Ranks
    must take collective decisions — such as repeating a run — together."""
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        return bool(flag)   # single process (or ranks emulated in one process by the tests)
    t = torch.tensor([1.0 if flag else 0.0], device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return bool(t.item())


def plan_vote(native, has_work, device, group=None):
    """(engine usable, fp16 flavour, incremental first layer) as the MINIMUM over the ranks — one
    tiny all-reduce — or None outside a process group (single process, or ranks emulated in one
    process by the tests).  A rank without work votes "don't care"."""
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        return None
    if has_work:
        mine = [int(native is not None), int(native is not None and native.use_f16),
                int(native is not None and native.incremental)]
    else:
        mine = [1, 1, 1]
    t = torch.tensor(mine, dtype=torch.int32, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
    return tuple(bool(v) for v in t.tolist())


def apply_vote(native, outcome):
    """The plan to run given what the ranks agreed on."""
    if native is None or outcome is None:
        return native
    usable, f16, incremental = outcome
    if not usable:
        native.release()
        return None
    if not f16:
        native.use_f16 = False
    if not incremental and native.incremental:
        native.release()   # drops the per-batch contributions; the first layer runs per mask
    return native


def agree_on_plan(native, has_work, device, group=None):
    """Make the execution plan of one (layer, batch) the same on every rank.  Ranks verify their
    plans on their own masks, and a verification that lands a hair on either side of its
    tolerance must not leave one rank on the library path (measured: one such rank made an 8-GPU
    step wait 700 ms at the all-reduce), nor may the tables mix roundings.  Returns the plan to
    use."""
    return apply_vote(native, plan_vote(native, has_work, device, group))


def sharded_steps(head, z, start, rank, world, group=None):
    """Outputs of the head's steps ``start`` .. last for the whole batch, each rank computing the
    rows of ``ceil(B / world)`` examples and one all-gather per step putting the batch together.
    Per-example arithmetic does not depend on which rank (or which slice) an example is in, so
    every rank ends up with the bits a single GPU computes.  Without this the once-per-batch
    forward to the evaluation module is replicated on every rank and does not shrink as GPUs are
    added (the serial fraction of the strong-scaling job)."""
    import torch.distributed as dist

    B = z.shape[0]
    if world <= 1 or B < 2 or not (dist.is_available() and dist.is_initialized()):
        return head.run_steps(z, start)
    chunk = -(-B // world)
    lo, hi = min(rank * chunk, B), min((rank + 1) * chunk, B)
    outs = []
    for i in range(start, len(head.exact)):
        mine = head.run_steps(z[lo:hi].contiguous(), i, i + 1) if hi > lo else []
        if mine:
            shape = tuple(mine[0].shape[1:])
        else:   # a rank without rows: the shape comes from the step itself
            e = head.exact[i]
            H, W = (z.shape[2] // 2, z.shape[3] // 2) if e["pool_before"] else (z.shape[2], z.shape[3])
            shape = (e["cout"], H, W)
        part = torch.zeros((chunk,) + shape, dtype=torch.float32, device=z.device)
        if mine:
            part[:hi - lo] = mine[0]
        gathered = torch.empty((world * chunk,) + shape, dtype=torch.float32, device=z.device)
        dist.all_gather_into_tensor(gathered, part, group=group)
        z = gathered[:B]
        outs.append(z)
    return outs


_WARNED_GENERIC = weakref.WeakSet()


def _warn_generic_path(model, module):
    """Say once per model that its layers after ``module`` are evaluated through the model's own
    ``forward_partial`` (library kernels, measured ~20x slower than the engine on VGG16)."""
    if model in _WARNED_GENERIC:
        return
    _WARNED_GENERIC.add(model)
    logging.warning(
        "torchpruner_b200: the layers after %s are outside the tensor-core engine's vocabulary "
        "(chains of 3x3/stride 1/pad 1 Conv2d, BatchNorm2d, ReLU/LeakyReLU, MaxPool2d(2), Flatten, "
        "Dropout, Linear in eval mode) or did not verify against forward_partial; Shapley masks "
        "are evaluated through the model's forward_partial instead (same results, slower)",
        type(module).__name__)


class ShapleyAttributionMetric(_base._AttributionMetric):
    """Monte-Carlo permutation-sampling Shapley values (reference:
    attributions/methods/shapley_values.py).

    Same estimator, same permutations (drawn from the global numpy RNG in the same call order),
    same sign convention (positive = removing the unit raises the loss), same fp32 loss
    arithmetic and fp64 accumulation order.  What changes is the schedule.  The reference
    evaluates one prefix mask at a time and synchronises with the host after each one
    (shapley_values.py:55-61); here, per batch and permutation,

      1. ``tp_prefix_mask_stack`` writes P cumulatively masked copies of the cached activation
         in one launch,
      2. the layers after the evaluation module run once on the stacked [P*B, ...] batch
         (the user's ``forward_partial`` — per-example maths is independent of its batch
         neighbours as long as BatchNorm/Dropout are in eval mode; otherwise each mask gets its
         own forward, in the reference's order),
      3. ``tp_loss_delta_scatter`` forms the adjacent-prefix loss differences and adds them to
         the fp64 score table that lives on the GPU,

    and nothing is copied to the host until the whole table is done.

    Optional sharding over the GPUs of one node: set ``TORCHPRUNER_B200_SHARD=1`` (or assign
    ``metric.process_group``) with ``torch.distributed`` initialised and every rank holding the
    same model, data and numpy seed.  (j, k) work items are split into contiguous spans, each
    rank evaluates one extra "halo" prefix so that loss differences stay local, and a single
    all-reduce (NCCL over NVLink) of the score table ends the run.
    """

    def __init__(self, *args, sv_samples=5, **kwargs):
        super().__init__(*args, **kwargs)
        self.samples = sv_samples
        self.mask_indices = []
        self.process_group = None
        self.shard = os.environ.get("TORCHPRUNER_B200_SHARD", "0") == "1"
        # Bytes of stacked masked activations per launch.  On the engine 2 GiB (measured 3 % faster
        # than 512 MiB on the VGG16 job: fewer, larger launches; the kernels index with 64 bits).
        # Through the model's own forward_partial 512 MiB: tensors of 2^31 bytes and more push
        # cuDNN onto a 25x slower direct kernel (profiles/r01_launches_bench_core_v1.csv).
        self.max_stack_bytes = int(os.environ.get("TORCHPRUNER_B200_STACK_MB", "512")) << 20
        self.max_stack_bytes_engine = int(os.environ.get(
            "TORCHPRUNER_B200_ENGINE_STACK_MB",
            os.environ.get("TORCHPRUNER_B200_STACK_MB", "2048"))) << 20
        self.max_stack = int(os.environ.get("TORCHPRUNER_B200_MAX_STACK", "512"))
        # "auto": run recognised conv/BN/ReLU/pool/linear tails on the tcgen05 3xTF32 engine after
        # checking it against forward_partial; False: always call forward_partial.
        self.native_tail = os.environ.get("TORCHPRUNER_B200_NATIVE", "1") != "0" and "auto"
        # first layer of the tail evaluated incrementally along each permutation (one channel's
        # contribution subtracted per mask) instead of once per mask
        self.incremental_first = os.environ.get("TORCHPRUNER_B200_INCREMENTAL", "1") != "0"
        # the layers BEFORE the evaluation module (input -> cached activation) on the engine as
        # well, when they are a recognised chain: the model's own fp32 forward on the GPU is the
        # largest single source of distance to the reference's tables (native_tail.head_plan)
        self.native_head = os.environ.get("TORCHPRUNER_B200_HEAD", "1") != "0"
        self.last_run_info = {}

    def run(self, module, sv_samples=None, **kwargs):
        module = super().run(module, **kwargs)
        sv_samples = sv_samples if sv_samples is not None else self.samples
        if hasattr(self.model, "forward_partial"):
            return self.run_module_with_partial(module, sv_samples)
        logging.warning("Consider adding a 'forward_partial' method to your model to speed-up "
                        "Shapley values computation")
        return self.run_module(module, sv_samples)

    # ------------------------------------------------------------------ helpers
    def _dist(self):
        """(group, world_size, rank) when sharding is requested and possible, else (None, 1, 0)."""
        if not (self.shard or self.process_group is not None):
            return None, 1, 0
        import torch.distributed as dist

        if not (dist.is_available() and dist.is_initialized()):
            return None, 1, 0
        group = self.process_group
        return group, dist.get_world_size(group), dist.get_rank(group)

    def _agreed(self, kind, plan, info, has_work, world, z, module, group):
        """(plan to run, agreed outcome or None): the ranks' agreement on this plan, by a
        collective the first time and from the plan cache afterwards (every rank stored the same
        outcome, so every rank takes the same branch here)."""
        key = (world, bool(getattr(self, "_f16_banned", False)), bool(self.incremental_first))
        outcome = ((info or {}).get("agreed") or {}).get(key)
        if outcome is None:
            outcome = plan_vote(plan, has_work, z.device, group)
            if outcome is not None:
                note_agreement(self.model, module, z, kind, key, outcome)
        return apply_vote(plan, outcome), outcome

    def _stack_safe(self):
        """Evaluating several masks as one batch is only the same computation if no layer mixes
        examples of a batch or draws per-call randomness: BatchNorm/Dropout must be in eval."""
        return not any(m.training and isinstance(m, (_BatchNorm, _DropoutNd))
                       for m in self.model.modules())

    @staticmethod
    def _per_example(loss, batch):
        if loss.numel() != batch:
            raise ValueError(
                f"criterion(..., reduction='none') returned {tuple(loss.shape)} for a batch of "
                f"{batch}; one loss per example is required")
        return loss.reshape(batch).float().contiguous()

    def _fused_ce_ok(self, logits, y, loss):
        """The fused cross-entropy path is taken only when it reproduces the user's criterion on
        this batch's unmasked logits."""
        if not (logits.dim() == 2 and y.dim() == 1 and y.dtype == torch.int64
                and logits.dtype == torch.float32 and y.numel() == logits.shape[0]):
            return False
        # (also for F.cross_entropy itself: targets equal to its ignore_index have loss 0 there and
        # are out of range for the kernel, which is exactly what the probe catches)
        B, K = logits.shape
        probe = torch.empty(B, dtype=torch.float32, device=logits.device)
        scratch_prev = torch.zeros(B, dtype=torch.float32, device=logits.device)
        dummy_sv = torch.zeros((B, 1), dtype=torch.float64, device=logits.device)
        dummy_perm = torch.zeros(1, dtype=torch.int64, device=logits.device)
        _lib.loss_delta_scatter(logits.contiguous(), y, _lib.TP_LOSS_CE, scratch_prev,
                                torch.empty_like(scratch_prev), dummy_perm, 0, 1, B, K, 1.0,
                                dummy_sv, 0, scatter_from=1, losses_out=probe)
        return bool(torch.allclose(probe, loss, rtol=2e-6, atol=2e-6))

    # ------------------------------------------------------------------ fast path
    def run_module_with_partial(self, module, sv_samples):
        """Shapley sampling for models that provide ``forward_partial``
        (shapley_values.py:28-64)."""
        sv = self._scores_on_device(module, sv_samples)
        if sv is None:  # empty loader: the reference would hand None to numpy as well
            return self.aggregate_over_samples(None)
        return self.aggregate_over_samples(sv.cpu().numpy())

    def _scores_on_device(self, module, sv_samples):
        """The [d, C] float64 score table, left on the GPU (all-reduced when sharded)."""
        rng_state = np.random.get_state()
        sv = self._scores_once(module, sv_samples)
        info = self.last_run_info
        if sv is not None and info.get("f16_any_rank") and not bool(torch.isfinite(sv).all()):
            # An activation left fp16's exponent range under some mask (the per-tensor scales are
            # calibrated on the unmasked batch, with 2^10 of headroom).  Redo the run on the TF32
            # flavour, with the numpy RNG rewound so the same permutations are drawn.
            self._f16_banned = True
            np.random.set_state(rng_state)
            sv = self._scores_once(module, sv_samples)
        return sv

    def _scores_once(self, module, sv_samples):
        d = len(self.data_gen.dataset)
        group, world, rank = self._dist()
        sv = None
        perm_dev = inv_dev = None
        spans = None
        row = 0
        n_evals = 0
        stack_safe = self._stack_safe()
        native = head = tail_vote = None
        native_info, head_info = {}, {}

        with torch.no_grad():
            for x, y in self._batches():
                # the evaluation module's output for this batch (shapley_values.py:42): on the
                # engine when the layers up to it are a recognised chain (see head_plan), else
                # through the model's forward_partial
                head, head_info = None, {}
                sig = model_signature(self.model) if (self.native_tail and stack_safe) else None
                if self.native_tail and self.native_head and stack_safe:
                    head, head_info = head_plan(self.model, module, x, sig=sig)
                if world > 1:
                    head, _ = self._agreed("head", head, head_info, True, world, x, module, group)
                if head is not None:
                    original_z = cached_activation(
                        self.model, sig, head, x,
                        lambda zz, start: sharded_steps(head, zz, start, rank, world, group))
                else:
                    original_z, _ = self.run_forward_partial(x, to_module=module)
                z = original_z.detach().float().contiguous()
                B, n = z.shape[0], z.shape[1]
                if perm_dev is None:
                    # Same permutations for all batches; same RNG calls as the reference (:47).
                    permutations = [np.random.permutation(n) for _ in range(sv_samples)]
                    perm_host = np.stack(permutations).astype(np.int64) if sv_samples else \
                        np.zeros((0, n), np.int64)
                    inv_host = np.empty_like(perm_host, dtype=np.int32)
                    for j in range(sv_samples):
                        inv_host[j, perm_host[j]] = np.arange(n, dtype=np.int32)
                    perm_dev = torch.from_numpy(perm_host).to(z.device)
                    inv_dev = torch.from_numpy(inv_host).to(z.device)
                    spans = shard_spans(sv_samples, n, world, rank)
                if sv is None:
                    sv = torch.zeros((d, n), dtype=torch.float64, device=z.device)

                # --- how the layers after `module` are executed for this batch
                # (in a sharded job even a rank without work keeps a plan: what the ranks agreed on
                # is stored with it, and every rank must find — or miss — that together)
                use_engine = bool(self.native_tail and stack_safe and sv_samples > 0 and
                                  (spans or world > 1))
                native, P_max, buf_hi, buf_lo = None, None, None, None
                banned = getattr(self, "_f16_banned", False)

                def engine_buffers():
                    per_copy = (4 if native.use_f16 else 8) * native.stack_elems(B, 1)  # hi + lo
                    p_max = int(max(1, min(self.max_stack, n,
                                           self.max_stack_bytes_engine // per_copy)))
                    hi = torch.empty(native.stack_elems(B, max(p_max, min(2, n))),
                                     dtype=native.stack_dtype, device=z.device)
                    return p_max, hi, torch.empty_like(hi)

                if use_engine:
                    found, native, info = cached_plan(self.model, module, z, sig=sig)
                    if found:
                        native_info = dict(info or {}, cached=True)
                        if native is not None:
                            native.use_f16 = native.use_f16 and not banned
                            native.prepare_batch(z, allow_incremental=self.incremental_first)
                    else:
                        native = NativeTail.build(self.model, module, z)
                        native_info = {"verified": False}
                        if native is not None:
                            native.use_f16 = native.use_f16 and not banned
                            native.prepare_batch(z, allow_incremental=self.incremental_first)
                            P_max, buf_hi, buf_lo = engine_buffers()
                            j0, k0v, _ = spans[0] if spans else (0, 0, 0)
                            ok, verr = native.verify(self.model, module, z, perm_dev[j0], inv_dev[j0],
                                                     k0v, min(2, n), buf_hi, buf_lo)
                            if not ok and native.use_f16:   # out of fp16 range? TF32 flavour
                                native.use_f16 = False
                                P_max, buf_hi, buf_lo = engine_buffers()
                                ok, verr = native.verify(self.model, module, z, perm_dev[j0],
                                                         inv_dev[j0], k0v, min(2, n), buf_hi, buf_lo)
                            native_info = {"verified": ok, "verify_rel_err": verr,
                                           "f16": native.use_f16}
                            if not ok:
                                native.release()
                                native = None
                        remember_plan(self.model, module, z, native, native_info, sig=sig)
                    if native is None:
                        _warn_generic_path(self.model, module)
                tail_vote = None
                if world > 1:
                    # every rank must run the same flavour, or the slowest plan sets the pace of
                    # the whole job at the all-reduce (and tables would mix roundings)
                    native, tail_vote = self._agreed("tail", native, native_info, True, world, z,
                                                     module, group)

                if native is not None:
                    # A plan verified against forward_partial for this very model state: the
                    # unmasked loss comes from it too, so the whole ladder is self-consistent and
                    # the library forward of the tail is not needed at all.
                    if buf_hi is None or buf_hi.dtype != native.stack_dtype:
                        P_max, buf_hi, buf_lo = engine_buffers()
                    logits0 = native.unmasked_logits(z, inv_dev[0], buf_hi, buf_lo)
                    original_loss = self.criterion(logits0, y, reduction="none")
                else:
                    buf_hi = buf_lo = None
                    logits0, original_loss = self.run_forward_partial(original_z, y_true=y,
                                                                      from_module=module)

                loss0 = self._per_example(original_loss, B)
                fused_ce = self._fused_ce_ok(logits0, y, loss0)
                prev_a = torch.empty(B, dtype=torch.float32, device=z.device)
                prev_b = torch.empty_like(prev_a)
                y_stack = None
                if native is None:
                    bytes_per_copy = max(1, z.numel() * 4)
                    P_max = int(max(1, min(self.max_stack, n, self.max_stack_bytes // bytes_per_copy)))
                    stack = torch.empty((P_max * B,) + tuple(z.shape[1:]), dtype=torch.float32,
                                        device=z.device)

                def evaluate(j, k0, P, prev_in, prev_out, scatter_from):
                    nonlocal y_stack
                    if native is not None:
                        out = native.logits(z, inv_dev[j], k0, P, buf_hi, buf_lo)
                    else:
                        masked = _lib.prefix_mask_stack(z, inv_dev[j], k0, P, out=stack)[:P * B]
                        if stack_safe:
                            out = self.model.forward_partial(masked, to_module=None,
                                                             from_module=module)
                        else:
                            out = torch.cat([
                                self.model.forward_partial(masked[p * B:(p + 1) * B], to_module=None,
                                                           from_module=module) for p in range(P)], 0)
                    if fused_ce:
                        _lib.loss_delta_scatter(out.float().contiguous(), y, _lib.TP_LOSS_CE,
                                                prev_in, prev_out, perm_dev[j], k0, P, B,
                                                out.shape[1], sv_samples, sv, row, scatter_from)
                    else:
                        if y_stack is None or y_stack.shape[0] != P * B:
                            y_stack = y.repeat((P,) + (1,) * (y.dim() - 1))
                        losses = self.criterion(out, y_stack, reduction="none")
                        _lib.loss_delta_scatter(self._per_example(losses, P * B), None,
                                                _lib.TP_LOSS_GIVEN, prev_in, prev_out, perm_dev[j],
                                                k0, P, B, 1, sv_samples, sv, row, scatter_from)

                self.set_deterministic()
                for (j, k_begin, k_end) in spans:
                    prev_in, prev_out = prev_a, prev_b
                    if native is not None:
                        native.begin(perm_dev[j], max(k_begin - 1, 0))
                    if k_begin == 0:
                        prev_in.copy_(loss0)
                    else:  # halo: loss with the first k_begin channels of permutation j removed
                        evaluate(j, k_begin - 1, 1, prev_out, prev_in, scatter_from=1)
                        n_evals += 1
                    for k0 in range(k_begin, k_end, P_max):
                        P = min(P_max, k_end - k0)
                        evaluate(j, k0, P, prev_in, prev_out, scatter_from=0)
                        prev_in, prev_out = prev_out, prev_in
                        n_evals += P
                self.restore_deterministic()
                row += B

            incremental_used = bool(native is not None and native.incremental)
            if native is not None:
                native.release()
            if sv is None:
                return None
            used_f16 = bool(native is not None and native.use_f16)
            f16_somewhere = used_f16
            if world > 1:
                allreduce_scores(sv, group)
                # a NaN from one rank's fp16 overflow is in everybody's table now; whether the
                # run is repeated must not depend on this rank's own flavour
                # (after an agreement every rank knows the job's flavour without asking)
                f16_somewhere = bool(tail_vote[0] and tail_vote[1]) if tail_vote is not None \
                    else any_rank(used_f16, sv.device, group)
            self.last_run_info = {"evaluations": n_evals, "world_size": world, "rank": rank,
                                  "stacked": stack_safe, "native_tail": native is not None,
                                  "incremental_first": incremental_used, "f16": used_f16,
                                  "f16_any_rank": f16_somewhere, "native": native_info,
                                  "native_head": head is not None, "head": head_info}
            return sv

    # ------------------------------------------------------------------ slow path
    def run_module(self, module, samples):
        """Shapley sampling without ``forward_partial`` (shapley_values.py:66-90): every marginal
        evaluation is a forward pass of the whole model over the whole loader with the growing
        set of masked channels zero-filled by a forward hook.  The fill is ``tp_index_fill_dim1``
        and the score update ``tp_loss_delta_scatter``; no per-evaluation host sync."""
        with torch.no_grad():
            self.mask_indices = []
            mask_dev = {"idx": None}

            def forward_hook(mod, _inputs, output):
                mod._tp_prune_dim = output.shape[1]
                if mask_dev["idx"] is not None and mask_dev["idx"].numel() > 0:
                    if output.is_contiguous():
                        _lib.index_fill_dim1_(output, mask_dev["idx"], _ZERO_BITS)
                    else:
                        filled = _lib.index_fill_dim1_(output.contiguous(), mask_dev["idx"],
                                                       _ZERO_BITS)
                        output.copy_(filled)
                return output

            handle = module.register_forward_hook(forward_hook)
            try:
                original_loss = self.run_all_forward()
                n = module._tp_prune_dim
                N = original_loss.shape[0]
                dev = original_loss.device
                sv = torch.zeros((N, n), dtype=torch.float64, device=dev)
                prev_a = torch.empty(N, dtype=torch.float32, device=dev)
                prev_b = torch.empty_like(prev_a)
                for _ in range(samples):
                    self.mask_indices = []
                    prev_in, prev_out = prev_a, prev_b
                    prev_in.copy_(self._per_example(original_loss, N))
                    perm = np.random.permutation(n)  # drawn per sample, as in the reference (:83)
                    perm_dev = torch.from_numpy(perm.astype(np.int64)).to(dev)
                    for k, i in enumerate(perm):
                        self.mask_indices.append(i)
                        mask_dev["idx"] = perm_dev[:k + 1]
                        new_loss = self._per_example(self.run_all_forward(), N)
                        _lib.loss_delta_scatter(new_loss, None, _lib.TP_LOSS_GIVEN, prev_in,
                                                prev_out, perm_dev, k, 1, N, 1, samples, sv, 0)
                        prev_in, prev_out = prev_out, prev_in
                    mask_dev["idx"] = None
            finally:
                handle.remove()
            return self.aggregate_over_samples(sv.cpu().numpy())
