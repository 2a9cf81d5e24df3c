"""The allocation-free NaN-detection entry point against the regular one.

`tp_nan_any_dim1_static` is meant for callers that cannot allocate device memory around the call
(a hook inside another party's forward pass, a process whose memory belongs to a pool, a host
that forbids allocation after start-up): its flag scratch is a fixed array inside the library.
Here it is driven through the C-ABI twice in a row with different inputs — the second call finds
the first call's flags in the shared array and must not be influenced by them — and must list
exactly the channels `tp_nan_any_dim1` lists (index lists are compared exactly, as everywhere in
this suite).  This file sorts last so that a failure here cannot hide another test's result.
"""
import numpy as np
import pytest
import torch

from torchpruner_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _regular(x):
    C = x.shape[1]
    flags = torch.empty(C, dtype=torch.int32, device=DEV)
    idx = torch.full((C,), -1, dtype=torch.int64, device=DEV)
    cnt = torch.full((1,), -1, dtype=torch.int64, device=DEV)
    _lib.nan_any_dim1(x, flags, idx, cnt)
    return idx[: int(cnt.item())].cpu().numpy()


def _make(shape, poisoned, seed):
    x = torch.randn(shape, generator=torch.Generator().manual_seed(seed))
    for c in poisoned:
        x[seed % shape[0], c] = float("nan")
    return x.to(DEV)


def test_static_scratch_variant_matches_regular_entry_point_twice():
    shape = (2, 300, 3, 3)
    first = _make(shape, [0, 17, 123, 299], seed=1)
    second = _make(shape, [5, 200], seed=2)          # disjoint from the first call's channels
    idx = torch.empty(shape[1], dtype=torch.int64, device=DEV)   # outputs are the caller's
    cnt = torch.empty(1, dtype=torch.int64, device=DEV)
    for x in (first, second):
        idx.fill_(-1)
        cnt.fill_(-1)
        _lib.nan_any_dim1_static(x, idx, cnt)
        got = idx[: int(cnt.item())].cpu().numpy()
        np.testing.assert_array_equal(got, _regular(x))
    np.testing.assert_array_equal(got, [5, 200])

    # a wider tensor after a narrow one, no NaN at all, and both infinities meeting in a channel
    wide = torch.zeros((2, 5000), device=DEV)
    wide[0, 4999], wide[1, 4999] = float("inf"), -float("inf")
    wide[0, 7] = float("inf")
    idx = torch.empty(5000, dtype=torch.int64, device=DEV)
    _lib.nan_any_dim1_static(wide, idx, cnt)
    np.testing.assert_array_equal(idx[: int(cnt.item())].cpu().numpy(), _regular(wide))
    assert int(cnt.item()) == 1 and int(idx[0].item()) == 4999
    _lib.nan_any_dim1_static(torch.ones((3, 40, 2), device=DEV), idx, cnt)
    assert int(cnt.item()) == 0


def test_static_scratch_variant_rejects_oversize():
    C = _lib.TP_NAN_STATIC_MAX_C + 1
    x = torch.zeros((1, C), device=DEV)
    idx = torch.full((C,), -3, dtype=torch.int64, device=DEV)
    cnt = torch.full((1,), -3, dtype=torch.int64, device=DEV)
    with pytest.raises(_lib.TorchPrunerB200Error, match="exceeds the built-in scratch"):
        _lib.nan_any_dim1_static(x, idx, cnt)
    assert int(cnt.item()) == -3 and int(idx.min().item()) == -3   # nothing was written
    # the documented maximum itself is accepted
    x = torch.zeros((1, _lib.TP_NAN_STATIC_MAX_C), device=DEV)
    x[0, -1] = float("nan")
    _lib.nan_any_dim1_static(x, idx, cnt)
    assert int(cnt.item()) == 1 and int(idx[0].item()) == _lib.TP_NAN_STATIC_MAX_C - 1


def test_static_scratch_variant_rejects_buffers_that_are_not_on_the_current_gpu():
    """The built-in work area belongs to the GPU that is current during the call.  Output buffers
    somewhere else (here: pinned host memory; on a multi-GPU box, another device) are refused with
    an error instead of being paired with the wrong device's array."""
    x = torch.zeros((2, 8, 3), device=DEV)
    idx = torch.full((8,), -5, dtype=torch.int64, device=DEV)
    host_cnt = torch.full((1,), -5, dtype=torch.int64).pin_memory()
    with pytest.raises(_lib.TorchPrunerB200Error, match="current device"):
        _lib.nan_any_dim1_static(x, idx, host_cnt)
    assert int(host_cnt[0]) == -5 and int(idx.min().item()) == -5
    if torch.cuda.device_count() > 1:
        other = torch.device("cuda:1")
        x1 = torch.zeros((2, 8, 3), device=other)
        x1[1, 3, 0] = float("nan")
        idx1 = torch.empty(8, dtype=torch.int64, device=other)
        cnt1 = torch.empty(1, dtype=torch.int64, device=other)
        _lib.nan_any_dim1_static(x1, idx1, cnt1)    # the wrapper makes x's device current
        assert int(cnt1.item()) == 1 and int(idx1[0].item()) == 3
        with torch.cuda.device(DEV):                # through the C-ABI with the wrong device current
            rc = _lib.load().tp_nan_any_dim1_static(
                x1.data_ptr(), 2, 8, 3, idx1.data_ptr(), cnt1.data_ptr(),
                torch.cuda.current_stream(DEV).cuda_stream)
            assert rc != 0
