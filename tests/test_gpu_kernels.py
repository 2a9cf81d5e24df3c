"""Parity of every CUDA kernel against the CPU oracle and the reference's golden vectors,
calling through the C-ABI (ctypes) with raw device pointers.  Bars: bit-exact for counts,
indices and copies; <= 1e-5 relative (normalised by the largest score of the tensor) for the
floating-point reductions, with a float64 recomputation as arbiter."""
import os

import numpy as np
import pytest
import torch

from oracle import tp_oracle as oracle
from tests import cases
from torchpruner_b200 import _lib

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
FP_TOL = 1e-5  # north_star: "within 1e-5 relative tolerance for the floating-point scores"
DEV = "cuda:0"


def gold(name):
    return np.load(os.path.join(GOLDEN, name))


def close(got, want, tol=FP_TOL):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    assert got.shape == want.shape
    scale = max(np.abs(want).max() if want.size else 0.0, 1e-30)
    err = np.abs(got - want).max() if want.size else 0.0
    assert err <= tol * scale, f"max abs err {err:.3e} vs scale {scale:.3e}"


@pytest.fixture(params=[1, 2], ids=["direct-load", "tma-tiles"])
def variant(request):
    _lib.set_tuning("reduce_variant", request.param)
    yield request.param
    _lib.set_tuning("reduce_variant", 0)


# ------------------------------------------------------------------ K1-K4
@pytest.mark.parametrize("name", sorted(cases.OP_SHAPES))
def test_row_reductions_match_reference_vectors(name, variant):
    g_ = gold("ops.npz")
    z, g = cases.op_inputs(name)
    zc, gc = z.to(DEV), g.to(DEV)
    close(_lib.taylor_reduce(zc, gc).cpu().numpy(), g_[f"{name}/taylor"])
    close(_lib.taylor_reduce(zc, gc, signed=True).cpu().numpy(), g_[f"{name}/taylor_signed"])
    close(_lib.abs_sum(gc).cpu().numpy(), g_[f"{name}/sensitivity"])
    np.testing.assert_array_equal(_lib.apoz_count(zc).cpu().numpy(), g_[f"{name}/apoz"])
    # and against the float64 arbiter (tighter than the fp32 reference itself can be)
    close(_lib.taylor_reduce(zc, gc, signed=True).cpu().numpy(), oracle.taylor_rows_f64(z, g, True))
    close(_lib.abs_sum(gc).cpu().numpy(), oracle.abs_sum_rows_f64(g))


def test_several_tensors_in_one_launch_match_the_oracle():
    """tp_reduce_rows_multi: the 13 activation shapes of VGG16 (scaled down), rows of 4 ... 1024
    floats, plus tensors it cannot take (49-float rows, rows longer than a tile, an empty one)
    that the wrapper sends through the single-tensor entry points — all against the oracle."""
    shapes = [(6, 64, 32, 32), (6, 64, 32, 32), (6, 128, 16, 16), (5, 256, 8, 8), (7, 512, 4, 4),
              (9, 512, 2, 2), (3, 512, 2, 2), (4, 33, 12), (2, 5, 4096), (11, 7, 4),
              (4, 16, 7, 7), (2, 3, 8192), (0, 8, 4, 4), (3, 2048)]
    gen = torch.Generator().manual_seed(5)
    zs = [torch.relu(torch.randn(s, generator=gen)) for s in shapes]
    gs = [torch.randn(s, generator=gen) / 7 for s in shapes]
    zd, gd = [z.to(DEV) for z in zs], [g.to(DEV) for g in gs]
    prof = _lib.profile_begin(timed=False)
    tay = _lib.reduce_rows_multi("taylor", list(zip(zd, gd)), signed=True)
    launches = {k: v[0] for k, v in prof.records.items()}
    _lib.profile_end()
    # 9 tensors in ONE launch; 49-float rows, 4096- and 8192-float rows (longer than the 1024 the
    # wrapper batches), [3, 2048] viewed as rows of one element and the empty tensor go through
    # the single-tensor entry point
    assert launches == {"reduce_rows_multi[taylor]": 1, "taylor_reduce": 5}, launches
    apoz = _lib.reduce_rows_multi("apoz", zd)
    sens = _lib.reduce_rows_multi("abs_sum", gd)
    unsigned = _lib.reduce_rows_multi("taylor", list(zip(zd, gd)))
    for i, (z, g) in enumerate(zip(zs, gs)):
        if z.numel() == 0:
            assert tay[i].numel() == 0 and apoz[i].shape == (0, 8)
            continue
        close(tay[i].cpu().numpy(), oracle.taylor_rows_f64(z, g, True))
        close(unsigned[i].cpu().numpy(), oracle.taylor_rows(z, g))
        close(sens[i].cpu().numpy(), oracle.abs_sum_rows_f64(g))
        np.testing.assert_array_equal(apoz[i].cpu().numpy(), oracle.apoz_rows(z))
    # more tensors than one descriptor table holds
    many = [torch.relu(torch.randn((3, 4, 8), generator=gen)).to(DEV) for _ in range(70)]
    for got, z in zip(_lib.reduce_rows_multi("apoz", many), many):
        np.testing.assert_array_equal(got.cpu().numpy(), oracle.apoz_rows(z.cpu()))
    # through the C-ABI: limits are errors, nothing is launched
    bad = (_lib.ReduceDesc * 1)(_lib.ReduceDesc(zd[10].data_ptr(), 0, apoz[10].data_ptr(), 64, 49))
    assert _lib.load().tp_reduce_rows_multi(1, bad, 1, 0, None) == -4      # TP_ERR_LIMIT
    assert b"multiple" in _lib.load().tp_last_error()
    assert _lib.load().tp_reduce_rows_multi(1, bad, 65, 0, None) == -4
    assert _lib.load().tp_reduce_rows_multi(7, bad, 1, 0, None) == -1      # unknown op


@pytest.mark.parametrize("name", sorted(cases.WEIGHT_SHAPES))
def test_weight_norm_matches_reference_vectors(name, variant):
    w = cases.weight_input(name)
    close(_lib.weight_abs_sum(w.to(DEV)).cpu().numpy(), gold("ops.npz")[f"{name}/weight_norm"])
    close(_lib.weight_abs_sum(w.to(DEV)).cpu().numpy(), oracle.weight_norm_f64(w))


@pytest.mark.parametrize("shape", [
    (256, 64, 32, 32),    # VGG features.0-3 (config 2)
    (256, 512, 2, 2),     # VGG features.34-40
    (64, 512, 7, 7),      # ResNet-50 layer4: 49-float rows, not 16-byte aligned
    (16, 64, 112, 112),   # ResNet-50 stem: 50 KB rows, longer than a shared-memory tile
    (37, 101, 14, 14),    # ragged everything
    (3, 5, 1031),         # odd long rows
])
def test_row_reductions_at_scale_vs_oracle(shape, variant):
    gen = torch.Generator().manual_seed(sum(shape))
    z = torch.relu(torch.randn(shape, generator=gen))
    g = torch.randn(shape, generator=gen) / shape[0]
    zc, gc = z.to(DEV), g.to(DEV)
    np.testing.assert_array_equal(_lib.apoz_count(zc).cpu().numpy(), oracle.apoz_rows(z))
    close(_lib.taylor_reduce(zc, gc, signed=True).cpu().numpy(), oracle.taylor_rows_f64(z, g, True))
    close(_lib.taylor_reduce(zc, gc).cpu().numpy(), oracle.taylor_rows(z, g))
    close(_lib.abs_sum(gc).cpu().numpy(), oracle.abs_sum_rows_f64(g))


def test_row_reduction_properties_at_full_size(variant):
    """Size-independent properties at BASELINE config 2/4 sizes (no CPU recomputation)."""
    gen = torch.Generator(device=DEV).manual_seed(3)
    for shape in [(256, 64, 32, 32), (256, 256, 56, 56)]:
        z = torch.relu(torch.randn(shape, generator=gen, device=DEV))
        g = torch.randn(shape, generator=gen, device=DEV)
        t = _lib.taylor_reduce(z, g, signed=True)
        # exact linearity in power-of-two scalings, sign symmetry, abs consistency
        assert torch.equal(_lib.taylor_reduce(z, g * 4.0, signed=True), t * 4.0)
        assert torch.equal(_lib.taylor_reduce(z, -g, signed=True), -t)
        assert torch.equal(_lib.taylor_reduce(z, g), t.abs())
        # determinism
        assert torch.equal(_lib.taylor_reduce(z, g, signed=True), t)
        # counts: exact against an independent integer count, and bounded by HW
        cnt = _lib.apoz_count(z)
        assert torch.equal(cnt, (z > 0).flatten(2).sum(-1).float())
        assert float(cnt.max()) <= shape[2] * shape[3]
        # abs-sum of a non-negative tensor equals minus the Taylor sum against all-ones
        ones = torch.ones_like(z)
        assert torch.equal(_lib.abs_sum(z), -_lib.taylor_reduce(ones, z, signed=True))
        close(_lib.abs_sum(g).cpu().numpy(),
              g.abs().double().flatten(2).sum(-1).cpu().numpy())
        del z, g, t, cnt, ones
        torch.cuda.empty_cache()


def test_row_reduction_edge_cases(variant):
    # empty batch, empty rows, unaligned views, non-contiguous input, NaN / inf / -0.0
    empty = torch.zeros((0, 4, 3, 3), device=DEV)
    assert _lib.taylor_reduce(empty, empty).shape == (0, 4)
    norow = torch.zeros((2, 3, 0), device=DEV)
    assert torch.equal(_lib.apoz_count(norow), torch.zeros(2, 3, device=DEV))
    base = torch.randn(4 * 6 * 64 + 3, device=DEV)
    for off in (1, 2, 3):
        z = base[off:off + 4 * 6 * 64].view(4, 6, 8, 8)  # data_ptr not 16-byte aligned
        assert z.data_ptr() % 16 != 0
        close(_lib.abs_sum(z).cpu().numpy(), oracle.abs_sum_rows_f64(z.cpu()))
        np.testing.assert_array_equal(_lib.apoz_count(z).cpu().numpy(), oracle.apoz_rows(z.cpu()))
    nc = torch.randn(5, 8, 6, 4, device=DEV).transpose(2, 3)
    close(_lib.abs_sum(nc).cpu().numpy(), oracle.abs_sum_rows_f64(nc.cpu()))
    special = torch.tensor([[[float("nan"), 1.0, -0.0, 0.0, float("inf"), -float("inf"), 1e-45, -1.0]]],
                           device=DEV)
    np.testing.assert_array_equal(_lib.apoz_count(special).cpu().numpy(), oracle.apoz_rows(special.cpu()))
    assert np.isnan(_lib.abs_sum(special).cpu().numpy()).all()
    with pytest.raises(_lib.TorchPrunerB200Error):
        _lib.taylor_reduce(special, special[:, :, :4])
    with pytest.raises(_lib.TorchPrunerB200Error):
        _lib.apoz_count(special.double())


def test_tma_tuning_space(variant):
    """Every ring depth / tile size the tuner may pick computes the same numbers."""
    if variant != 2:
        pytest.skip("TMA variant only")
    z, g = cases.op_inputs("hw5184")
    big = torch.relu(torch.randn((64, 48, 14, 14), generator=torch.Generator().manual_seed(1)))
    try:
        for stages in (2, 3, 8):
            for tile in (256, 1024, 4096, 8192):
                for ctas in (1, 2, 4):
                    _lib.set_tuning("reduce_stages", stages)
                    _lib.set_tuning("reduce_tile_floats", tile)
                    _lib.set_tuning("reduce_ctas_per_sm", ctas)
                    for t in (z, big):
                        tc = t.to(DEV)
                        np.testing.assert_array_equal(_lib.apoz_count(tc).cpu().numpy(),
                                                      oracle.apoz_rows(t))
                        close(_lib.abs_sum(tc).cpu().numpy(), oracle.abs_sum_rows_f64(t))
                        close(_lib.taylor_reduce(tc, tc, signed=True).cpu().numpy(),
                              oracle.taylor_rows_f64(t, t, True))
    finally:
        _lib.set_tuning("reduce_stages", 4)
        _lib.set_tuning("reduce_tile_floats", 4096)
        _lib.set_tuning("reduce_ctas_per_sm", 2)


# ------------------------------------------------------------------ K5
@pytest.mark.parametrize("shape,k0,P", [((3, 6, 4, 4), 0, 6), ((2, 9, 3, 3), 2, 5), ((4, 16), 5, 11),
                                        ((2, 5, 7), 4, 1), ((1, 64, 2, 2), 10, 54)])
def test_prefix_mask_stack_bit_exact(shape, k0, P):
    gen = torch.Generator().manual_seed(P)
    z = torch.randn(shape, generator=gen)
    z[0, 0] = float("inf")  # a masked inf must become +0.0, not NaN (select, not multiply)
    C = shape[1]
    perm = np.random.RandomState(k0).permutation(C)
    inv = np.empty(C, np.int32)
    inv[perm] = np.arange(C, dtype=np.int32)
    got = _lib.prefix_mask_stack(z.to(DEV), torch.from_numpy(inv).to(DEV), k0, P)
    want = oracle.prefix_mask_stack(z, perm, k0, P)
    got = got.view((P,) + tuple(shape)).cpu().numpy()
    np.testing.assert_array_equal(got.view(np.uint32), want.view(np.uint32))


def test_index_fill_dim1_bit_exact():
    gen = torch.Generator().manual_seed(2)
    for shape, idx in [((2, 7, 3, 3), [0, 6, 3]), ((3, 5), [4]), ((2, 4, 5), [1, 1, 2]), ((2, 3, 2, 2), [])]:
        x = torch.randn(shape, generator=gen)
        got = _lib.index_fill_dim1_(x.to(DEV), torch.tensor(idx, dtype=torch.int64, device=DEV), 0x7FC00000)
        want = oracle.nan_fill(x, idx) if idx else x.numpy()
        np.testing.assert_array_equal(got.cpu().numpy().view(np.uint32), want.view(np.uint32))
        got0 = _lib.index_fill_dim1_(x.to(DEV), torch.tensor(idx, dtype=torch.int64, device=DEV), 0)
        want0 = x.clone()
        if idx:
            want0.index_fill_(1, torch.tensor(idx), 0.0)
        np.testing.assert_array_equal(got0.cpu().numpy().view(np.uint32), want0.numpy().view(np.uint32))


# ------------------------------------------------------------------ K6
@pytest.mark.parametrize("kind", ["given", "ce"])
@pytest.mark.parametrize("P,B,K,C,k0,scatter_from", [(5, 8, 10, 12, 3, 0), (1, 3, 4, 6, 0, 0),
                                                     (7, 33, 10, 64, 10, 2), (1, 5, 3, 4, 2, 1)])
def test_loss_delta_scatter(kind, P, B, K, C, k0, scatter_from):
    rs = np.random.RandomState(P * 100 + B)
    logits = torch.from_numpy(rs.randn(P, B, K).astype(np.float32) * 3)
    target = torch.from_numpy(rs.randint(0, K, size=B).astype(np.int64))
    perm = rs.permutation(C).astype(np.int64)
    losses = oracle.cross_entropy_rows(logits.view(P * B, K), target.repeat(P)).reshape(P, B)
    prev = rs.rand(B).astype(np.float32) + 1.0
    d, row0, samples = B + 4, 3, 5
    sv0 = rs.randn(d, C)
    want_sv = sv0.copy()
    want_prev = oracle.loss_delta_scatter(losses, prev, perm, k0, samples, want_sv, row0, scatter_from)

    sv = torch.from_numpy(sv0.copy()).to(DEV)
    prev_in = torch.from_numpy(prev).to(DEV)
    prev_out = torch.full((B,), -7.0, device=DEV)
    losses_out = torch.empty(P * B, device=DEV)
    src = logits.to(DEV) if kind == "ce" else torch.from_numpy(losses).to(DEV)
    _lib.loss_delta_scatter(src, target.to(DEV), _lib.TP_LOSS_CE if kind == "ce" else _lib.TP_LOSS_GIVEN,
                            prev_in, prev_out, torch.from_numpy(perm).to(DEV), k0, P, B, K, samples,
                            sv, row0, scatter_from, losses_out)
    got_losses = losses_out.cpu().numpy().reshape(P, B)
    if kind == "given":
        np.testing.assert_array_equal(got_losses, losses)
        np.testing.assert_array_equal(sv.cpu().numpy(), want_sv)  # same fp32 sub/div, same fp64 adds
        np.testing.assert_array_equal(prev_out.cpu().numpy(), want_prev)
    else:
        close(got_losses, losses, tol=1e-6)  # per-evaluation losses, well conditioned
        scale = np.abs(losses).max()
        assert np.abs(sv.cpu().numpy() - want_sv).max() <= 4 * np.finfo(np.float32).eps * scale
        close(prev_out.cpu().numpy(), want_prev, tol=1e-6)
    untouched = np.ones((d, C), bool)
    untouched[row0:row0 + B][:, perm[k0 + scatter_from:k0 + P]] = False
    np.testing.assert_array_equal(sv.cpu().numpy()[untouched], sv0[untouched])


# ------------------------------------------------------------------ K7
def test_multi_gather_bit_exact():
    rs = np.random.RandomState(0)
    specs = [  # (shape, axis, dtype)
        ((20, 7, 3, 3), 0, torch.float32), ((20, 7, 3, 3), 1, torch.float32), ((20,), 0, torch.float32),
        ((9, 33), 1, torch.float32), ((9, 33), 0, torch.float16), ((5, 6, 7), 1, torch.float64),
        ((12, 5), 1, torch.uint8), ((64, 64, 3, 3), 0, torch.float32), ((4, 0, 3), 0, torch.float32),
        ((33, 16, 1, 1), 1, torch.bfloat16),
    ]
    srcs, dsts, descs, keeps, wants, off = [], [], [], [], [], 0
    for shape, axis, dtype in specs:
        t = torch.from_numpy(rs.randint(0, 255, size=shape).astype(np.uint8)).to(dtype)
        n = shape[axis]
        drop = rs.choice(n, size=n // 3, replace=False) if n else []
        keep = oracle.keep_indices(n, drop)
        wants.append(t.index_select(axis, torch.from_numpy(keep)))
        src = t.to(DEV)
        new_shape = list(shape)
        new_shape[axis] = len(keep)
        dst = torch.full(new_shape, 1, dtype=dtype, device=DEV)
        outer = int(np.prod(shape[:axis], dtype=np.int64))
        inner = int(np.prod(shape[axis + 1:], dtype=np.int64)) * t.element_size()
        descs.append(_lib.GatherDesc(src.data_ptr(), dst.data_ptr(), outer, n, len(keep), inner, off))
        keeps.append(keep)
        off += len(keep)
        srcs.append(src)
        dsts.append(dst)
    keep_dev = torch.from_numpy(np.concatenate(keeps)).to(DEV)
    _lib.multi_gather(descs, keep_dev, torch.device(DEV))
    for dst, want in zip(dsts, wants):
        assert dst.shape == want.shape
        assert torch.equal(dst.cpu().view(torch.uint8), want.contiguous().view(torch.uint8))
    # more descriptors than fit one launch
    many = [descs[0]] * 70
    _lib.multi_gather(many, keep_dev, torch.device(DEV))
    assert torch.equal(dsts[0].cpu(), wants[0])
    bad = _lib.GatherDesc(srcs[0].data_ptr(), dsts[0].data_ptr(), 1, 20, 14, 252, 10 ** 6)
    with pytest.raises(_lib.TorchPrunerB200Error):
        _lib.multi_gather([bad], keep_dev, torch.device(DEV))


# ------------------------------------------------------------------ K8
@pytest.mark.parametrize("shape", [(2, 12), (2, 3, 2, 2), (2, 70, 5, 5), (3, 2500), (2, 1, 4), (2, 40000)])
def test_nan_any_dim1(shape):
    rs = np.random.RandomState(shape[1])
    x = torch.from_numpy(rs.randn(*shape).astype(np.float32))
    C = shape[1]
    poisoned = sorted(rs.choice(C, size=max(1, C // 5), replace=False).tolist())
    x = torch.from_numpy(oracle.nan_fill(x, poisoned))
    if C > 8:  # +inf and -inf meeting in one channel also sum to NaN; +inf alone does not
        clean = [c for c in range(C) if c not in poisoned]
        x[0, clean[0]] = float("inf")
        x[1, clean[0]] = -float("inf")
        x[0, clean[1]] = float("inf")
    want = oracle.nan_indices(x)
    flags = torch.empty(C, dtype=torch.int32, device=DEV)
    idx = torch.full((C,), -1, dtype=torch.int64, device=DEV)
    cnt = torch.full((1,), -1, dtype=torch.int64, device=DEV)
    for _ in range(2):  # flags are cleared by the call itself
        _lib.nan_any_dim1(x.to(DEV), flags, idx, cnt)
        n = int(cnt.item())
        np.testing.assert_array_equal(idx[:n].cpu().numpy(), want)
    clean_x = torch.from_numpy(rs.randn(*shape).astype(np.float32)).to(DEV)
    _lib.nan_any_dim1(clean_x, flags, idx, cnt)
    assert int(cnt.item()) == 0
