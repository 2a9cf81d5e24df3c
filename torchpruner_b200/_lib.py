"""ctypes binding of the C-ABI in include/torchpruner_b200.h.

Everything that reaches a GPU goes through this module: tensors are reduced to raw device
pointers, sizes and the current CUDA stream handle before the call.  There is deliberately no
CPU implementation behind these wrappers — if the shared library is missing, or a tensor does
not live on a CUDA device, the call raises.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libtorchpruner_b200.so")

TP_LOSS_GIVEN = 0
TP_LOSS_CE = 1
TP_GATHER_MAX_DESCS = 32
TP_NAN_STATIC_MAX_C = 65536

_i64 = ctypes.c_int64
_vp = ctypes.c_void_p


class GatherDesc(ctypes.Structure):
    """Mirror of `tp_gather_desc`."""

    _fields_ = [
        ("src", _vp),
        ("dst", _vp),
        ("outer", _i64),
        ("n_src", _i64),
        ("n_keep", _i64),
        ("inner_bytes", _i64),
        ("keep_offset", _i64),
    ]


class ReduceDesc(ctypes.Structure):
    """Mirror of `tp_reduce_desc`."""

    _fields_ = [("a", _vp), ("b", _vp), ("out", _vp), ("rows", _i64), ("len", _i64)]


TP_REDUCE_MAX_DESCS = 64
REDUCE_OPS = {"taylor": 0, "apoz": 1, "abs_sum": 2}

_SIGNATURES = {
    "tp_version": (ctypes.c_int, []),
    "tp_last_error": (ctypes.c_char_p, []),
    "tp_device_check": (ctypes.c_int, [ctypes.c_int]),
    "tp_set_tuning": (ctypes.c_int, [ctypes.c_char_p, _i64]),
    "tp_taylor_reduce": (ctypes.c_int, [_vp, _vp, _i64, _i64, _i64, ctypes.c_int, _vp, _vp]),
    "tp_apoz_count": (ctypes.c_int, [_vp, _i64, _i64, _i64, _vp, _vp]),
    "tp_abs_sum": (ctypes.c_int, [_vp, _i64, _i64, _i64, _vp, _vp]),
    "tp_weight_abs_sum": (ctypes.c_int, [_vp, _i64, _i64, _vp, _vp]),
    "tp_reduce_rows_multi": (ctypes.c_int, [ctypes.c_int, ctypes.POINTER(ReduceDesc), _i64,
                                            ctypes.c_int, _vp]),
    "tp_prefix_mask_stack": (ctypes.c_int, [_vp, _vp, _i64, _i64, _i64, _i64, _i64, _vp, _vp]),
    "tp_index_fill_dim1": (ctypes.c_int, [_vp, _i64, _i64, _i64, _vp, _i64, ctypes.c_uint32, _vp]),
    "tp_loss_delta_scatter": (
        ctypes.c_int,
        [_vp, _vp, ctypes.c_int, _vp, _vp, _vp, _i64, _i64, _i64, _i64, ctypes.c_float, _vp,
         _i64, _i64, _i64, _vp, _vp],
    ),
    "tp_ce_mean_acc": (ctypes.c_int, [_vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp]),
    "tp_multi_gather": (ctypes.c_int, [ctypes.POINTER(GatherDesc), _i64, _vp, _i64, _vp]),
    "tp_nan_any_dim1": (ctypes.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp]),
    "tp_nan_any_dim1_static": (ctypes.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _vp]),
    "tp_conv_tf32x3": (ctypes.c_int, [_vp, _vp, _i64, _i64, _i64, _i64, _vp, _vp, ctypes.c_int, _i64,
                                      _vp, _vp, ctypes.c_int, ctypes.c_float, _vp, _vp, _vp]),
    "tp_conv_tf32x3_flops": (_i64, [_i64, _i64, _i64, _i64, _i64, ctypes.c_int]),
    "tp_tf32_split": (ctypes.c_int, [_vp, _i64, _vp, _vp, _vp]),
    "tp_mask_stack_nhwc": (ctypes.c_int, [_vp, _vp, _i64, _i64, _i64, _i64, _i64, _i64, ctypes.c_int,
                                          _i64, _vp, _vp, _vp]),
    "tp_maxpool2_split_nhwc": (ctypes.c_int, [_vp, _i64, _i64, _i64, _i64, _vp, _vp, _vp]),
    "tp_conv_f16x3": (ctypes.c_int, [_vp, _vp, _i64, _i64, _i64, _i64, _vp, _vp, ctypes.c_int, _i64,
                                     _vp, _vp, ctypes.c_int, ctypes.c_float, ctypes.c_float, _vp, _vp,
                                     _vp]),
    "tp_conv_f16x3_pool2": (ctypes.c_int, [_vp, _vp, _i64, _i64, _i64, _i64, _vp, _vp, _i64,
                                           _vp, _vp, ctypes.c_int, ctypes.c_float, ctypes.c_float,
                                           _vp, _vp, _vp]),
    "tp_conv_pool2_fusable": (ctypes.c_int, [_i64, _i64, _i64, _i64, _i64]),
    "tp_linear_f16x3_pool4": (ctypes.c_int, [_vp, _vp, _i64, _i64, _vp, _vp, _i64, _vp, _vp,
                                             ctypes.c_int, ctypes.c_float, ctypes.c_float, _vp, _vp,
                                             _vp]),
    "tp_f16_split": (ctypes.c_int, [_vp, _i64, ctypes.c_float, _vp, _vp, _vp]),
    "tp_mask_stack_nhwc_f16": (ctypes.c_int, [_vp, _vp, _i64, _i64, _i64, _i64, _i64, _i64,
                                              ctypes.c_int, ctypes.c_float, _i64, _vp, _vp, _vp]),
    "tp_maxpool2_split_nhwc_f16": (ctypes.c_int, [_vp, _i64, _i64, _i64, _i64, ctypes.c_float, _vp,
                                                  _vp, _vp]),
    "tp_prefix_contrib_stack_f16": (ctypes.c_int, [_vp, _vp, _i64, _i64, _vp, _vp, _vp, _i64, _i64,
                                                   _i64, _i64, ctypes.c_int, ctypes.c_int,
                                                   ctypes.c_float, ctypes.c_float, _vp, _vp, _vp]),
    "tp_channel_major": (ctypes.c_int, [_vp, _i64, _i64, _i64, _i64, ctypes.c_int, _vp, _vp]),
    "tp_prefix_contrib_direct": (ctypes.c_int, [_vp, _vp, _vp, _i64, _i64, _vp, _vp, _vp, _i64, _i64,
                                                _i64, _i64, ctypes.c_int, _i64, ctypes.c_int,
                                                ctypes.c_int, ctypes.c_float, ctypes.c_float,
                                                ctypes.c_float, ctypes.c_int, _vp, _vp, _vp]),
    "tp_channel_contrib": (ctypes.c_int, [_vp, _vp, _i64, _i64, _i64, _i64, ctypes.c_int, ctypes.c_int,
                                          _i64, _vp, _vp]),
    "tp_contrib_sum": (ctypes.c_int, [_vp, _i64, _i64, _vp, _vp]),
    "tp_prefix_contrib_stack": (ctypes.c_int, [_vp, _vp, _i64, _i64, _vp, _vp, _vp, _i64, _i64, _i64,
                                               _i64, ctypes.c_int, ctypes.c_int, ctypes.c_float, _vp,
                                               _vp, _vp]),
}

_lib = None
_checked_devices = set()


class TorchPrunerB200Error(RuntimeError):
    pass


def load():
    """Load (once) and return the ctypes handle. Raises if the library has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise TorchPrunerB200Error(
            f"{LIB_PATH} is missing: the CUDA library has not been built. Run "
            "`python -m torchpruner_b200.csrc.build` (needs nvcc). There is no CPU fallback."
        )
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here means header and library disagree
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def exported_symbols():
    return sorted(_SIGNATURES)


def _check(rc, what):
    if rc != 0:
        msg = load().tp_last_error().decode("utf-8", "replace")
        raise TorchPrunerB200Error(f"{what} failed (code {rc}): {msg}")


def _stream(t):
    return torch.cuda.current_stream(t.device).cuda_stream


class LaunchProfile:
    """Optional per-kernel accounting used by bench.py: while active, every C-ABI call is
    bracketed by CUDA events on the stream it is launched on and its algorithmic bytes are
    recorded.  Inactive (the default) it costs one attribute test per call."""

    def __init__(self):
        self.records = {}   # name -> [launches, bytes, [(start_event, stop_event), ...], flops]
        self.timed = True

    def summary(self):
        """{name: {"launches", "bytes", "ms"}}; synchronises the device."""
        torch.cuda.synchronize()
        out = {}
        for name, (n, nbytes, pairs, flops) in self.records.items():
            ms = sum(a.elapsed_time(b) for a, b in pairs)
            out[name] = {"launches": n, "bytes": nbytes, "ms": ms, "flops": flops}
        return out


_profile = None


def profile_begin(timed=True):
    global _profile
    _profile = LaunchProfile()
    _profile.timed = timed
    return _profile


def profile_end():
    global _profile
    p, _profile = _profile, None
    return p


class _launch:
    """with _launch(name, algorithmic_bytes, tensor, kernels): <C-ABI call>"""

    __slots__ = ("name", "nbytes", "t", "kernels", "ev", "flops")

    def __init__(self, name, nbytes, t, kernels=1, flops=0):
        self.name, self.nbytes, self.t, self.kernels, self.flops = name, nbytes, t, kernels, flops

    def __enter__(self):
        self.ev = None
        if _profile is not None and _profile.timed:
            self.ev = torch.cuda.Event(enable_timing=True)
            self.ev.record(torch.cuda.current_stream(self.t.device))

    def __exit__(self, et, ev, tb):
        if _profile is not None and et is None:
            rec = _profile.records.setdefault(self.name, [0, 0, [], 0])
            rec[0] += self.kernels
            rec[1] += int(self.nbytes)
            rec[3] += int(self.flops)
            if self.ev is not None:
                stop = torch.cuda.Event(enable_timing=True)
                stop.record(torch.cuda.current_stream(self.t.device))
                rec[2].append((self.ev, stop))
        return False


def _cuda_f32(t, name):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise TorchPrunerB200Error(
            f"{name} must be a CUDA tensor: torchpruner_b200 has no CPU path (got "
            f"{getattr(t, 'device', type(t))})"
        )
    if t.dtype != torch.float32:
        raise TorchPrunerB200Error(f"{name} must be float32, got {t.dtype}")
    if t.device.index not in _checked_devices:
        _check(load().tp_device_check(t.device.index), "tp_device_check")
        _checked_devices.add(t.device.index)
    return t if t.is_contiguous() else t.contiguous()


def _cuda_half(t, name):
    if not isinstance(t, torch.Tensor) or not t.is_cuda or t.dtype != torch.float16:
        raise TorchPrunerB200Error(f"{name} must be a CUDA float16 tensor")
    return t if t.is_contiguous() else t.contiguous()


def _bchw(t):
    """(B, C, HW) of a [B, C, *spatial] tensor; HW = 1 for 2-D tensors."""
    if t.dim() < 2:
        raise TorchPrunerB200Error(f"expected a tensor with >= 2 dims, got shape {tuple(t.shape)}")
    B, C = t.shape[0], t.shape[1]
    hw = 1
    for s in t.shape[2:]:
        hw *= s
    return B, C, hw


def set_tuning(key, value):
    _check(load().tp_set_tuning(key.encode(), int(value)), "tp_set_tuning")


# ---------------------------------------------------------------- K1-K4 row reductions
def taylor_reduce(z, g, signed=False, out=None):
    z = _cuda_f32(z, "z")
    g = _cuda_f32(g, "g")
    if z.shape != g.shape:
        raise TorchPrunerB200Error(f"activation {tuple(z.shape)} and gradient {tuple(g.shape)} differ")
    B, C, hw = _bchw(z)
    if out is None:
        out = torch.empty((B, C), dtype=torch.float32, device=z.device)
    with torch.cuda.device(z.device), _launch("taylor_reduce", 8 * z.numel() + 4 * B * C, z, 1):
        _check(load().tp_taylor_reduce(z.data_ptr(), g.data_ptr(), B, C, hw, int(bool(signed)),
                                       out.data_ptr(), _stream(z)), "tp_taylor_reduce")
    return out


def apoz_count(z, out=None):
    z = _cuda_f32(z, "z")
    B, C, hw = _bchw(z)
    if out is None:
        out = torch.empty((B, C), dtype=torch.float32, device=z.device)
    with torch.cuda.device(z.device), _launch("apoz_count", 4 * z.numel() + 4 * B * C, z, 1):
        _check(load().tp_apoz_count(z.data_ptr(), B, C, hw, out.data_ptr(), _stream(z)),
               "tp_apoz_count")
    return out


def abs_sum(g, out=None):
    g = _cuda_f32(g, "g")
    B, C, hw = _bchw(g)
    if out is None:
        out = torch.empty((B, C), dtype=torch.float32, device=g.device)
    with torch.cuda.device(g.device), _launch("abs_sum", 4 * g.numel() + 4 * B * C, g, 1):
        _check(load().tp_abs_sum(g.data_ptr(), B, C, hw, out.data_ptr(), _stream(g)), "tp_abs_sum")
    return out


def reduce_rows_multi(op, tensors, signed=False):
    """The per-row reduction ``op`` ("taylor": pairs (z, g); "apoz" / "abs_sum": single tensors,
    viewed as [shape[0] * shape[1], rest] — or [shape[0], rest] with ``rows_dim0`` semantics for
    2-D weights, see ``weight_abs_sum_many``) of several tensors in as few launches as possible:
    everything ``tp_reduce_rows_multi`` accepts goes into launches of up to 64 tensors, the rest
    (rows that are no multiple of 4 floats, or longer than a tile) through the single-tensor
    entry points.  Returns the [B, C] results in order."""
    code = REDUCE_OPS[op]
    outs, batch = [None] * len(tensors), []
    tile = 4096
    for i, item in enumerate(tensors):
        a, b = (item[1], item[0]) if op == "taylor" else (item, None)   # a = gradient, b = z
        a = _cuda_f32(a, "tensor")
        B, C, hw = _bchw(a)
        if b is not None:
            b = _cuda_f32(b, "tensor")
            if b.shape != a.shape:
                raise TorchPrunerB200Error("activation and gradient shapes differ")
        # (rows longer than 1024 floats leave most warps of a CTA without a row in a 16 KB tile and
        # are large enough to run at speed on their own: profiles/r02_tune_multi_single.txt)
        ok = hw % 4 == 0 and 0 < hw <= min(tile, 1024) and a.data_ptr() % 16 == 0 and \
            (b is None or b.data_ptr() % 16 == 0) and B * C > 0
        if not ok:
            outs[i] = taylor_reduce(b, a, signed) if op == "taylor" else \
                (apoz_count(a) if op == "apoz" else abs_sum(a))
            continue
        out = torch.empty((B, C), dtype=torch.float32, device=a.device)
        outs[i] = out
        batch.append((a, b, out, B * C, hw))
    for start in range(0, len(batch), TP_REDUCE_MAX_DESCS):
        part = batch[start:start + TP_REDUCE_MAX_DESCS]
        descs = (ReduceDesc * len(part))(*[
            ReduceDesc(a.data_ptr(), b.data_ptr() if b is not None else 0, out.data_ptr(), rows, hw)
            for a, b, out, rows, hw in part])
        first = part[0][0]
        nbytes = sum((8 if b is not None else 4) * a.numel() + 4 * rows for a, b, _, rows, _ in part)
        with torch.cuda.device(first.device), _launch(f"reduce_rows_multi[{op}]", nbytes, first, 1):
            _check(load().tp_reduce_rows_multi(code, descs, len(part), int(bool(signed)),
                                               _stream(first)), "tp_reduce_rows_multi")
    return outs


def weight_abs_sum(w):
    w = _cuda_f32(w, "weight")
    cout = w.shape[0]
    k = w.numel() // cout if cout else 0
    out = torch.empty((cout,), dtype=torch.float32, device=w.device)
    with torch.cuda.device(w.device), _launch("weight_abs_sum", 4 * w.numel() + 4 * cout, w, 1):
        _check(load().tp_weight_abs_sum(w.data_ptr(), cout, k, out.data_ptr(), _stream(w)),
               "tp_weight_abs_sum")
    return out


# ---------------------------------------------------------------- K5 masks
def prefix_mask_stack(z, inv_perm, k0, P, out=None):
    """out[p] = z with every channel c such that inv_perm[c] <= k0+p zeroed; shape [P*B, C, ...]."""
    z = _cuda_f32(z, "z")
    B, C, hw = _bchw(z)
    if inv_perm.dtype != torch.int32 or not inv_perm.is_cuda or inv_perm.numel() != C:
        raise TorchPrunerB200Error("inv_perm must be a CUDA int32 tensor with one entry per channel")
    if out is None:
        out = torch.empty((P * B,) + tuple(z.shape[1:]), dtype=torch.float32, device=z.device)
    elif out.numel() < P * z.numel() or not out.is_contiguous():
        raise TorchPrunerB200Error("prefix_mask_stack: `out` is too small or not contiguous")
    with torch.cuda.device(z.device), _launch("prefix_mask_stack", 4 * z.numel() * (int(P) + 1), z, 1):
        _check(load().tp_prefix_mask_stack(z.data_ptr(), inv_perm.data_ptr(), int(k0), int(P), B, C,
                                           hw, out.data_ptr(), _stream(z)), "tp_prefix_mask_stack")
    return out


def index_fill_dim1_(x, idx, value_bits):
    """In-place x[:, idx] = value (given as fp32 bit pattern). idx: CUDA int64 tensor in [0, C)."""
    if not x.is_contiguous():
        raise TorchPrunerB200Error("index_fill_dim1_: tensor must be contiguous to be filled in place")
    x = _cuda_f32(x, "x")
    if idx.dtype != torch.int64 or not idx.is_cuda:
        raise TorchPrunerB200Error("index_fill_dim1_: idx must be a CUDA int64 tensor")
    B, C, inner = _bchw(x)
    with torch.cuda.device(x.device), _launch("index_fill_dim1", 4 * B * idx.numel() * inner, x, 1):
        _check(load().tp_index_fill_dim1(x.data_ptr(), B, C, inner, idx.data_ptr(), idx.numel(),
                                         int(value_bits), _stream(x)), "tp_index_fill_dim1")
    return x


# ---------------------------------------------------------------- K6 loss deltas
def loss_delta_scatter(src, target, kind, prev_in, prev_out, perm, k0, P, B, K, sv_samples, sv,
                       row0, scatter_from=0, losses_out=None):
    src = _cuda_f32(src, "src")
    if sv.dtype != torch.float64 or not sv.is_cuda or not sv.is_contiguous() or sv.dim() != 2:
        raise TorchPrunerB200Error("sv must be a contiguous CUDA float64 [d, C] tensor")
    if perm.dtype != torch.int64 or not perm.is_cuda:
        raise TorchPrunerB200Error("perm must be a CUDA int64 tensor")
    if row0 + B > sv.shape[0]:
        raise TorchPrunerB200Error(
            f"score table has {sv.shape[0]} rows but rows [{row0}, {row0 + B}) were addressed")
    tptr = 0
    if kind == TP_LOSS_CE:
        if target.dtype != torch.int64 or not target.is_cuda or target.numel() != B:
            raise TorchPrunerB200Error("cross-entropy targets must be a CUDA int64 [B] tensor")
        tptr = target.data_ptr()
    with torch.cuda.device(src.device), _launch("loss_delta_scatter", 4 * src.numel() + 24 * int(P) * int(B), src, 1):
        _check(load().tp_loss_delta_scatter(
            src.data_ptr(), tptr, int(kind), prev_in.data_ptr(), prev_out.data_ptr(),
            perm.data_ptr(), int(k0), int(P), int(B), int(K), float(sv_samples), sv.data_ptr(),
            sv.shape[1], int(row0), int(scatter_from),
            losses_out.data_ptr() if losses_out is not None else 0, _stream(src)),
            "tp_loss_delta_scatter")


# ---------------------------------------------------------------- K7 gather
def multi_gather(descs, keep, device):
    """descs: list of GatherDesc (host); keep: CUDA int64 tensor holding every keep list."""
    n = len(descs)
    if n == 0:
        return
    arr = (GatherDesc * n)(*descs)
    nbytes = 2 * sum(d.outer * d.n_keep * d.inner_bytes for d in descs) + 8 * keep.numel()
    launches = (n + TP_GATHER_MAX_DESCS - 1) // TP_GATHER_MAX_DESCS
    with torch.cuda.device(device), _launch("multi_gather", nbytes, keep, launches):
        _check(load().tp_multi_gather(arr, n, keep.data_ptr(), keep.numel(),
                                      torch.cuda.current_stream(device).cuda_stream),
               "tp_multi_gather")


# ---------------------------------------------------------------- K8 NaN detection
def nan_any_dim1(x, flags, out_idx, out_count):
    x = _cuda_f32(x, "x")
    B, C, inner = _bchw(x)
    if flags.numel() < C or out_idx.numel() < C:
        raise TorchPrunerB200Error("nan_any_dim1: work buffers smaller than the channel count")
    with torch.cuda.device(x.device), _launch("nan_any_dim1", 4 * x.numel(), x, 2):
        _check(load().tp_nan_any_dim1(x.data_ptr(), B, C, inner, flags.data_ptr(),
                                      out_idx.data_ptr(), out_count.data_ptr(), _stream(x)),
               "tp_nan_any_dim1")


def nan_any_dim1_static(x, out_idx, out_count):
    """`nan_any_dim1` for callers that may not allocate: the flag scratch is a fixed array inside
    the library (at most TP_NAN_STATIC_MAX_C channels), so only the outputs are passed in.  Calls
    on one device share that array — keep them ordered on one stream — and the first call on a
    device does a one-off address lookup, so make it before capturing any CUDA graph."""
    x = _cuda_f32(x, "x")
    B, C, inner = _bchw(x)
    if out_idx.numel() < C:
        raise TorchPrunerB200Error("nan_any_dim1_static: out_idx is smaller than the channel count")
    with torch.cuda.device(x.device), _launch("nan_any_dim1_static", 4 * x.numel(), x, 2):
        _check(load().tp_nan_any_dim1_static(x.data_ptr(), B, C, inner, out_idx.data_ptr(),
                                             out_count.data_ptr(), _stream(x)),
               "tp_nan_any_dim1_static")


# ---------------------------------------------------------------- fp32-accurate layer engine
def tf32_split(x):
    """(hi, lo) with hi = x truncated to TF32 and lo = x - hi (both fp32, x = hi + lo exactly)."""
    x = _cuda_f32(x, "x")
    hi, lo = torch.empty_like(x), torch.empty_like(x)
    with torch.cuda.device(x.device), _launch("tf32_split", 12 * x.numel(), x):
        _check(load().tp_tf32_split(x.data_ptr(), x.numel(), hi.data_ptr(), lo.data_ptr(), _stream(x)),
               "tp_tf32_split")
    return hi, lo


def _act_code(act):
    """(code, slope) of an activation spec: False/None, True/"relu", or ("leaky", slope)."""
    if isinstance(act, tuple):
        return 2, float(act[1])
    return (1 if act else 0), 0.0


def f16_split(x, scale):
    """fp16 pair (hi, lo) of x * scale: hi = rn_fp16(x*scale), lo = rn_fp16(x*scale - hi)."""
    x = _cuda_f32(x, "x")
    hi = torch.empty(x.shape, dtype=torch.float16, device=x.device)
    lo = torch.empty_like(hi)
    with torch.cuda.device(x.device), _launch("f16_split", 8 * x.numel(), x):
        _check(load().tp_f16_split(x.data_ptr(), x.numel(), float(scale), hi.data_ptr(),
                                   lo.data_ptr(), _stream(x)), "tp_f16_split")
    return hi, lo


def conv_pool2_fusable(N, H, W, Cin, Cout):
    """True where tp_conv_f16x3_pool2 can run the 3x3 layer and the 2x2 max pool as one kernel."""
    return bool(load().tp_conv_pool2_fusable(int(N), int(H), int(W), int(Cin), int(Cout)))


def conv_tf32x3(x_hi, x_lo, w_hi, w_lo, taps, scale, shift, relu, split=True, out=None,
                out_scale=1.0, pool=False):
    """y = act(conv(x, w) * scale + shift) on tcgen05 with 3xTF32 products.  x_*: NHWC
    [N,H,W,Cin]; w_*: [Cout, taps*Cin].  Returns (y_hi, y_lo) or (y, None) when split=False.
    pool=True (fp16 flavour, where conv_pool2_fusable): MaxPool2d(2) fused, outputs
    [N,H/2,W/2,Cout]."""
    half = x_hi.dtype == torch.float16   # fp16x3 flavour: operands are fp16 pairs
    if pool and not (half and split and taps == 9):
        raise TorchPrunerB200Error("conv: fused pooling needs the fp16 flavour, 3x3 and split outputs")
    if half:
        x_hi, x_lo = _cuda_half(x_hi, "x_hi"), _cuda_half(x_lo, "x_lo")
        if w_hi.dtype != torch.float16 or w_lo.dtype != torch.float16:
            raise TorchPrunerB200Error("conv: fp16 activations need fp16 weights")
    else:
        x_hi, x_lo = _cuda_f32(x_hi, "x_hi"), _cuda_f32(x_lo, "x_lo")
    N, H, W, Cin = x_hi.shape
    Cout = w_hi.shape[0]
    if w_hi.shape[1] != taps * Cin or w_lo.shape != w_hi.shape or x_lo.shape != x_hi.shape:
        raise TorchPrunerB200Error("conv_tf32x3: operand shapes do not match")
    if out is None:
        odt = torch.float16 if (half and split) else torch.float32
        oshape = (N, H // 2, W // 2, Cout) if pool else (N, H, W, Cout)
        y_hi = torch.empty(oshape, dtype=odt, device=x_hi.device)
        y_lo = torch.empty_like(y_hi) if split else None
    else:
        y_hi, y_lo = out
        if pool and (y_hi.numel() != N * (H // 2) * (W // 2) * Cout or y_lo is None
                     or y_lo.numel() != y_hi.numel()):
            raise TorchPrunerB200Error("conv: pooled outputs must be [N, H/2, W/2, Cout] pairs")
    flops = load().tp_conv_tf32x3_flops(N, H, W, Cin, Cout, int(taps)) if _profile is not None else 0
    esz = 2 if half else 4
    nbytes = esz * (2 * x_hi.numel() + 2 * w_hi.numel()) + \
        y_hi.element_size() * (2 if split else 1) * y_hi.numel()
    name = "conv_f16x3" if half else "conv_tf32x3"
    if _profile is not None and getattr(_profile, "by_shape", False):
        name = f"{name}[{H}x{W} {Cin}->{Cout} t{taps}{' pool' if pool else ''}]"
    with torch.cuda.device(x_hi.device), _launch(name, nbytes, x_hi, 1, flops):
        if pool:
            _check(load().tp_conv_f16x3_pool2(
                x_hi.data_ptr(), x_lo.data_ptr(), N, H, W, Cin, w_hi.data_ptr(), w_lo.data_ptr(),
                Cout, scale.data_ptr(), shift.data_ptr(), *_act_code(relu), float(out_scale),
                y_hi.data_ptr(), y_lo.data_ptr(), _stream(x_hi)), "tp_conv_f16x3_pool2")
        elif half:
            _check(load().tp_conv_f16x3(
                x_hi.data_ptr(), x_lo.data_ptr(), N, H, W, Cin, w_hi.data_ptr(), w_lo.data_ptr(),
                int(taps), Cout, scale.data_ptr(), shift.data_ptr(), *_act_code(relu),
                float(out_scale), y_hi.data_ptr(), y_lo.data_ptr() if y_lo is not None else 0,
                _stream(x_hi)), "tp_conv_f16x3")
        else:
            _check(load().tp_conv_tf32x3(
                x_hi.data_ptr(), x_lo.data_ptr(), N, H, W, Cin, w_hi.data_ptr(), w_lo.data_ptr(),
                int(taps), Cout, scale.data_ptr(), shift.data_ptr(), *_act_code(relu),
                y_hi.data_ptr(), y_lo.data_ptr() if y_lo is not None else 0, _stream(x_hi)),
                "tp_conv_tf32x3")
    return y_hi, y_lo


def linear_pool4(x_hi, x_lo, w_hi, w_lo, scale, shift, relu, out_scale=1.0):
    """Dense form of conv3x3 + affine + activation + MaxPool2d(2) on a 2x2 image, fp16 flavour:
    x_* [N, Cin] (any shape with N leading), w_* [4*Cout, Cin] group-major (one group per output
    pixel), scale/shift [Cout].  Returns the fp16 pair [N, 1, 1, Cout] of the maximum over the four
    groups, times out_scale."""
    x_hi, x_lo = _cuda_half(x_hi, "x_hi"), _cuda_half(x_lo, "x_lo")
    if w_hi.dtype != torch.float16 or w_lo.dtype != torch.float16:
        raise TorchPrunerB200Error("linear_pool4: fp16 activations need fp16 weights")
    N = x_hi.shape[0]
    Cin = x_hi.numel() // max(N, 1)
    Cout = w_hi.shape[0] // 4
    if w_hi.shape[0] != 4 * Cout or w_hi.shape[1] != Cin or w_lo.shape != w_hi.shape or \
            x_lo.shape != x_hi.shape or scale.numel() != Cout or shift.numel() != Cout:
        raise TorchPrunerB200Error("linear_pool4: operand shapes do not match")
    y_hi = torch.empty((N, 1, 1, Cout), dtype=torch.float16, device=x_hi.device)
    y_lo = torch.empty_like(y_hi)
    flops = load().tp_conv_tf32x3_flops(N, 1, 1, Cin, 4 * Cout, 1) if _profile is not None else 0
    nbytes = 2 * (2 * x_hi.numel() + 2 * w_hi.numel()) + 2 * 2 * y_hi.numel()
    name = "conv_f16x3"
    if _profile is not None and getattr(_profile, "by_shape", False):
        name = f"{name}[1x1 {Cin}->{4 * Cout} t1 pool4]"
    with torch.cuda.device(x_hi.device), _launch(name, nbytes, x_hi, 1, flops):
        _check(load().tp_linear_f16x3_pool4(
            x_hi.data_ptr(), x_lo.data_ptr(), N, Cin, w_hi.data_ptr(), w_lo.data_ptr(), Cout,
            scale.data_ptr(), shift.data_ptr(), *_act_code(relu), float(out_scale),
            y_hi.data_ptr(), y_lo.data_ptr(), _stream(x_hi)), "tp_linear_f16x3_pool4")
    return y_hi, y_lo


def mask_stack_nhwc(z, inv_perm, k0, P, pool, out_hi, out_lo, scale=1.0, c_pad=None):
    """NHWC TF32-split stack of P prefix-masked copies of the NCHW activation z (optionally
    2x2 max-pooled first), channels zero-padded to c_pad.  out_*: preallocated, at least
    P*B*H'*W'*c_pad elements."""
    z = _cuda_f32(z, "z")
    if z.dim() != 4:
        raise TorchPrunerB200Error("mask_stack_nhwc expects a [B,C,H,W] activation")
    B, C, H, W = z.shape
    Cp = C if c_pad is None else int(c_pad)
    if Cp < C:
        raise TorchPrunerB200Error("mask_stack_nhwc: c_pad must be at least the channel count")
    Ho, Wo = (H // 2, W // 2) if pool else (H, W)
    need = P * B * Ho * Wo * Cp
    if out_hi.numel() < need or out_lo.numel() < need:
        raise TorchPrunerB200Error("mask_stack_nhwc: output buffers too small")
    esz = out_hi.element_size()
    with torch.cuda.device(z.device), _launch("mask_stack_nhwc", 4 * z.numel() + 2 * esz * need, z):
        if out_hi.dtype == torch.float16:
            _check(load().tp_mask_stack_nhwc_f16(
                z.data_ptr(), inv_perm.data_ptr(), int(k0), int(P), B, C, H, W, int(bool(pool)),
                float(scale), Cp, out_hi.data_ptr(), out_lo.data_ptr(), _stream(z)),
                "tp_mask_stack_nhwc_f16")
        else:
            _check(load().tp_mask_stack_nhwc(
                z.data_ptr(), inv_perm.data_ptr(), int(k0), int(P), B, C, H, W, int(bool(pool)),
                Cp, out_hi.data_ptr(), out_lo.data_ptr(), _stream(z)), "tp_mask_stack_nhwc")
    shape = (P * B, Ho, Wo, Cp)
    return out_hi[:need].view(shape), out_lo[:need].view(shape)


def maxpool2_split_nhwc(x, half=False, scale=1.0):
    x = _cuda_f32(x, "x")
    N, H, W, C = x.shape
    hi = torch.empty((N, H // 2, W // 2, C), dtype=torch.float16 if half else torch.float32,
                     device=x.device)
    lo = torch.empty_like(hi)
    nbytes = 4 * x.numel() + 2 * hi.element_size() * hi.numel()
    with torch.cuda.device(x.device), _launch("maxpool2_split_nhwc", nbytes, x):
        if half:
            _check(load().tp_maxpool2_split_nhwc_f16(x.data_ptr(), N, H, W, C, float(scale),
                                                     hi.data_ptr(), lo.data_ptr(), _stream(x)),
                   "tp_maxpool2_split_nhwc_f16")
        else:
            _check(load().tp_maxpool2_split_nhwc(x.data_ptr(), N, H, W, C, hi.data_ptr(),
                                                 lo.data_ptr(), _stream(x)), "tp_maxpool2_split_nhwc")
    return hi, lo


def ce_mean_acc(logits, target, P, B):
    """(mean cross entropy [P] fp32, correct top-1 count [P] int32) of P stacked batches."""
    logits = _cuda_f32(logits, "logits")
    K = logits.shape[-1]
    if logits.numel() != P * B * K or target.dtype != torch.int64 or target.numel() != B:
        raise TorchPrunerB200Error("ce_mean_acc: shapes do not match")
    mean_loss = torch.empty(P, dtype=torch.float32, device=logits.device)
    correct = torch.empty(P, dtype=torch.int32, device=logits.device)
    with torch.cuda.device(logits.device), _launch("ce_mean_acc", 4 * logits.numel(), logits):
        _check(load().tp_ce_mean_acc(logits.data_ptr(), target.data_ptr(), int(P), int(B), K,
                                     mean_loss.data_ptr(), correct.data_ptr(), _stream(logits)),
               "tp_ce_mean_acc")
    return mean_loss, correct


# ---------------------------------------------------------------- incremental first layer
def channel_contrib(z4, wt, pool, mode, out):
    """out[c] = first layer applied to channel c of z4 alone ([C, pixels*C1] fp32)."""
    z4 = _cuda_f32(z4, "z")
    B, C, H, W = z4.shape
    C1 = wt.shape[2]
    Hp, Wp = (H // 2, W // 2) if pool else (H, W)
    pixels = B * Hp * Wp if mode == 0 else B
    if out.numel() < C * pixels * C1:
        raise TorchPrunerB200Error("channel_contrib: output buffer too small")
    flops = 2 * C * pixels * C1 * wt.shape[1]
    with torch.cuda.device(z4.device), _launch("channel_contrib", 4 * C * pixels * C1, z4, 1, flops):
        _check(load().tp_channel_contrib(z4.data_ptr(), wt.data_ptr(), B, C, H, W, int(bool(pool)),
                                         int(mode), C1, out.data_ptr(), _stream(z4)),
               "tp_channel_contrib")
    return pixels


def channel_major(z4, pool, out=None):
    """[C, B, H', W'] copy of the NCHW activation (2x2 max-pooled first when pool)."""
    z4 = _cuda_f32(z4, "z")
    B, C, H, W = z4.shape
    Hp, Wp = (H // 2, W // 2) if pool else (H, W)
    if out is None or out.numel() < C * B * Hp * Wp:
        out = torch.empty(C * B * Hp * Wp, dtype=torch.float32, device=z4.device)
    with torch.cuda.device(z4.device), _launch("channel_major", 4 * z4.numel() + 4 * C * B * Hp * Wp, z4):
        _check(load().tp_channel_major(z4.data_ptr(), B, C, H, W, int(bool(pool)), out.data_ptr(),
                                       _stream(z4)), "tp_channel_major")
    return out


def prefix_contrib_direct(zc, wt, perm, k0, P, base, scale, shift, B, C, Hp, Wp, mode, pool, relu,
                          sign=-1.0, out_hi=None, out_lo=None, out_scale=1.0):
    """The prefix ladder with contributions recomputed on the fly (see the header).  Returns the
    elements of one mask's output."""
    C1 = wt.shape[2]
    pixels = B * Hp * Wp if mode == 0 else B
    m_out = (pixels // 4 if pool else pixels) * C1
    if out_hi is not None and (out_hi.numel() < P * m_out or out_lo.numel() < P * m_out):
        raise TorchPrunerB200Error("prefix_contrib_direct: output buffers too small")
    if base.numel() < pixels * C1 or base.dtype != torch.float64:
        raise TorchPrunerB200Error("prefix_contrib_direct: base must hold pixels*C1 float64")
    half = out_hi is not None and out_hi.dtype == torch.float16
    nbytes = ((4 if half else 8) * P * m_out if out_hi is not None else 0) + 16 * pixels * C1
    flops = 2 * P * pixels * C1 * wt.shape[1]
    name = "prefix_contrib_direct" if perm is not None else "contrib_sum_direct"
    with torch.cuda.device(zc.device), _launch(name, nbytes, zc, 1, flops):
        _check(load().tp_prefix_contrib_direct(
            zc.data_ptr(), wt.data_ptr(), perm.data_ptr() if perm is not None else 0, int(k0), int(P),
            base.data_ptr(), scale.data_ptr(), shift.data_ptr(), int(B), int(C), int(Hp), int(Wp),
            int(mode), int(C1), int(bool(pool)), *_act_code(relu), float(sign), float(out_scale),
            int(half), out_hi.data_ptr() if out_hi is not None else 0,
            out_lo.data_ptr() if out_lo is not None else 0, _stream(zc)),
            "tp_prefix_contrib_direct")
    return m_out


def contrib_sum(T, C, M1, s_all):
    with torch.cuda.device(T.device), _launch("contrib_sum", 4 * C * M1, T):
        _check(load().tp_contrib_sum(T.data_ptr(), int(C), int(M1), s_all.data_ptr(), _stream(T)),
               "tp_contrib_sum")


def prefix_contrib_stack(T, perm, k0, P, base, scale, shift, pixels, C1, Hp, Wp, pool, relu,
                         out_hi=None, out_lo=None, out_scale=1.0):
    m_out = (pixels // 4 if pool else pixels) * C1
    if out_hi is not None and (out_hi.numel() < P * m_out or out_lo.numel() < P * m_out):
        raise TorchPrunerB200Error("prefix_contrib_stack: output buffers too small")
    half = out_hi is not None and out_hi.dtype == torch.float16
    nbytes = 4 * P * pixels * C1 + ((4 if half else 8) * P * m_out if out_hi is not None else 0)
    with torch.cuda.device(T.device), _launch("prefix_contrib_stack", nbytes, T):
        if half:
            _check(load().tp_prefix_contrib_stack_f16(
                T.data_ptr(), perm.data_ptr(), int(k0), int(P), base.data_ptr(), scale.data_ptr(),
                shift.data_ptr(), int(pixels), int(C1), int(Hp), int(Wp), int(bool(pool)),
                *_act_code(relu), float(out_scale), out_hi.data_ptr(), out_lo.data_ptr(),
                _stream(T)), "tp_prefix_contrib_stack_f16")
        else:
            _check(load().tp_prefix_contrib_stack(
                T.data_ptr(), perm.data_ptr(), int(k0), int(P), base.data_ptr(), scale.data_ptr(),
                shift.data_ptr(), int(pixels), int(C1), int(Hp), int(Wp), int(bool(pool)),
                *_act_code(relu), out_hi.data_ptr() if out_hi is not None else 0,
                out_lo.data_ptr() if out_lo is not None else 0, _stream(T)),
                "tp_prefix_contrib_stack")
    return m_out
