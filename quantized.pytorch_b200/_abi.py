"""ctypes binding of libqb200.so (C ABI: include/qb200.h) + thin tensor-level wrappers.

PyTorch is plumbing here: it owns device memory and streams; every compute call goes through
the C ABI with raw pointers and the *current* CUDA stream of the calling thread (the autograd
engine runs backward on its own thread, so the stream is looked up per call).
There is NO fallback: if the library or a CUDA device is missing the call raises.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libqb200.so")

QB_SCALE_DOUBLE, QB_SCALE_FLOAT = 0, 1

_p = C.c_void_p
_i64, _i32, _u64, _dbl, _flt, _sz = C.c_int64, C.c_int, C.c_uint64, C.c_double, C.c_float, C.c_size_t

# name -> (restype, argtypes); must list every function declared in include/qb200.h (tests check it)
PROTOTYPES = {
    "qb_last_error": (C.c_char_p, []),
    "qb_version": (_i32, []),
    "qb_device_info": (_i32, [_i32, C.POINTER(_i32), C.POINTER(_i32), C.POINTER(_i32), C.POINTER(_sz)]),
    "qb_measure_ws_bytes": (_sz, [_i64]),
    "qb_measure_stats": (_i32, [_p, _i64, _i64, _p, _p, _sz, _p]),
    "qb_relu_measure": (_i32, [_p, _p, _i64, _i64, _p, _p, _sz, _p]),
    "qb_add_relu_measure": (_i32, [_p, _p, _p, _i64, _i64, _p, _p, _sz, _p]),
    "qb_rangebn_apply_relu_measure": (_i32, [_p, _p, _p, _i64, _i64, _i64, _p, _p, _p, _p, _p, _i32, _p, _p, _p, _sz, _p]),
    "qb_relu_bwd_minmax": (_i32, [_p, _p, _p, _i64, _p, _p, _sz, _p]),
    "qb_measure_finalize": (_i32, [_p, _p, _p, _p, _p, _p, _dbl, _i32, _i32, _i32, _p, _p]),
    "qb_fakequant_affine": (_i32, [_p, _p, _p, _i64, _i32, _p, _dbl, _dbl, _i32, _p]),
    "qb_fakequant_affine16": (_i32, [_p, _p, _p, _p, _i64, _i32, _p, _dbl, _dbl, _i32, _p]),
    "qb_fakequant_affine_stochastic": (_i32, [_p, _p, _i64, _i32, _p, _dbl, _dbl, _i32, _p, _u64, _u64, _p, _p]),
    "qb_selftest_div": (_i32, [_p, _flt, _p, _p, _i64, _p]),
    "qb_fakequant_rows": (_i32, [_p, _p, _p, _i64, _i64, _i32, _p, _p, _i32, _p]),
    "qb_minmax_ws_bytes": (_sz, []),
    "qb_minmax": (_i32, [_p, _i64, _p, _p, _sz, _p]),
    "qb_fakequant_grad": (_i32, [_p, _p, _p, _i64, _i32, _p, _dbl, _dbl, _p, _i32, _u64, _u64, _p, _p]),
    "qb_fakequant_grad_fused": (_i32, [_p, _p, _p, _i64, _i32, _p, _p, _i32, _u64, _u64, _p, _p, _sz, _p]),
    "qb_rangebn_ws_bytes": (_sz, [_i64, _i32]),
    "qb_rangebn_stats": (_i32, [_p, _i64, _i64, _i64, _i32, _p, _p, _p, _p, _sz, _p]),
    "qb_rangebn_finalize": (_i32, [_p, _i64, _i32, _dbl, _dbl, _flt, _dbl, _i32, _i32, _p, _p, _p, _p, _p, _p, _p, _p,
                                   _p, _p]),
    "qb_rangebn_apply": (_i32, [_p, _p, _i64, _i64, _i64, _p, _p, _p, _p, _p]),
    "qb_rangebn_bwd_ws_bytes": (_sz, [_i64]),
    "qb_rangebn_bwd_reduce": (_i32, [_p, _p, _i64, _i64, _i64, _p, _p, _p, _sz, _p]),
    "qb_rangebn_bwd_coeffs": (_i32, [_p, _p, _i64, _i32, _dbl, _dbl, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "qb_rangebn_bwd_apply": (_i32, [_p, _p, _i64, _i64, _i64, _i32, _p, _p, _p, _p, _p, _p, _p]),
    "qb_rangebn_nhwc_ws_bytes": (_sz, [_i64, _i32]),
    "qb_rangebn_stats_nhwc": (_i32, [_p, _i64, _i64, _i64, _i32, _p, _p, _p, _p, _p, _p, _i32, _p, _p, _sz, _p]),
    "qb_rangebn_stats_finalize_nhwc": (_i32, [_p, _i64, _i64, _i64, _i32, _p, _p, _p, _p, _p, _p, _i32, _p, _flt, _dbl, _i32,
                                              _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "qb_rangebn_apply_nhwc": (_i32, [_p, _p, _p, _i64, _i64, _i64, _p, _p, _p, _p, _p, _i32, _p]),
    "qb_rangebn_bwd_nhwc_ws_bytes": (_sz, [_i64]),
    "qb_rangebn_bwd_reduce_nhwc": (_i32, [_p, _p, _p, _i64, _i64, _i64, _p, _p, _p, _i32, _p, _i32, _p, _i32, _u64, _u64, _p,
                                          _p, _p, _p, _sz, _p]),
    "qb_rangebn_bwd_reduce_coeffs_nhwc": (_i32, [_p, _p, _p, _i64, _i64, _i64, _p, _p, _p, _i32, _p, _i32, _p, _i32, _u64, _u64, _p,
                                                 _p, _p, _i32, _dbl, _dbl, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "qb_relu_bwd_minmax_bits": (_i32, [_p, _p, _p, _i64, _p, _p, _sz, _p]),
    "qb_rangebn_bwd_apply_fused_nhwc": (_i32, [_p, _p, _i64, _i64, _i64, _i32, _p, _p, _p, _p, _p, _p, _p, _i32, _p, _i32,
                                               _u64, _u64, _p, _p, _p, _p, _sz, _p]),
    "qb_rangebn_bwd_apply_nhwc": (_i32, [_p, _p, _i64, _i64, _i64, _i32, _p, _p, _p, _p, _p, _p, _p]),
    "qb_rowsum_u8": (_i32, [_p, _i64, _i64, _p, _p]),
    "qb_fakequant2": (_i32, [_p, _p, _i64, _i64, _p, _p, _i32, _i32, _i32, _i32, _p, _i32, _u64, _u64, _p, _p]),
    "qb_fakequant2_codes": (_i32, [_p, _p, _p, _p, _i64, _p, _p, _i32, _p]),
    "qb_row_minmax": (_i32, [_p, _i64, _i64, _p, _p, _p]),
    "qb_rangebn2_finalize": (_i32, [_p, _i64, _i32, _dbl, _dbl, _flt, _dbl, _i32, _p, _p, _p, _p, _p, _p, _p]),
    "qb_rangebn2_apply": (_i32, [_p, _p, _i64, _i64, _i64, _i32, _p, _p, _p, _p, _p]),
    "qb_gemm_i8": (_i32, [_p, _p, _p, _i64, _i64, _i64, _i64, _p, _i32, _p, _p, _i32, _p, _p, _p, _i32, _p]),
    "qb_fakequant_rows_conv": (_i32, [_p, _p, _p, _p, _i64, _i64, _i32, _i32, _p, _p, _i32, _p]),
    "qb_allreduce_small_buffer_bytes": (_sz, [_i32, _sz]),
    "qb_allreduce_small": (_i32, [_p, _i64, _i32, _i32, _p, _i32, _i32, _sz, _p]),
    "qb_maxpool2d_nhwc_fwd": (_i32, [_p, _p, _p, _i64, _i64, _i64, _i64, _i32, _i32, _i32, _p]),
    "qb_maxpool2d_nhwc_bwd": (_i32, [_p, _p, _p, _i64, _i64, _i64, _i64, _i32, _i32, _i32, _p]),
    "qb_rangebn_apply_maxpool_nhwc": (_i32, [_p, _p, _p, _i64, _i64, _i64, _i64, _i32, _i32, _i32, _p, _p, _p, _p, _p, _i32, _p]),
    "qb_avgpool2d_nhwc_fwd": (_i32, [_p, _p, _i64, _i64, _i64, _i64, _i32, _p]),
    "qb_avgpool2d_nhwc_bwd": (_i32, [_p, _p, _i64, _i64, _i64, _i64, _i32, _p]),
    "qb_cross_entropy_ws_bytes": (_sz, [_i64]),
    "qb_cross_entropy_fwd": (_i32, [_p, _p, _i64, _i64, _i64, _p, _p, _p, _p, _sz, _p]),
    "qb_cross_entropy_bwd": (_i32, [_p, _p, _p, _p, _p, _p, _i64, _i64, _i64, _p]),
    "qb_maxpool2d_relu_bwd_ws_bytes": (_sz, []),
    "qb_maxpool2d_relu_nhwc_bwd": (_i32, [_p, _p, _p, _p, _i64, _i64, _i64, _i64, _i32, _i32, _i32, _p, _p, _sz, _p]),
    "qb_conv_stem_supported": (_i32, [_i64, _i64, _i64, _i64, _i32, _i32, _i32, _i32]),
    "qb_conv_stem_ws_bytes": (_sz, [_i64]),
    "qb_conv_stem_wgrad_ws_bytes": (_sz, [_i64, _i64, _i32, _i32]),
    "qb_conv_stem_wgrad": (_i32, [_p, _p, _p, _i64, _i64, _i64, _i64, _i64, _i32, _i32, _i32, _i32, _p, _i32, _p, _sz, _p]),
    "qb_conv_stem_fwd": (_i32, [_p, _p, _p, _i64, _i64, _i64, _i64, _i64, _i32, _i32, _i32, _i32, _p, _i32, _p, _p, _i32,
                                _p, _p, _p, _sz, _p]),
    "qb_conv_dw_supported": (_i32, [_i64, _i32, _i32, _i32, _i32]),
    "qb_conv_dw_dgrad": (_i32, [_p, _p, _p, _i64, _i64, _i64, _i64, _i32, _i32, _i32, _i32, _p]),
    "qb_conv_dw_wgrad_ws_bytes": (_sz, [_i64, _i64, _i64, _i64, _i32, _i32]),
    "qb_conv_dw_wgrad": (_i32, [_p, _p, _p, _p, _i64, _i64, _i64, _i64, _i32, _i32, _i32, _i32, _p, _sz, _p]),
    "qb_conv_dw_fwd": (_i32, [_p, _p, _p, _i64, _i64, _i64, _i64, _i32, _i32, _i32, _i32, _p, _i32, _p, _p, _i32, _p, _p, _p,
                              _sz, _p]),
    "qb_bwd_pack": (_i32, [_p, _i64, _p, _p, _p, _i32, _p, _p, _i32, _u64, _u64, _p, _p]),
    "qb_conv_stem_wgrad16_ws_bytes": (_sz, [_i64, _i64, _i64, _i64, _i32, _i32, _i32]),
    "qb_conv_stem_wgrad16": (_i32, [_p, _p, _p, _p, _i64, _i64, _i64, _i64, _i32, _i32, _i32, _p, _i32, _p, _sz, _p]),
    "qb_conv_generic_fwd": (_i32, [_p, _p, _p, _p, _i64, _i64, _i64, _i64, _i64, _p, _p, _p, _p]),
    "qb_conv_generic_dgrad": (_i32, [_p, _p, _p, _i64, _i64, _i64, _i64, _i64, _p, _p, _p, _p]),
    "qb_conv_generic_wgrad": (_i32, [_p, _p, _p, _p, _i64, _i64, _i64, _i64, _i64, _p, _p, _p, _p]),
    "qb_mt_ws_bytes": (_sz, [_i32, _p]),
    "qb_l2_sumsq": (_i32, [_p, _p, _i32, _dbl, _p, _i32, _p, _sz, _p]),
    "qb_l2_sumsq_noalloc_max_blocks": (_i64, []),
    "qb_l2_sumsq_noalloc_setup": (_i32, []),
    "qb_l2_sumsq_noalloc": (_i32, [_p, _p, _i32, _dbl, _p, _i32, _p]),
    "qb_l2_grad": (_i32, [_p, _p, _p, _i32, _dbl, _p, _p]),
    "qb_sgd_step": (_i32, [_p, _p, _p, _p, _i32, _dbl, _dbl, _dbl, _i32, _p]),
    "qb_debug_umma_shift": (_i32, [_p, _p, _p, _i32, _i32, _i32, _p]),
    "qb_debug_umma_rate": (_i32, [_p, _i32, _i32, _i32, _i32, _p]),
    "qb_wt_pack": (_i32, [_p, _p, _i64, _i64, _i32, _p]),
    "qb_conv_dgrad": (_i32, [_p, _p, _p, _p, _i64, _i64, _i64, _i64, _i64, _i32, _i32, _i32, _i32, _p, _i32, _flt, _p]),
    "qb_fakequant_affine_nchw3": (_i32, [_p, _p, _i64, _i64, _i32, _p, _p]),
    "qb_fakequant_affine_rowsum_supported": (_i32, [_i64, _i32]),
    "qb_fakequant_affine_rowsum": (_i32, [_p, _p, _p, _i64, _i64, _i32, _p, _p, _p]),
    "qb_conv_dgrad_acc": (_i32, [_p, _p, _p, _i64, _i64, _i64, _i64, _i64, _i32, _i32, _i32, _i32, _p, _i32, _flt, _p]),
    "qb_conv_wgrad_ws_bytes": (_sz, [_i64, _i64, _i64, _i64, _i64, _i32, _i32, _i32, _i32]),
    "qb_conv_wgrad": (_i32, [_p, _p, _p, _p, _p, _p, _i64, _i64, _i64, _i64, _i64, _i32, _i32, _i32, _i32, _p, _i32, _p, _sz, _p]),
    "qb_conv_i8_fwd": (_i32, [_p, _p, _p, _i64, _i64, _i64, _i64, _i64, _i32, _i32, _i32, _i32, _p, _i32, _p, _p, _i32,
                              _p, _p, _p, _p, _p, _p, _sz, _p]),
}

_lib = None


class QB200Error(RuntimeError):
    pass


def load():
    """Load libqb200.so (built by ``__graft_entry__.build()`` / ``make -C csrc``).  Raises if absent."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise QB200Error("libqb200.so not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                             "(expected at %s); there is no CPU fallback" % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def last_error():
    return load().qb_last_error().decode("utf-8", "replace")


def _check(rc, what):
    if rc != 0:
        raise QB200Error("%s failed (%d): %s" % (what, rc, last_error()))


LAUNCHES = 0          # kernels launched through the C ABI by this process (bench.py reports it)
_profile = None       # list of (name, algorithmic_bytes, start_event, end_event) while profiling

# kernels enqueued per C call
_NLAUNCH = {"qb_rangebn_stats": 2, "qb_rangebn_bwd_reduce": 2, "qb_rangebn_bwd_apply": 2,
            "qb_conv_i8_fwd": 2, "qb_conv_wgrad": 3, "qb_conv_dw_wgrad": 2, "qb_conv_stem_fwd": 2, "qb_conv_stem_wgrad": 2, "qb_rangebn_stats_nhwc": 2, "qb_rangebn_stats_finalize_nhwc": 2, "qb_rangebn_bwd_reduce_nhwc": 2, "qb_rangebn_bwd_reduce_coeffs_nhwc": 2, "qb_rangebn_apply_relu_measure": 1, "qb_rangebn_bwd_apply_nhwc": 2}


FUSE_RANGEBN_COEFFS = os.environ.get("QB_FUSE_RBN_COEFFS", "1") != "0"    # rangebn_backward without a cross-rank hook: sums + coefficients in one call (tests flip it)
_SYNC_EACH = os.environ.get("QB_SYNC_EACH", "") == "1"
_next_ops = 0          # integer operations (2*M*N*K) of the next tensor-core call, set by the conv / GEMM wrappers
PROFILE_OPS = {}       # name -> operations recorded between profile_begin() and profile_end()


def _run(name, nbytes, *args):
    """Call one C-ABI entry point; count its launches; under profile_begin() bracket it with CUDA events on the
    launching stream."""
    global LAUNCHES, _next_ops
    if _profile is not None and _next_ops:
        PROFILE_OPS[name] = PROFILE_OPS.get(name, 0) + _next_ops
    _next_ops = 0
    fn = getattr(load(), name)
    LAUNCHES += _NLAUNCH.get(name, 1)
    if _profile is None:
        rc = fn(*args)
    else:
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        rc = fn(*args)
        b.record()
        _profile.append((name, nbytes, a, b))
    if rc != 0:
        raise QB200Error("%s failed (%d): %s" % (name, rc, last_error()))
    if _SYNC_EACH:                                   # debugging aid (QB_SYNC_EACH=1): name the entry point whose kernel faults
        try:
            torch.cuda.synchronize()
        except Exception as e:  # noqa: BLE001
            raise QB200Error("%s: device fault after this call: %s | args %r" % (name, str(e).splitlines()[0], args[-12:])) from e


def profile_begin():
    global _profile
    _profile = []
    PROFILE_OPS.clear()


def profile_end():
    """-> [(name, algorithmic_bytes, ms)] for every call since profile_begin()."""
    global _profile
    recs, _profile = _profile or [], None
    torch.cuda.synchronize()
    return [(n, nb, a.elapsed_time(b)) for n, nb, a, b in recs]


def device_info(device=0):
    sm, maj, mnr, l2 = _i32(), _i32(), _i32(), _sz()
    _check(load().qb_device_info(device, C.byref(sm), C.byref(maj), C.byref(mnr), C.byref(l2)), "qb_device_info")
    return dict(sm_count=sm.value, cc=(maj.value, mnr.value), l2_bytes=l2.value)


# ----------------------------------------------------------------------------------------------
# tensor-level helpers
# ----------------------------------------------------------------------------------------------
def _stream():
    return torch.cuda.current_stream().cuda_stream


def _ptr(t):
    return None if t is None else t.data_ptr()


def _f32c(t, name):
    if not (t.is_cuda and t.dtype == torch.float32):
        raise QB200Error("%s must be a CUDA float32 tensor (got %s on %s); there is no CPU fallback"
                         % (name, t.dtype, t.device))
    if t.is_contiguous() or (t.dim() == 4 and t.is_contiguous(memory_format=torch.channels_last)):
        return t            # dense in either memory format: elementwise kernels work on the raw storage order
    return t.contiguous()


def is_nhwc(t):
    """True if a 4-D tensor is laid out channels-last (and not equally describable as NCHW-contiguous)."""
    return t.dim() == 4 and t.is_contiguous(memory_format=torch.channels_last) and not t.is_contiguous()


_ws_cache = {}


def workspace(kind, nbytes, device):
    """Zero-initialised scratch, one per (kind, device, stream, thread-of-use).  The kernels leave their
    workspace clean, so it is zeroed only at creation."""
    key = (kind, device.index, torch.cuda.current_stream(device).cuda_stream)
    ws = _ws_cache.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.zeros(max(nbytes, 1), dtype=torch.uint8, device=device)
        _ws_cache[key] = ws
    return ws


def measure_stats(x, stats=None):
    """x [B, ...] -> double[8] stats tensor (see qb_measure_stats)."""
    lib = load()
    x = _f32c(x, "x")
    B = x.size(0)
    inner = x.numel() // max(B, 1)
    if stats is None:
        stats = torch.empty(8, dtype=torch.float64, device=x.device)
    nb = lib.qb_measure_ws_bytes(B)
    ws = workspace("measure", nb, x.device)
    _run("qb_measure_stats", 4 * x.numel(), _ptr(x), B, inner, _ptr(stats), _ptr(ws), ws.numel(), _stream())
    return stats


def relu_measure(x, out=None):
    """out = relu(x) (out=None: in place) and the QuantMeasure statistics (double[8]) of the result, one pass."""
    lib = load()
    x = _f32c(x, "x")
    if out is None:
        out = x
    B = x.size(0)
    inner = x.numel() // max(B, 1)
    stats = torch.empty(8, dtype=torch.float64, device=x.device)
    ws = workspace("measure", lib.qb_measure_ws_bytes(B), x.device)
    _run("qb_relu_measure", 8 * x.numel(), _ptr(x), _ptr(out), B, inner, _ptr(stats), _ptr(ws), ws.numel(), _stream())
    return out, stats


def relu_bwd_minmax(g, out):
    """(dx, range) with dx = where(out <= 0, 0, g) and range = device float[2] {min(dx), max(dx)}; g and out must share
    shape and (dense) memory layout."""
    lib = load()
    dx = torch.empty_like(out)
    rng = torch.empty(2, dtype=torch.float32, device=g.device)
    ws = workspace("minmax", lib.qb_minmax_ws_bytes(), g.device)
    _run("qb_relu_bwd_minmax", 12 * g.numel(), _ptr(g), _ptr(out), _ptr(dx), g.numel(), _ptr(rng), _ptr(ws), ws.numel(),
         _stream())
    return dx, rng


def relu_bwd_minmax_bits(g, bits, want_dx=True):
    """(dx | None, range): ReLU backward from the bit mask of qb_rangebn_apply_relu_measure; ``want_dx=False`` yields only
    the range of the masked gradient (the consumer applies the mask itself)."""
    lib = load()
    dx = torch.empty_like(g) if want_dx else None
    rng = torch.empty(2, dtype=torch.float32, device=g.device)
    ws = workspace("minmax", lib.qb_minmax_ws_bytes(), g.device)
    _run("qb_relu_bwd_minmax_bits", (8 if want_dx else 4) * g.numel() + g.numel() // 8, _ptr(g), _ptr(bits), _ptr(dx), g.numel(),
         _ptr(rng), _ptr(ws), ws.numel(), _stream())
    return dx, rng


def add_relu_measure(x, r, out=None):
    """out = relu(x + r) (out=None: in place in x) and the QuantMeasure statistics (double[8]) of the result, one pass.
    x and r must share shape and memory layout."""
    lib = load()
    x, r = _f32c(x, "x"), _f32c(r, "residual")
    if x.shape != r.shape or x.stride() != r.stride():
        raise QB200Error("add_relu_measure: x and residual differ in shape or layout")
    if out is None:
        out = x
    B = x.size(0)
    inner = x.numel() // max(B, 1)
    stats = torch.empty(8, dtype=torch.float64, device=x.device)
    ws = workspace("measure", lib.qb_measure_ws_bytes(B), x.device)
    _run("qb_add_relu_measure", 12 * x.numel(), _ptr(x), _ptr(r), _ptr(out), B, inner, _ptr(stats), _ptr(ws), ws.numel(),
         _stream())
    return out, stats


def rangebn_apply_relu_measure(x_codes, like, B, Cn, HW, vec, has_gamma, has_beta, q_range, q_bits, residual=None, want_bits=False):
    """-> (out, stats[, bits]): relu(RangeBN-normalise(codes) [+ residual]) and the QuantMeasure statistics of the result in
    one pass (qb_rangebn_apply_relu_measure); ``like`` gives shape / layout of the output (the raw RangeBN input);
    ``want_bits``: also the 1-bit-per-element mask "out is not <= 0" for the ReLU backward."""
    lib = load()
    out = torch.empty_like(like)
    if residual is not None:
        residual = _f32c(residual, "residual")
        if residual.shape != out.shape or residual.stride() != out.stride():
            raise QB200Error("rangebn_apply_relu_measure: residual differs from the output in shape or layout")
    stats = torch.empty(8, dtype=torch.float64, device=like.device)
    ws = workspace("measure", lib.qb_measure_ws_bytes(B), like.device)
    n = B * Cn * HW
    bits = torch.empty(n // 8, dtype=torch.uint8, device=like.device) if (want_bits and (HW * Cn) % 8 == 0) else None
    _run("qb_rangebn_apply_relu_measure", (9 if residual is not None else 5) * n, _ptr(x_codes), _ptr(residual), _ptr(out), B, Cn, HW,
         _ptr(vec[0]), _ptr(vec[2]), _ptr(vec[3]) if has_gamma else None, _ptr(vec[4]) if has_beta else None, _ptr(q_range),
         int(q_bits), _ptr(bits), _ptr(stats), _ptr(ws), ws.numel(), _stream())
    if want_bits:
        return out, stats, bits
    return out, stats


def measure_finalize(stats, running, momentum, training, method, bits, range_out=None):
    """running: tuple of the five [1] buffers (min, max, num_measurements, var, mean)."""
    lib = load()
    rmin, rmax, nmeas, rvar, rmean = running
    if range_out is None:
        range_out = torch.empty(2, dtype=torch.float32, device=rmin.device)
    _run("qb_measure_finalize", 0, _ptr(stats), _ptr(rmin), _ptr(rmax), _ptr(nmeas), _ptr(rvar), _ptr(rmean),
                                   # quantize.py:217 `momentum or self.momentum or n/(n+1)`: a falsy momentum (None, 0, 0.0)
                                   # selects the cumulative average
                                   -1.0 if not momentum else float(momentum), int(training), int(method), int(bits),
                                   _ptr(range_out), _stream())
    return range_out


def fakequant_affine(x, bits, range_dev=None, lo=0.0, hi=0.0, scale_mode=QB_SCALE_DOUBLE, out=None, codes=None,
                     want_y=True, codes16=None):
    lib = load()
    x = _f32c(x, "x")
    if want_y and out is None:
        out = torch.empty_like(x)
    if codes16 is not None:
        nb = 4 + (4 if want_y else 0) + (1 if codes is not None else 0) + 2
        _run("qb_fakequant_affine16", nb * x.numel(), _ptr(x), _ptr(out) if want_y else None, _ptr(codes), _ptr(codes16), x.numel(),
             int(bits), _ptr(range_dev), float(lo), float(hi), int(scale_mode), _stream())
        return out
    _run("qb_fakequant_affine", (4 + (4 if want_y else 0) + (1 if codes is not None else 0)) * x.numel(), _ptr(x), _ptr(out) if want_y else None, _ptr(codes), x.numel(), int(bits),
                                   _ptr(range_dev), float(lo), float(hi), int(scale_mode), _stream())
    return out


def fakequant_affine_nchw3(x, bits, range_dev):
    """x fp32 [B, 3, H, W] NCHW-contiguous -> uint8 codes, logical [B, 3, H, W] stored channels-last (qb_fakequant_affine_nchw3)."""
    B, Cn, H, W = x.shape
    codes = torch.empty((B, Cn, H, W), dtype=torch.uint8, device=x.device, memory_format=torch.channels_last)
    _run("qb_fakequant_affine_nchw3", 5 * x.numel(), _ptr(x), _ptr(codes), B, H * W, int(bits), _ptr(range_dev), _stream())
    return codes


def fakequant_affine_rowsum_supported(cols, bits):
    return bool(load().qb_fakequant_affine_rowsum_supported(int(cols), int(bits)))


def fakequant_affine_rowsum(x, rows, cols, bits, range_dev, codes, codes16=None):
    """Quantizer codes (+ bf16 codes) of x viewed as [rows, cols] and the int32 row sums of the codes, one pass
    (qb_fakequant_affine_rowsum).  Returns the row sums."""
    rowsum = torch.empty(rows, dtype=torch.int32, device=x.device)
    _run("qb_fakequant_affine_rowsum", (5 + (2 if codes16 is not None else 0)) * x.numel(), _ptr(x), _ptr(codes), _ptr(codes16),
         int(rows), int(cols), int(bits), _ptr(range_dev), _ptr(rowsum), _stream())
    return rowsum


def fakequant_affine_stochastic(x, bits, range_dev=None, lo=0.0, hi=0.0, scale_mode=QB_SCALE_DOUBLE, noise=None, seed=0,
                                step_dev=None, out=None):
    """Min-anchored fake-quant with stochastic rounding (qb_fakequant_affine_stochastic)."""
    x = _f32c(x, "x")
    if out is None:
        out = torch.empty_like(x)
    if noise is not None:
        noise = _f32c(noise, "noise")
        if noise.numel() != x.numel():
            raise QB200Error("noise has %d elements, input %d" % (noise.numel(), x.numel()))
    _run("qb_fakequant_affine_stochastic", (8 + (4 if noise is not None else 0)) * x.numel(), _ptr(x), _ptr(out), x.numel(), int(bits),
         _ptr(range_dev), float(lo), float(hi), int(scale_mode), _ptr(noise), int(seed) & (2 ** 64 - 1), 0, _ptr(step_dev), _stream())
    return out


def fakequant_rows(w, bits, row_min=None, row_max=None, compute_range=True, out=None, codes=None, want_y=True):
    lib = load()
    w = _f32c(w, "w")
    rows = w.size(0)
    cols = w.numel() // max(rows, 1)
    if row_min is None:
        row_min = torch.empty(rows, dtype=torch.float32, device=w.device)
        row_max = torch.empty(rows, dtype=torch.float32, device=w.device)
    if want_y and out is None:
        out = torch.empty_like(w)
    _run("qb_fakequant_rows", (4 + (4 if want_y else 0) + (1 if codes is not None else 0)) * w.numel(), _ptr(w), _ptr(out) if want_y else None, _ptr(codes), rows, cols, int(bits),
                                 _ptr(row_min), _ptr(row_max), int(compute_range), _stream())
    return out, row_min, row_max


def minmax(x, range_out=None):
    lib = load()
    x = _f32c(x, "x")
    if range_out is None:
        range_out = torch.empty(2, dtype=torch.float32, device=x.device)
    ws = workspace("minmax", lib.qb_minmax_ws_bytes(), x.device)
    _run("qb_minmax", 4 * x.numel(), _ptr(x), x.numel(), _ptr(range_out), _ptr(ws), ws.numel(), _stream())
    return range_out


def selftest_div(a, b):
    """-> (qb_div(a, b), __fdiv_rn(a, b)) elementwise; see include/qb200.h."""
    a = _f32c(a, "a")
    fast, ieee = torch.empty_like(a), torch.empty_like(a)
    _run("qb_selftest_div", 0, _ptr(a), float(b), _ptr(fast), _ptr(ieee), a.numel(), _stream())
    return fast, ieee


def fakequant_grad(g, bits, range_dev=None, lo=0.0, hi=0.0, noise=None, stochastic=True, seed=0, offset=0,
                   step_dev=None, out=None, codes=None, want_y=True):
    lib = load()
    g = _f32c(g, "g")
    if noise is not None:
        noise = _f32c(noise, "noise")
        if noise.numel() != g.numel():
            raise QB200Error("noise has %d elements, gradient %d" % (noise.numel(), g.numel()))
    if want_y and out is None:
        out = torch.empty_like(g)
    _run("qb_fakequant_grad", (4 + (4 if want_y else 0) + (1 if codes is not None else 0) + (4 if noise is not None else 0)) * g.numel(), _ptr(g), _ptr(out) if want_y else None, _ptr(codes), g.numel(), int(bits),
                                 _ptr(range_dev), float(lo), float(hi), _ptr(noise), int(stochastic),
                                 int(seed) & (2 ** 64 - 1), int(offset), _ptr(step_dev), _stream())
    return out


def fakequant_grad_fused(g, bits, noise=None, stochastic=True, seed=0, offset=0, step_dev=None, out=None, codes=None,
                         range_out=None, want_y=True):
    """UniformQuantizeGrad.backward in one cooperative launch.  Returns (out, range_out)."""
    lib = load()
    g = _f32c(g, "g")
    if noise is not None:
        noise = _f32c(noise, "noise")
        if noise.numel() != g.numel():
            raise QB200Error("noise has %d elements, gradient %d" % (noise.numel(), g.numel()))
    if want_y and out is None:
        out = torch.empty_like(g)
    if range_out is None:
        range_out = torch.empty(2, dtype=torch.float32, device=g.device)
    ws = workspace("minmax", lib.qb_minmax_ws_bytes(), g.device)
    _run("qb_fakequant_grad_fused", (8 + (4 if want_y else 0) + (1 if codes is not None else 0) + (4 if noise is not None else 0)) * g.numel(), _ptr(g), _ptr(out) if want_y else None, _ptr(codes), g.numel(), int(bits),
                                       _ptr(range_out), _ptr(noise), int(stochastic), int(seed) & (2 ** 64 - 1),
                                       int(offset), _ptr(step_dev), _ptr(ws), ws.numel(), _stream())
    return out, range_out


# ----------------------------------------------------------------------------------------------
# RangeBN
# ----------------------------------------------------------------------------------------------
def rangebn_stats(x, B, Cn, HW, num_chunks, want_arg=True, nhwc=False, q_range=None, q_bits=8, x_codes_out=None):
    """x: RangeBN input (already fake-quantized, or RAW with ``q_range``: channels-last only, the quantizer then runs
    inside the kernel).  Returns (red double[3*C], amax int32 | None, amin | None).  For channels-last, amax / amin
    are [2, C*K]: row 0 in the reference's [c][k] order, row 1 transposed to [k][c] for the fused backward.
    ``x_codes_out`` (uint8, x.numel() elements): also receives x_hat as integer codes (see qb200.h)."""
    lib = load()
    dev = x.device
    red = torch.empty(3 * Cn, dtype=torch.float64, device=dev)
    sfx = "_nhwc" if nhwc else ""
    ws = workspace("rangebn" + sfx, (lib.qb_rangebn_nhwc_ws_bytes if nhwc else lib.qb_rangebn_ws_bytes)(Cn, num_chunks), dev)
    if nhwc:
        amax = torch.empty(2, Cn * num_chunks, dtype=torch.int32, device=dev) if want_arg else None
        amin = torch.empty(2, Cn * num_chunks, dtype=torch.int32, device=dev) if want_arg else None
        _run("qb_rangebn_stats_nhwc", (4 + (1 if x_codes_out is not None else 0)) * B * Cn * HW, _ptr(x), B, Cn, HW, num_chunks,
             _ptr(amax[0]) if want_arg else None, _ptr(amin[0]) if want_arg else None,
             _ptr(amax[1]) if want_arg else None, _ptr(amin[1]) if want_arg else None, _ptr(red),
             _ptr(q_range), int(q_bits), _ptr(x_codes_out), _ptr(ws), ws.numel(), _stream())
        return red, amax, amin
    if x_codes_out is not None:
        raise QB200Error("rangebn_stats: x_codes_out is a channels-last feature")
    if q_range is not None:
        raise QB200Error("rangebn_stats: on-the-fly input quantization is a channels-last feature")
    amax = torch.empty(Cn * num_chunks, dtype=torch.int32, device=dev) if want_arg else None
    amin = torch.empty(Cn * num_chunks, dtype=torch.int32, device=dev) if want_arg else None
    _run("qb_rangebn_stats", 4 * B * Cn * HW, _ptr(x), B, Cn, HW, num_chunks, _ptr(amax), _ptr(amin), _ptr(red), _ptr(ws),
         ws.numel(), _stream())
    return red, amax, amin


def rangebn_stats_finalize_nhwc(x, B, Cn, HW, num_chunks, eps, momentum, bits, gamma, beta, running_mean, running_var,
                                want_arg=True, q_range=None, q_bits=8, x_codes_out=None):
    """rangebn_stats(nhwc=True) + rangebn_finalize(training=True) in one call (qb_rangebn_stats_finalize_nhwc).
    Returns (vec float[5, C], amax, amin)."""
    lib = load()
    dev = x.device
    red = torch.empty(3 * Cn, dtype=torch.float64, device=dev)
    vec = torch.empty(5, Cn, dtype=torch.float32, device=dev)
    ws = workspace("rangebn_nhwc", lib.qb_rangebn_nhwc_ws_bytes(Cn, num_chunks), dev)
    amax = torch.empty(2, Cn * num_chunks, dtype=torch.int32, device=dev) if want_arg else None
    amin = torch.empty(2, Cn * num_chunks, dtype=torch.int32, device=dev) if want_arg else None
    _run("qb_rangebn_stats_finalize_nhwc", (4 + (1 if x_codes_out is not None else 0)) * B * Cn * HW, _ptr(x), B, Cn, HW,
         num_chunks, _ptr(amax[0]) if want_arg else None, _ptr(amin[0]) if want_arg else None,
         _ptr(amax[1]) if want_arg else None, _ptr(amin[1]) if want_arg else None, _ptr(red), _ptr(q_range), int(q_bits),
         _ptr(x_codes_out), float(eps), float(momentum), int(bits), _ptr(gamma), _ptr(beta), _ptr(running_mean),
         _ptr(running_var), _ptr(vec[0]), _ptr(vec[1]), _ptr(vec[2]), _ptr(vec[3]) if gamma is not None else None,
         _ptr(vec[4]) if beta is not None else None, _ptr(ws), ws.numel(), _stream())
    return vec, amax, amin


def rangebn_finalize(red, Cn, num_chunks, count, chunk_len, eps, momentum, training, bits, gamma, beta, running_mean,
                     running_var):
    """Returns vec float[5, C]: rows mean, scale_raw, qscale, qgamma, qbeta."""
    lib = load()
    vec = torch.empty(5, Cn, dtype=torch.float32, device=running_mean.device)
    _run("qb_rangebn_finalize", 0, _ptr(red), Cn, int(num_chunks), float(count), float(chunk_len), float(eps),
                                   float(momentum), int(training), int(bits), _ptr(gamma), _ptr(beta),
                                   _ptr(running_mean), _ptr(running_var), _ptr(vec[0]), _ptr(vec[1]), _ptr(vec[2]),
                                   _ptr(vec[3]) if gamma is not None else None,
                                   _ptr(vec[4]) if beta is not None else None, _stream())
    return vec


def rangebn_apply(x, B, Cn, HW, vec, has_gamma, has_beta, out=None, nhwc=False, q_range=None, q_bits=8, x_codes=None):
    """``x_codes`` (channels-last): x_hat as the uint8 codes rangebn_stats wrote; x is then used for its shape only."""
    lib = load()
    if out is None:
        out = torch.empty_like(x)
    args = [_ptr(x)] + ([_ptr(x_codes)] if nhwc else []) + [
        _ptr(out), B, Cn, HW, _ptr(vec[0]), _ptr(vec[2]), _ptr(vec[3]) if has_gamma else None,
        _ptr(vec[4]) if has_beta else None]
    if nhwc:
        args += [_ptr(q_range), int(q_bits)]
    elif q_range is not None or x_codes is not None:
        raise QB200Error("rangebn_apply: on-the-fly input quantization / x_codes are channels-last features")
    _run("qb_rangebn_apply" + ("_nhwc" if nhwc else ""), (5 if x_codes is not None else 8) * B * Cn * HW, *args, _stream())
    return out


def rangebn_backward(gq, xq, B, Cn, HW, num_chunks, count, chunk_len, training, vec, has_gamma, amax, amin,
                     reduce_hook=None, chunks_global=None, nhwc=False, q_range=None, q_bits=8, grad_quant=None,
                     want_dx_range=False, x_codes=None, relu_bits=None):
    """Returns (dx, dgamma, dbeta[, dx_range]).  ``reduce_hook(sums)`` may all-reduce the double[2*C] sums in place
    (batch-sharded training); ``num_chunks`` is this rank's chunk count (layout of amax/amin),
    ``chunks_global`` the count the chunk-mean was taken over.

    Channels-last fusion: ``q_range`` - xq is the RAW RangeBN input, quantized on the fly; ``grad_quant`` =
    dict(range=float[2] device tensor, bits, noise, stochastic, seed, step_dev) - gq is the RAW output gradient,
    quantized on the fly in both passes; ``want_dx_range`` - also return min/max of dx (float[2]); ``x_codes`` - x_hat
    as the uint8 codes the forward statistics pass wrote (xq may then be None)."""
    lib = load()
    dev = gq.device
    if chunks_global is None:
        chunks_global = num_chunks
    sums = torch.empty(2 * Cn, dtype=torch.float64, device=dev)
    sfx = "_nhwc" if nhwc else ""
    ws = workspace("rangebn_bwd" + sfx, (lib.qb_rangebn_bwd_nhwc_ws_bytes if nhwc else lib.qb_rangebn_bwd_ws_bytes)(Cn), dev)
    fused = nhwc and Cn % 8 == 0 and gq.numel() < 2 ** 32
    if (grad_quant is not None or want_dx_range) and not fused:
        raise QB200Error("rangebn_backward: gradient-quantizer / dx-range fusion needs channels-last with C % 8 == 0")
    gqa = [None, 0, None, 0, 0, 0, None]
    if grad_quant is not None:
        gqa = [_ptr(grad_quant["range"]), int(grad_quant["bits"]), _ptr(grad_quant.get("noise")),
               int(bool(grad_quant.get("stochastic", True))), int(grad_quant.get("seed", 0)) & (2 ** 64 - 1), 0,
               _ptr(grad_quant.get("step_dev"))]
    codes = None
    if fused and grad_quant is not None and int(grad_quant["bits"]) <= 8:
        codes = torch.empty(gq.numel(), dtype=torch.uint8, device=dev)   # Q(g) handed from the reduce to the apply pass
    coef = torch.empty(4, Cn, dtype=torch.float32, device=dev)     # dgamma, dbeta, a, e
    coef_args = (int(chunks_global), float(count), float(chunk_len), int(training), _ptr(vec[0]), _ptr(vec[1]), _ptr(vec[2]),
                 _ptr(vec[3]) if has_gamma else None, _ptr(coef[0]), _ptr(coef[1]), _ptr(coef[2]), _ptr(coef[3]))
    have_coef = False
    if nhwc:
        nb = 4 + (1 if x_codes is not None else 4) + (1 if codes is not None else 0)
        red_args = (_ptr(gq), _ptr(xq), _ptr(x_codes), B, Cn, HW, _ptr(sums[:Cn]), _ptr(sums[Cn:]), _ptr(q_range), int(q_bits),
                    *gqa, _ptr(codes), _ptr(relu_bits))
        if reduce_hook is None and FUSE_RANGEBN_COEFFS:      # nothing to exchange between the sums and the coefficients
            _run("qb_rangebn_bwd_reduce_coeffs_nhwc", nb * B * Cn * HW, *red_args, *coef_args, _ptr(ws), ws.numel(), _stream())
            have_coef = True
        else:
            _run("qb_rangebn_bwd_reduce_nhwc", nb * B * Cn * HW, *red_args, _ptr(ws), ws.numel(), _stream())
    else:
        if q_range is not None or x_codes is not None:
            raise QB200Error("rangebn_backward: on-the-fly input quantization / x_codes are channels-last features")
        _run("qb_rangebn_bwd_reduce", 8 * B * Cn * HW, _ptr(gq), _ptr(xq), B, Cn, HW, _ptr(sums[:Cn]), _ptr(sums[Cn:]),
             _ptr(ws), ws.numel(), _stream())
    if reduce_hook is not None:
        reduce_hook(sums)
    if not have_coef:
        _run("qb_rangebn_bwd_coeffs", 0, _ptr(sums[:Cn]), _ptr(sums[Cn:]), Cn, *coef_args, _stream())
    dx = torch.empty_like(gq)
    dx_range = None
    if fused:
        mm = None
        if want_dx_range:
            dx_range = torch.empty(2, dtype=torch.float32, device=dev)
            mm = workspace("minmax", lib.qb_minmax_ws_bytes(), dev)
        _run("qb_rangebn_bwd_apply_fused_nhwc", (5 if codes is not None else 8) * B * Cn * HW, _ptr(gq), _ptr(dx), B, Cn, HW,
             int(num_chunks) if training else 0, _ptr(vec[2]), _ptr(vec[3]) if has_gamma else None, _ptr(coef[2]),
             _ptr(coef[3]) if training else None, _ptr(amax[1]) if training else None,
             _ptr(amin[1]) if training else None, *gqa, _ptr(codes), _ptr(dx_range), _ptr(mm),
             mm.numel() if mm is not None else 0, _stream())
    else:
        a0 = amax[0] if (nhwc and amax is not None) else amax
        i0 = amin[0] if (nhwc and amin is not None) else amin
        _run("qb_rangebn_bwd_apply" + sfx, 8 * B * Cn * HW, _ptr(gq), _ptr(dx), B, Cn, HW, int(num_chunks) if training else 0,
             _ptr(vec[2]), _ptr(vec[3]) if has_gamma else None, _ptr(coef[2]), _ptr(coef[3]) if training else None,
             _ptr(a0) if training else None, _ptr(i0) if training else None, _stream())
    if want_dx_range:
        return dx, coef[0], coef[1], dx_range
    return dx, coef[0], coef[1]

def rowsum_u8(codes, rows, cols, out=None):
    if out is None:
        out = torch.empty(rows, dtype=torch.int32, device=codes.device)
    _run("qb_rowsum_u8", rows * cols, _ptr(codes), rows, cols, _ptr(out), _stream())
    return out


def gemm_i8(a_codes, b_codes, M, N, K, a_range, a_bits, b_row_min, b_row_max, b_bits, a_rowsum, b_rowsum, bias=None,
            out=None, b_offset=0):
    """out[M,N] fp32 = dequantised (A_hat @ B_hat^T + bias) from uint8 codes (see qb_gemm_i8)."""
    global _next_ops
    if out is None:
        out = torch.empty(M, N, dtype=torch.float32, device=a_codes.device)
    _next_ops = 2 * M * N * K
    _run("qb_gemm_i8", M * K + N * K + 4 * M * N, _ptr(a_codes), _ptr(b_codes), _ptr(out), M, N, K, out.stride(0),
         _ptr(a_range), int(a_bits), _ptr(b_row_min), _ptr(b_row_max), int(b_bits), _ptr(a_rowsum), _ptr(b_rowsum),
         _ptr(bias), int(b_offset), _stream())
    return out


def fakequant_rows_conv(w, bits, row_min, row_max, compute_range=True, want_y=True):
    """w [Cout, Cin, kh, kw] contiguous -> (wq | None, codes [Cout, kh*kw, Cin] uint8, tapsum [Cout, kh*kw] int32)."""
    w = w if w.is_contiguous() else w.contiguous()
    Cout, Cin = w.shape[0], w.shape[1]
    taps = w.numel() // (Cout * Cin)
    wq = torch.empty_like(w) if want_y else None
    codes = torch.empty(Cout, taps, Cin, dtype=torch.uint8, device=w.device)
    tapsum = torch.empty(Cout, taps, dtype=torch.int32, device=w.device)
    _run("qb_fakequant_rows_conv", (4 + (4 if want_y else 0) + 1) * w.numel(), _ptr(w), _ptr(wq), _ptr(codes), _ptr(tapsum),
         Cout, Cin, taps, int(bits), _ptr(row_min), _ptr(row_max), int(compute_range), _stream())
    return wq, codes, tapsum


def conv_i8_fwd(x_codes, w_codes, B, H, W, Cin, Cout, kh, kw, stride, pad, a_range, a_bits, row_min, row_max, w_bits, pixsum,
                tapsum, bias=None, out_stats=None):
    """-> fp32 output, logical shape [B, Cout, Ho, Wo] stored channels-last.  ``out_stats``: double[8] tensor that
    receives the QuantMeasure statistics of the output (fused range reduction for the next layer)."""
    Ho, Wo = (H + 2 * pad - kh) // stride + 1, (W + 2 * pad - kw) // stride + 1
    out = torch.empty((B, Cout, Ho, Wo), dtype=torch.float32, device=x_codes.device, memory_format=torch.channels_last)
    boxsum = torch.empty(B * Ho * Wo, dtype=torch.int32, device=x_codes.device)
    ws = workspace("measure", load().qb_measure_ws_bytes(B), x_codes.device) if out_stats is not None else None
    global _next_ops
    _next_ops = 2 * B * Ho * Wo * Cout * kh * kw * Cin
    _run("qb_conv_i8_fwd", B * H * W * Cin + Cout * kh * kw * Cin + 4 * B * Ho * Wo * Cout, _ptr(x_codes), _ptr(w_codes),
         _ptr(out), B, H, W, Cin, Cout, kh, kw, stride, pad, _ptr(a_range), int(a_bits), _ptr(row_min), _ptr(row_max),
         int(w_bits), _ptr(pixsum), _ptr(boxsum), _ptr(tapsum), _ptr(bias), _ptr(out_stats), _ptr(ws),
         ws.numel() if ws is not None else 0, _stream())
    return out


def conv_dw_supported(C, kh, kw, stride, pad):
    return bool(load().qb_conv_dw_supported(int(C), int(kh), int(kw), int(stride), int(pad)))


def conv_dw_fwd(x_codes, w_codes, B, H, W, C, kh, kw, stride, pad, a_range, a_bits, row_min, row_max, w_bits, bias=None,
                out_stats=None):
    """Depthwise 3x3 convolution on the integer codes -> fp32 [B, C, Ho, Wo] stored channels-last (qb_conv_dw_fwd)."""
    Ho, Wo = (H + 2 * pad - kh) // stride + 1, (W + 2 * pad - kw) // stride + 1
    out = torch.empty((B, C, max(Ho, 0), max(Wo, 0)), dtype=torch.float32, device=x_codes.device,
                      memory_format=torch.channels_last)
    ws = workspace("measure", load().qb_measure_ws_bytes(B), x_codes.device) if out_stats is not None else None
    _run("qb_conv_dw_fwd", B * H * W * C + 9 * C + 4 * B * Ho * Wo * C, _ptr(x_codes), _ptr(w_codes), _ptr(out), B, H, W, C,
         kh, kw, stride, pad, _ptr(a_range), int(a_bits), _ptr(row_min), _ptr(row_max), int(w_bits), _ptr(bias),
         _ptr(out_stats), _ptr(ws), ws.numel() if ws is not None else 0, _stream())
    return out


def conv_dw_bwd_supported(g, x, w, stride, pad):
    """Channels-last fp32 tensors of a depthwise 3x3 convolution the native backward kernels take."""
    C = x.shape[1]
    return (g.dim() == 4 and x.dim() == 4 and tuple(w.shape) == (C, 1, 3, 3) and C % 4 == 0 and 256 % (C // 4) == 0
            and conv_dw_supported(C, 3, 3, stride, pad) and x.numel() < 2 ** 32 and g.dtype == torch.float32
            and x.dtype == torch.float32 and w.dtype == torch.float32 and is_nhwc(x) and x.data_ptr() % 16 == 0)


def conv_dw_dgrad(grad_q, w_hat, x_shape, stride, pad):
    """Input gradient of the depthwise 3x3 convolution: grad_q [B,C,Ho,Wo] (channels-last), w_hat [C,1,3,3] -> dx."""
    B, C, H, W = x_shape
    g = grad_q if is_nhwc(grad_q) else grad_q.contiguous(memory_format=torch.channels_last)
    w = w_hat if w_hat.is_contiguous() else w_hat.contiguous()
    dx = torch.empty((B, C, H, W), dtype=torch.float32, device=g.device, memory_format=torch.channels_last)
    _run("qb_conv_dw_dgrad", 4 * g.numel() + 4 * dx.numel(), _ptr(g), _ptr(w), _ptr(dx), B, H, W, C, 3, 3, int(stride), int(pad),
         _stream())
    return dx


def conv_dw_wgrad(grad_out, x_hat, stride, pad, want_bias):
    """Weight (and bias) gradient of the depthwise 3x3 convolution from channels-last grad_out / x_hat -> (dw, db|None)."""
    B, C, H, W = x_hat.shape
    g = grad_out if is_nhwc(grad_out) else grad_out.contiguous(memory_format=torch.channels_last)
    dw = torch.empty((C, 1, 3, 3), dtype=torch.float32, device=g.device)
    db = torch.empty(C, dtype=torch.float32, device=g.device) if want_bias else None
    nbytes = load().qb_conv_dw_wgrad_ws_bytes(B, H, W, C, int(stride), int(pad))
    ws = workspace("conv_dw_wgrad", nbytes, g.device)
    _run("qb_conv_dw_wgrad", 4 * g.numel() + 4 * x_hat.numel(), _ptr(g), _ptr(x_hat), _ptr(dw), _ptr(db), B, H, W, C, 3, 3,
         int(stride), int(pad), _ptr(ws), ws.numel(), _stream())
    return dw, db


def bwd_pack(g, range_dev=None, bits=8, noise=None, stochastic=True, seed=0, step_dev=None, want_codes=True, want_split=True):
    """One pass over the fp32 gradient g -> (gq16 | None, g_hi | None, g_lo | None): bf16 tensors shaped / laid out like g
    (qb_bwd_pack).  gq16 = integer codes (q - zp) of the gradient quantizer over ``range_dev``; hi / lo = split of raw g."""
    g = _f32c(g, "grad")
    gq = torch.empty_like(g, dtype=torch.bfloat16) if want_codes else None
    hi = torch.empty_like(g, dtype=torch.bfloat16) if want_split else None
    lo = torch.empty_like(g, dtype=torch.bfloat16) if want_split else None
    if noise is not None:
        noise = _f32c(noise, "noise")
        if noise.numel() != g.numel():
            raise QB200Error("noise has %d elements, gradient %d" % (noise.numel(), g.numel()))
    nb = 4 + (2 if want_codes else 0) + (4 if want_split else 0) + (4 if noise is not None else 0)
    _run("qb_bwd_pack", nb * g.numel(), _ptr(g), g.numel(), _ptr(gq), _ptr(hi), _ptr(lo), int(bits), _ptr(range_dev), _ptr(noise),
         int(bool(stochastic)), int(seed) & (2 ** 64 - 1), 0, _ptr(step_dev), _stream())
    return gq, hi, lo


def wt_pack(w_hat):
    """w_hat fp32 [Cout, Cin, kh, kw] (or [Cout, Cin]) -> bf16 [2, Cin, taps, Cout] hi / lo planes (qb_wt_pack)."""
    w = w_hat if w_hat.is_contiguous() else w_hat.contiguous()
    Cout, Cin = w.shape[0], w.shape[1]
    taps = w.numel() // (Cout * Cin)
    wt = torch.empty(2, Cin, taps, Cout, dtype=torch.bfloat16, device=w.device)
    _run("qb_wt_pack", 8 * w.numel(), _ptr(w), _ptr(wt), Cout, Cin, taps, _stream())
    return wt


def conv_bwd_supported(Cin, Cout, kh, kw, stride, pad):
    return Cin % 16 == 0 and Cout % 8 == 0 and kh * kw <= 9 and stride in (1, 2)


def conv_dgrad(a0, a1, wt, x_shape, Cout, kh, kw, stride, pad, scale_range=None, scale_bits=8, scale_mul=1.0, accumulate_into=None):
    """dX [B, Cin, H, W] (channels-last fp32) of a convolution from the bf16 plane(s) of its output gradient (qb_conv_dgrad).
    ``accumulate_into``: a tensor of dX's shape and layout holding the other gradient of the layer's input (a residual skip);
    the result is added INTO it by the kernel's reduce-add stores (qb_conv_dgrad_acc) and that tensor is returned - or, where
    the accumulating kernel does not apply, by a separate in-place add."""
    global _next_ops
    B, Cin, H, W = x_shape
    if accumulate_into is not None:
        acc = accumulate_into
        ok = (a1 is None and acc.dtype == torch.float32 and tuple(acc.shape) == (B, Cin, H, W) and is_nhwc(acc)
              and acc.data_ptr() % 16 == 0)
        if ok:
            _next_ops = 2 * a0.numel() * kh * kw * Cin
            _NLAUNCH["qb_conv_dgrad_acc"] = stride * stride
            try:
                _run("qb_conv_dgrad_acc", 2 * a0.numel() + 4 * Cout * Cin * kh * kw + 8 * acc.numel(), _ptr(a0), _ptr(wt), _ptr(acc), B, H,
                     W, Cin, Cout, kh, kw, stride, pad, _ptr(scale_range), int(scale_bits), float(scale_mul), _stream())
                return acc
            except QB200Error:
                _next_ops = 0
        dx = conv_dgrad(a0, a1, wt, x_shape, Cout, kh, kw, stride, pad, scale_range, scale_bits, scale_mul)
        return acc.add_(dx) if acc.shape == dx.shape else dx + acc
    dx = torch.empty((B, Cin, H, W), dtype=torch.float32, device=a0.device, memory_format=torch.channels_last)
    Ho, Wo = (H + 2 * pad - kh) // stride + 1, (W + 2 * pad - kw) // stride + 1
    _next_ops = 2 * B * Ho * Wo * Cout * kh * kw * Cin          # useful operations (the kernel executes 2-3 bf16 planes of them)
    _NLAUNCH["qb_conv_dgrad"] = stride * stride if kh >= stride else 1
    _run("qb_conv_dgrad", (2 if a1 is None else 4) * a0.numel() + 4 * Cout * Cin * kh * kw + 4 * dx.numel(), _ptr(a0), _ptr(a1),
         _ptr(wt), _ptr(dx), B, H, W, Cin, Cout, kh, kw, stride, pad, _ptr(scale_range), int(scale_bits), float(scale_mul),
         _stream())
    return dx


def conv_wgrad(g_hi, g_lo, g_f32, x16, Cout, kh, kw, stride, pad, a_range, a_bits, want_bias=False):
    """(dW [Cout, Cin, kh, kw], db | None) from the bf16 planes of the output gradient and the bf16 input codes (qb_conv_wgrad)."""
    B, Cin, H, W = x16.shape
    dw = torch.empty((Cout, Cin, kh, kw), dtype=torch.float32, device=x16.device)
    db = torch.empty(Cout, dtype=torch.float32, device=x16.device) if want_bias else None
    nb = load().qb_conv_wgrad_ws_bytes(B, H, W, Cin, Cout, kh, kw, stride, pad)
    if nb == 0:
        raise QB200Error("conv_wgrad: unsupported geometry Cin=%d Cout=%d k=%dx%d stride=%d" % (Cin, Cout, kh, kw, stride))
    ws = workspace("conv_wgrad", nb, x16.device)
    global _next_ops
    _next_ops = 2 * g_hi.numel() * kh * kw * Cin                # useful operations (executed twice: hi + lo planes of g)
    _run("qb_conv_wgrad", (2 if g_lo is None else 4) * g_hi.numel() + 2 * x16.numel() + 4 * dw.numel(), _ptr(g_hi), _ptr(g_lo),
         _ptr(g_f32), _ptr(x16), _ptr(dw), _ptr(db), B, H, W, Cin, Cout, kh, kw, stride, pad, _ptr(a_range), int(a_bits), _ptr(ws),
         ws.numel(), _stream())
    return dw, db


# ----------------------------------------------------------------------------------------------
# generic direct convolution / linear (conv_generic.cu): every geometry the tensor-core kernels do not take
# ----------------------------------------------------------------------------------------------
def _gc_args(x_like, o_like, conv):
    """conv: (stride, padding, dilation, groups) with 2-tuples, or None for a linear layer."""
    def st4(t):
        if t.dim() == 4:
            return (C.c_int64 * 4)(*t.stride())
        return (C.c_int64 * 4)(t.stride(0), t.stride(1), 0, 0)
    return st4(x_like), st4(o_like)


def _gc_conv(w, conv):
    kh, kw = (w.shape[2], w.shape[3]) if w.dim() == 4 else (1, 1)
    if conv is None:
        return (C.c_int * 9)(1, 1, 1, 1, 0, 0, 1, 1, 1)
    (sh, sw), (ph, pw), (dh, dw), groups = conv
    return (C.c_int * 9)(kh, kw, sh, sw, ph, pw, dh, dw, groups)


def conv_generic_fwd(x, w, bias, conv):
    """F.conv2d(x, w, bias, *conv) / F.linear(x, w, bias) (conv=None; x [M, K]) on the native direct kernel."""
    x = _f32c(x, "x") if x.dim() == 4 else (x if x.is_contiguous() else x.contiguous())
    w = w if w.is_contiguous() else w.contiguous()
    if x.dim() == 4:
        B, Cin, H, W = x.shape
        (sh, sw), (ph, pw), (dh, dw), groups = conv
        Ho = (H + 2 * ph - dh * (w.shape[2] - 1) - 1) // sh + 1
        Wo = (W + 2 * pw - dw * (w.shape[3] - 1) - 1) // sw + 1
        out = torch.empty((B, w.shape[0], Ho, Wo), dtype=torch.float32, device=x.device,
                          memory_format=torch.channels_last if is_nhwc(x) else torch.contiguous_format)
    else:
        (B, Cin), H, W = x.shape, 1, 1
        out = torch.empty((B, w.shape[0]), dtype=torch.float32, device=x.device)
    xs, os_ = _gc_args(x, out, conv)
    _run("qb_conv_generic_fwd", 4 * (x.numel() + w.numel() + out.numel()), _ptr(x), _ptr(w), _ptr(bias), _ptr(out), B, Cin, H, W,
         w.shape[0], _gc_conv(w, conv), xs, os_, _stream())
    return out


def conv_generic_dgrad(g, w, x_shape, conv, like=None):
    g = g if (g.is_contiguous() or (g.dim() == 4 and g.is_contiguous(memory_format=torch.channels_last))) else g.contiguous()
    w = w if w.is_contiguous() else w.contiguous()
    nhwc = like is not None and like.dim() == 4 and is_nhwc(like)
    if len(x_shape) == 4:
        B, Cin, H, W = x_shape
        dx = torch.empty(tuple(x_shape), dtype=torch.float32, device=g.device,
                         memory_format=torch.channels_last if nhwc else torch.contiguous_format)
    else:
        (B, Cin), H, W = x_shape, 1, 1
        dx = torch.empty(tuple(x_shape), dtype=torch.float32, device=g.device)
    xs, os_ = _gc_args(dx, g, conv)
    _run("qb_conv_generic_dgrad", 4 * (g.numel() + w.numel() + dx.numel()), _ptr(g), _ptr(w), _ptr(dx), B, Cin, H, W, w.shape[0],
         _gc_conv(w, conv), xs, os_, _stream())
    return dx


def conv_generic_wgrad(g, x, w_shape, conv, want_w=True, want_bias=False):
    g = g if (g.is_contiguous() or (g.dim() == 4 and g.is_contiguous(memory_format=torch.channels_last))) else g.contiguous()
    x = x if (x.is_contiguous() or (x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last))) else x.contiguous()
    dw = torch.empty(tuple(w_shape), dtype=torch.float32, device=g.device) if want_w else None
    db = torch.empty(w_shape[0], dtype=torch.float32, device=g.device) if want_bias else None
    if x.dim() == 4:
        B, Cin, H, W = x.shape
    else:
        (B, Cin), H, W = x.shape, 1, 1
    xs, os_ = _gc_args(x, g, conv)
    kh, kw = (w_shape[2], w_shape[3]) if len(w_shape) == 4 else (1, 1)
    cv = (C.c_int * 9)(1, 1, 1, 1, 0, 0, 1, 1, 1) if conv is None else (C.c_int * 9)(kh, kw, conv[0][0], conv[0][1], conv[1][0], conv[1][1],
                                                                                       conv[2][0], conv[2][1], conv[3])
    _run("qb_conv_generic_wgrad", 4 * (g.numel() + x.numel()), _ptr(g), _ptr(x), _ptr(dw), _ptr(db), B, Cin, H, W, w_shape[0], cv, xs, os_,
         _stream())
    return dw, db


def conv_stem_supported(H, W, Cin, Cout, kh, kw, stride, pad):
    return bool(load().qb_conv_stem_supported(int(H), int(W), int(Cin), int(Cout), int(kh), int(kw), int(stride), int(pad)))


def conv_stem_fwd(x_codes, w_codes, B, H, W, Cin, Cout, kh, kw, stride, pad, a_range, a_bits, row_min, row_max, w_bits, tapsum,
                  bias=None):
    """First-layer convolution (3 input channels) on the integer codes -> fp32 [B, Cout, Ho, Wo] stored channels-last."""
    Ho, Wo = (H + 2 * pad - kh) // stride + 1, (W + 2 * pad - kw) // stride + 1
    out = torch.empty((B, Cout, max(Ho, 0), max(Wo, 0)), dtype=torch.float32, device=x_codes.device,
                      memory_format=torch.channels_last)
    ws = workspace("conv_stem", load().qb_conv_stem_ws_bytes(Cout), x_codes.device)
    global _next_ops
    _next_ops = 2 * B * Ho * Wo * Cout * kh * kw * Cin
    _run("qb_conv_stem_fwd", B * H * W * Cin + Cout * kh * kw * Cin + 4 * B * Ho * Wo * Cout, _ptr(x_codes), _ptr(w_codes),
         _ptr(out), B, H, W, Cin, Cout, kh, kw, stride, pad, _ptr(a_range), int(a_bits), _ptr(row_min), _ptr(row_max),
         int(w_bits), _ptr(tapsum), _ptr(bias), _ptr(ws), ws.numel(), _stream())
    return out


def conv_stem_wgrad(grad_out, x_codes, a_range, a_bits, Cout, kh, kw, stride, pad):
    """Weight gradient of the first-layer convolution from the (channels-last, fp32) output gradient and the input CODES:
    -> dw fp32 [Cout, Cin, kh, kw] (qb_conv_stem_wgrad; tf32 tensor cores, deterministic)."""
    B, Cin, H, W = x_codes.shape
    if not grad_out.is_contiguous(memory_format=torch.channels_last):
        raise QB200Error("conv_stem_wgrad: grad_out must be channels-last")
    dw = torch.empty((Cout, Cin, kh, kw), dtype=torch.float32, device=x_codes.device)
    ws = workspace("conv_stem_wgrad", load().qb_conv_stem_wgrad_ws_bytes(Cout, Cin, kh, kw), x_codes.device)
    global _next_ops
    _next_ops = 2 * grad_out.numel() * kh * kw * Cin
    _run("qb_conv_stem_wgrad", 4 * grad_out.numel() + x_codes.numel(), _ptr(grad_out), _ptr(x_codes), _ptr(dw), B, H, W, Cin, Cout,
         kh, kw, stride, pad, _ptr(a_range), int(a_bits), _ptr(ws), ws.numel(), _stream())
    return dw


def conv_stem_wgrad16_supported(B, H, W, Cout, k, stride, pad):
    return load().qb_conv_stem_wgrad16_ws_bytes(B, H, W, Cout, k, stride, pad) > 0


def conv_stem_wgrad16(g_hi, g_lo, x_codes, a_range, a_bits, Cout, k, stride, pad):
    """First-layer weight gradient on the bf16 wgrad kernel (qb_conv_stem_wgrad16) -> dw fp32 [Cout, 3, k, k]."""
    B, Cin, H, W = x_codes.shape
    nb = load().qb_conv_stem_wgrad16_ws_bytes(B, H, W, Cout, k, stride, pad)
    if nb == 0 or Cin != 3:
        raise QB200Error("conv_stem_wgrad16: unsupported geometry")
    ws = workspace("conv_stem_wgrad16", nb, x_codes.device)
    dw = torch.empty((Cout, 3, k, k), dtype=torch.float32, device=x_codes.device)
    global _next_ops
    _next_ops = 2 * g_hi.numel() * k * k * 3
    _NLAUNCH["qb_conv_stem_wgrad16"] = 3
    _run("qb_conv_stem_wgrad16", 4 * g_hi.numel() + x_codes.numel(), _ptr(g_hi), _ptr(g_lo), _ptr(x_codes), _ptr(dw), B, H, W, Cout, int(k),
         int(stride), int(pad), _ptr(a_range), int(a_bits), _ptr(ws), ws.numel(), _stream())
    return dw


def allreduce_small(t, mode, peer_ptrs, rank, world, slot_bytes):
    """In-place one-shot all-reduce of a small contiguous double / float tensor over peer memory (qb_allreduce_small)."""
    if t.dtype == torch.float64:
        dt = 0
    elif t.dtype == torch.float32:
        dt = 1
    else:
        raise QB200Error("allreduce_small: double or float tensors only")
    if not t.is_contiguous():
        raise QB200Error("allreduce_small: tensor must be contiguous")
    _run("qb_allreduce_small", 0, _ptr(t), t.numel(), dt, int(mode), peer_ptrs, int(rank), int(world), int(slot_bytes), _stream())
    return t


# ----------------------------------------------------------------------------------------------
# glue
# ----------------------------------------------------------------------------------------------
def maxpool2d_nhwc_fwd(x, k, stride, pad):
    """x: channels-last fp32 [B, C, H, W] -> (out channels-last, idx uint8 flat)."""
    B, Cn, H, W = x.shape
    Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    out = torch.empty((B, Cn, Ho, Wo), dtype=torch.float32, device=x.device, memory_format=torch.channels_last)
    idx = torch.empty(out.numel(), dtype=torch.uint8, device=x.device)
    _run("qb_maxpool2d_nhwc_fwd", 4 * x.numel() + 5 * out.numel(), _ptr(x), _ptr(out), _ptr(idx), B, H, W, Cn, int(k), int(stride),
         int(pad), _stream())
    return out, idx


def rangebn_apply_maxpool(x_codes, in_shape, vec, has_gamma, has_beta, q_range, q_bits, k, stride, pad):
    """maxpool(RangeBN-normalise(codes)) -> (out channels-last, idx uint8 flat) without the normalised tensor
    (qb_rangebn_apply_maxpool_nhwc); ``in_shape`` = logical [B, C, H, W] of the RangeBN input."""
    B, Cn, H, W = in_shape
    Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    out = torch.empty((B, Cn, Ho, Wo), dtype=torch.float32, device=x_codes.device, memory_format=torch.channels_last)
    idx = torch.empty(out.numel(), dtype=torch.uint8, device=x_codes.device)
    _run("qb_rangebn_apply_maxpool_nhwc", B * Cn * H * W + 5 * out.numel(), _ptr(x_codes), _ptr(out), _ptr(idx), B, H, W, Cn, int(k),
         int(stride), int(pad), _ptr(vec[0]), _ptr(vec[2]), _ptr(vec[3]) if has_gamma else None, _ptr(vec[4]) if has_beta else None,
         _ptr(q_range), int(q_bits), _stream())
    return out, idx


def maxpool2d_nhwc_bwd(g, idx, in_shape, k, stride, pad):
    B, Cn, H, W = in_shape
    dx = torch.empty(in_shape, dtype=torch.float32, device=g.device, memory_format=torch.channels_last)
    _run("qb_maxpool2d_nhwc_bwd", 5 * g.numel() + 4 * dx.numel(), _ptr(g), _ptr(idx), _ptr(dx), B, H, W, Cn, int(k), int(stride),
         int(pad), _stream())
    return dx


def avgpool2d_nhwc_fwd(x, k):
    B, Cn, H, W = x.shape
    out = torch.empty((B, Cn, H // k, W // k), dtype=torch.float32, device=x.device, memory_format=torch.channels_last)
    _run("qb_avgpool2d_nhwc_fwd", 4 * (x.numel() + out.numel()), _ptr(x), _ptr(out), B, H, W, Cn, int(k), _stream())
    return out


def avgpool2d_nhwc_bwd(g, in_shape, k):
    B, Cn, H, W = in_shape
    dx = torch.empty(in_shape, dtype=torch.float32, device=g.device, memory_format=torch.channels_last)
    _run("qb_avgpool2d_nhwc_bwd", 4 * (g.numel() + dx.numel()), _ptr(g), _ptr(dx), B, H, W, Cn, int(k), _stream())
    return dx


def cross_entropy_fwd(logits, target, ignore_index=-100):
    """(loss 0-dim tensor, row_state float[2*B], n_valid float[1]) of qb_cross_entropy_fwd."""
    lib = load()
    B, Cn = logits.shape
    out = torch.empty(2, dtype=torch.float32, device=logits.device)           # loss, n_valid
    row_state = torch.empty(2 * B, dtype=torch.float32, device=logits.device)
    ws = workspace("cross_entropy", lib.qb_cross_entropy_ws_bytes(B), logits.device)
    _NLAUNCH["qb_cross_entropy_fwd"] = 2
    _run("qb_cross_entropy_fwd", 4 * logits.numel(), _ptr(logits), _ptr(target), B, Cn, int(ignore_index), _ptr(out[0:1]),
         _ptr(row_state), _ptr(out[1:2]), _ptr(ws), ws.numel(), _stream())
    return out[0], row_state, out[1:2]


def cross_entropy_bwd(logits, target, row_state, n_valid, grad_out, ignore_index=-100):
    B, Cn = logits.shape
    dx = torch.empty_like(logits)
    _run("qb_cross_entropy_bwd", 8 * logits.numel(), _ptr(logits), _ptr(target), _ptr(row_state), _ptr(n_valid), _ptr(grad_out),
         _ptr(dx), B, Cn, int(ignore_index), _stream())
    return dx


def maxpool2d_relu_nhwc_bwd(g, idx, pool_out, in_shape, k, stride, pad):
    """Backward of ReLU -> MaxPool2d in one pass: (dx w.r.t. the ReLU input, device float[2] range of dx)."""
    B, Cn, H, W = in_shape
    dx = torch.empty(in_shape, dtype=torch.float32, device=g.device, memory_format=torch.channels_last)
    rng = torch.empty(2, dtype=torch.float32, device=g.device)
    ws = workspace("pool_range", load().qb_maxpool2d_relu_bwd_ws_bytes(), g.device)
    _run("qb_maxpool2d_relu_nhwc_bwd", 9 * g.numel() + 4 * dx.numel(), _ptr(g), _ptr(idx), _ptr(pool_out), _ptr(dx), B, H, W, Cn,
         int(k), int(stride), int(pad), _ptr(rng), _ptr(ws), ws.numel(), _stream())
    return dx, rng


# ----------------------------------------------------------------------------------------------
# multi-tensor optimizer / regulariser glue (glue_opt.cu)
# ----------------------------------------------------------------------------------------------
MT_MAX = 96


def _mt_arrays(*lists):
    n = len(lists[0])
    ptrs = [(C.c_void_p * n)(*[t.data_ptr() for t in lst]) if lst is not None else None for lst in lists]
    sizes = (C.c_int64 * n)(*[t.numel() for t in lists[0]])
    return ptrs, sizes, n


def _mt_check(ts, name):
    for t in ts:
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
            raise QB200Error("%s: tensors must be contiguous CUDA float32" % name)


def l2_sumsq(tensors, weight_decay, out=None):
    """-> device scalar 0.5 * wd * SUM w^2 over all tensors (qb_l2_sumsq; one launch pair per 96 tensors)."""
    _mt_check(tensors, "l2_sumsq")
    dev = tensors[0].device
    if out is None:
        out = torch.empty((), dtype=torch.float32, device=dev)
    for k in range(0, len(tensors), MT_MAX):
        (w,), sizes, n = _mt_arrays(tensors[k:k + MT_MAX])
        ws = workspace("mt_sumsq", load().qb_mt_ws_bytes(n, sizes), dev)
        _NLAUNCH["qb_l2_sumsq"] = 2
        _run("qb_l2_sumsq", 4 * sum(t.numel() for t in tensors[k:k + MT_MAX]), w, sizes, n, float(weight_decay), _ptr(out), int(k > 0),
             _ptr(ws), ws.numel(), _stream())
    return out


def l2_sumsq_noalloc(tensors, weight_decay, out, accumulate=False):
    """qb_l2_sumsq_noalloc: as l2_sumsq for <= 96 tensors, scratch in the library's static device array - no device
    allocation here or in the library (`out` is the caller's preallocated device scalar).  Added for callers that cannot
    allocate; no counterpart in the reference."""
    _mt_check(tensors, "l2_sumsq_noalloc")
    (w,), sizes, n = _mt_arrays(tensors)
    _NLAUNCH["qb_l2_sumsq_noalloc"] = 2
    _run("qb_l2_sumsq_noalloc", 4 * sum(t.numel() for t in tensors), w, sizes, n, float(weight_decay), _ptr(out), int(accumulate), _stream())
    return out


def l2_grad(tensors, weight_decay, upstream=None):
    """-> [wd * upstream * w for w in tensors] in ONE launch per 96 tensors (qb_l2_grad)."""
    _mt_check(tensors, "l2_grad")
    grads = [torch.empty_like(t) for t in tensors]
    for k in range(0, len(tensors), MT_MAX):
        (w, g), sizes, n = _mt_arrays(tensors[k:k + MT_MAX], grads[k:k + MT_MAX])
        _run("qb_l2_grad", 8 * sum(t.numel() for t in tensors[k:k + MT_MAX]), w, g, sizes, n, float(weight_decay), _ptr(upstream), _stream())
    return grads


def sgd_step(params, grads, bufs, lr, momentum, weight_decay, first):
    """torch.optim.SGD update of every parameter in one launch per 96 tensors (qb_sgd_step)."""
    _mt_check(params, "sgd_step")
    _mt_check(grads, "sgd_step")
    for k in range(0, len(params), MT_MAX):
        (g, p, b), sizes, n = _mt_arrays(grads[k:k + MT_MAX], params[k:k + MT_MAX], bufs[k:k + MT_MAX] if bufs is not None else None)
        _run("qb_sgd_step", 16 * sum(t.numel() for t in params[k:k + MT_MAX]), g, p, b, sizes, n, float(lr), float(momentum),
             float(weight_decay), int(bool(first)), _stream())


# ----------------------------------------------------------------------------------------------
# quantize2.py flavour (row A9)
# ----------------------------------------------------------------------------------------------
def fakequant2(x, zero_point, rng, bits, per_row=False, signed=False, dequantize=True, noise=None, stochastic=False, seed=0,
               step_dev=None, out=None):
    """x viewed as [rows, cols] (rows = x.shape[0] when per_row else 1); zero_point / rng: device fp32 tensors with
    rows (or 1) elements.  See qb_fakequant2."""
    x = _f32c(x, "x")
    rows = x.shape[0] if per_row else 1
    cols = x.numel() // max(rows, 1)
    if out is None:
        out = torch.empty_like(x)
    zp = zero_point.detach().reshape(-1).contiguous().float()
    rg = rng.detach().reshape(-1).contiguous().float()
    if zp.numel() != (rows if per_row else 1) or rg.numel() != zp.numel():
        raise QB200Error("fakequant2: qparams must hold one value per tensor or one per row of dim 0")
    _run("qb_fakequant2", 8 * x.numel(), _ptr(x), _ptr(out), rows, cols, _ptr(zp), _ptr(rg), int(per_row), int(bits),
         int(bool(signed)), int(bool(dequantize)), _ptr(noise), int(bool(stochastic)), int(seed) & (2 ** 64 - 1), 0,
         _ptr(step_dev), _stream())
    return out


def fakequant2_codes(x, zero_point, rng, bits, codes=None, codes16=None, want_y=False):
    """quantize2's per-tensor unsigned quantizer -> integer codes (uint8 / bf16) [+ dequantised y] (qb_fakequant2_codes)."""
    x = _f32c(x, "x")
    y = torch.empty_like(x) if want_y else None
    zp = zero_point.detach().reshape(-1).contiguous().float()
    rg = rng.detach().reshape(-1).contiguous().float()
    if zp.numel() != 1 or rg.numel() != 1:
        raise QB200Error("fakequant2_codes: per-tensor qparams only")
    nb = 4 + (4 if want_y else 0) + (1 if codes is not None else 0) + (2 if codes16 is not None else 0)
    _run("qb_fakequant2_codes", nb * x.numel(), _ptr(x), _ptr(y), _ptr(codes), _ptr(codes16), x.numel(), _ptr(zp), _ptr(rg), int(bits),
         _stream())
    return y


def row_minmax(x):
    """-> (row_min, row_max) fp32[x.shape[0]] over everything but dim 0."""
    x = _f32c(x, "x")
    x = x if x.is_contiguous() else x.contiguous()
    rows = x.shape[0]
    mn = torch.empty(rows, dtype=torch.float32, device=x.device)
    mx = torch.empty(rows, dtype=torch.float32, device=x.device)
    _run("qb_row_minmax", 4 * x.numel(), _ptr(x), rows, x.numel() // rows, _ptr(mn), _ptr(mx), _stream())
    return mn, mx


def rangebn2_finalize(red, Cn, num_chunks, count, chunk_len, eps, momentum, training, running_mean, running_var):
    """-> vec float[4, C]: rows mean, scale, den = scale + eps, rinv = 1 / den."""
    vec = torch.empty(4, Cn, dtype=torch.float32, device=running_mean.device)
    _run("qb_rangebn2_finalize", 0, _ptr(red), Cn, int(num_chunks), float(count), float(chunk_len), float(eps), float(momentum),
         int(training), _ptr(running_mean), _ptr(running_var), _ptr(vec[0]), _ptr(vec[1]), _ptr(vec[2]), _ptr(vec[3]), _stream())
    return vec


def rangebn2_apply(x, B, Cn, HW, vec, gamma, beta, nhwc=False):
    out = torch.empty_like(x)
    _run("qb_rangebn2_apply", 8 * x.numel(), _ptr(x), _ptr(out), B, Cn, HW, int(nhwc), _ptr(vec[0]), _ptr(vec[2]), _ptr(gamma),
         _ptr(beta), _stream())
    return out
