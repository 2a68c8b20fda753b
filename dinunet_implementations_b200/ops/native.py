"""ctypes loader for the in-tree native library ``_native/libdinunet_b200.so``.

The library is plain CUDA C++ with a C ABI (declarations in ``include/dinunet_b200.h``),
built by ``__graft_entry__.build()`` with ``nvcc -gencode arch=compute_100a,code=sm_100a``.
It has no PyTorch dependency: Python passes raw device pointers and the current stream.

Policy: on a machine with a CUDA device the library **must** load — GPU ops raise instead
of silently falling back to eager PyTorch.  On a CPU-only machine ``available()`` is False
and only the host-side entry points (file reducer, parsers) are used when present.
"""
from __future__ import annotations

import ctypes
import os
import threading

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "_native", "libdinunet_b200.so")

_lib = None
_lock = threading.Lock()
_load_error = None
launches = 0  # kernels launched through this module (bench.py reports it)


def _try_load():
    global _lib, _load_error
    with _lock:
        if _lib is not None or _load_error is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            _load_error = f"{LIB_PATH} not built (run python -c 'import __graft_entry__ as g; g.build()')"
            return None
        try:
            _lib = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
        except OSError as e:
            _load_error = f"cannot load {LIB_PATH}: {e}"
            return None
        return _lib


def lib():
    """The loaded library or a loud failure."""
    l = _try_load()
    if l is None:
        raise RuntimeError(f"native library unavailable: {_load_error}")
    return l


def available(cpu_only: bool = False) -> bool:
    """True when the library is loaded and (unless ``cpu_only``) a CUDA device exists."""
    if _try_load() is None:
        return False
    if cpu_only:
        return True
    import torch
    return torch.cuda.is_available()


def require_gpu():
    import torch
    if not torch.cuda.is_available():
        raise RuntimeError("this op needs a CUDA device (sm_100a)")
    return lib()


def check(code: int, what: str = ""):
    if code != 0:
        msg = ""
        try:
            f = lib().dnb_last_error
            f.restype = ctypes.c_char_p
            msg = (f() or b"").decode()
        except Exception:
            pass
        raise RuntimeError(f"native call {what} failed with code {code}: {msg}")


def count_launch(n: int = 1):
    global launches
    launches += n


def ptr(t):
    """Device/host pointer of a tensor (None → NULL)."""
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def stream():
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


# ----------------------------------------------------------------------- host-side entry points
def reduce_mean_npy(paths):
    """Mean of same-length flat ``.npy`` files with the native threaded reducer.

    Returns ``None`` when the library lacks the symbol (older build) so callers fall back.
    """
    l = _try_load()
    if l is None or not hasattr(l, "dnb_cpu_mean_f32"):
        return None
    arrs = [np.load(p, mmap_mode="r") for p in paths]
    n = int(arrs[0].shape[0])
    if any(a.dtype != arrs[0].dtype or a.shape[0] != n for a in arrs) or arrs[0].dtype not in (np.float32, np.float16):
        return None
    out = np.empty(n, dtype=np.float32)
    arrs = [np.ascontiguousarray(a) for a in arrs]
    ptrs = (ctypes.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])
    is_half = 1 if arrs[0].dtype == np.float16 else 0
    rc = l.dnb_cpu_mean_f32(ptrs, ctypes.c_int(len(arrs)), ctypes.c_longlong(n), ctypes.c_int(is_half),
                            ctypes.c_void_p(out.ctypes.data), ctypes.c_int(0))
    if rc != 0:
        return None
    return out


# ----------------------------------------------------------------------- ctypes mirrors of include/dinunet_b200.h
MAX_RANKS = 8
SIGNAL_BLOCKS = 160
SIGNAL_WORDS = 2 * SIGNAL_BLOCKS * MAX_RANKS


class DsgdDesc(ctypes.Structure):
    _fields_ = [
        ("grads", ctypes.c_void_p * MAX_RANKS),
        ("params", ctypes.c_void_p * MAX_RANKS),
        ("shadow_bf16", ctypes.c_void_p * MAX_RANKS),
        ("flags", ctypes.c_void_p * MAX_RANKS),
        ("m", ctypes.c_void_p), ("v", ctypes.c_void_p), ("out", ctypes.c_void_p),
        ("n", ctypes.c_longlong),
        ("nranks", ctypes.c_int), ("rank", ctypes.c_int),
        ("wire", ctypes.c_int), ("mode", ctypes.c_int), ("opt", ctypes.c_int),
        ("use_flags", ctypes.c_int), ("step", ctypes.c_int),
        ("lr", ctypes.c_float), ("beta1", ctypes.c_float), ("beta2", ctypes.c_float), ("eps", ctypes.c_float),
        ("weight_decay", ctypes.c_float), ("momentum", ctypes.c_float),
        ("grad_scale", ctypes.c_float),
        ("epoch", ctypes.c_uint32),
        ("dev_state", ctypes.c_void_p),
        ("stage", ctypes.c_void_p), ("zero_stage", ctypes.c_int),
        ("grads_mc", ctypes.c_void_p), ("params_mc", ctypes.c_void_p), ("shadow_mc", ctypes.c_void_p),
    ]


class GemmDesc(ctypes.Structure):
    _fields_ = [
        ("A", ctypes.c_void_p), ("lda", ctypes.c_longlong),
        ("B", ctypes.c_void_p), ("ldb", ctypes.c_longlong),
        ("M", ctypes.c_int), ("N", ctypes.c_int), ("K", ctypes.c_int),
        ("out", ctypes.c_void_p), ("ldo", ctypes.c_longlong),
        ("out_bf16", ctypes.c_int),
        ("outT", ctypes.c_void_p), ("ldt", ctypes.c_longlong),
        ("bias_n", ctypes.c_void_p), ("bias_m", ctypes.c_void_p),
        ("mask", ctypes.c_void_p), ("ldmask", ctypes.c_longlong),
        ("colsum", ctypes.c_void_p),
        ("alpha", ctypes.c_float),
        ("relu", ctypes.c_int), ("accumulate", ctypes.c_int), ("split_k", ctypes.c_int), ("block_n", ctypes.c_int),
        ("split_k_auto", ctypes.c_int),
        ("bn_gamma", ctypes.c_void_p), ("bn_beta", ctypes.c_void_p), ("bn_eps", ctypes.c_float),
        ("bn_xhat", ctypes.c_void_p), ("ldxh", ctypes.c_longlong), ("bn_invstd", ctypes.c_void_p),
    ]


def launch_count() -> int:
    l = _try_load()
    if l is None:
        return 0
    l.dnb_launch_count.restype = ctypes.c_longlong
    return int(l.dnb_launch_count())


def reset_launch_count():
    l = _try_load()
    if l is not None:
        l.dnb_reset_launch_count()


def _p(t):
    return 0 if t is None else t.data_ptr()


class LstmFwdDesc(ctypes.Structure):
    _fields_ = [
        ("PX", ctypes.c_void_p), ("Whh", ctypes.c_void_p * 2), ("bhh", ctypes.c_void_p * 2),
        ("C", ctypes.c_void_p), ("Hprev", ctypes.c_void_p), ("Hmean", ctypes.c_void_p),
        ("hT", ctypes.c_void_p), ("cT", ctypes.c_void_p), ("dbg", ctypes.c_void_p),
        ("B", ctypes.c_int), ("S", ctypes.c_int), ("H", ctypes.c_int), ("D", ctypes.c_int),
    ]


class LstmBwdDesc(ctypes.Structure):
    _fields_ = [
        ("P", ctypes.c_void_p), ("C", ctypes.c_void_p), ("Whh", ctypes.c_void_p * 2), ("WhhT", ctypes.c_void_p * 2),
        ("dHmean", ctypes.c_void_p), ("dP", ctypes.c_void_p),
        ("db_ih", ctypes.c_void_p * 2), ("db_hh", ctypes.c_void_p * 2),
        ("dbg", ctypes.c_void_p),
        ("accumulate", ctypes.c_int),
        ("B", ctypes.c_int), ("S", ctypes.c_int), ("H", ctypes.c_int), ("D", ctypes.c_int),
    ]


class HeadFwdDesc(ctypes.Structure):
    _fields_ = [
        ("x", ctypes.c_void_p), ("drop", ctypes.c_void_p),
        ("W1", ctypes.c_void_p), ("W2", ctypes.c_void_p), ("W3", ctypes.c_void_p),
        ("b1", ctypes.c_void_p), ("b2", ctypes.c_void_p), ("b3", ctypes.c_void_p),
        ("bn_w", ctypes.c_void_p), ("bn_b", ctypes.c_void_p),
        ("run_mean", ctypes.c_void_p), ("run_var", ctypes.c_void_p),
        ("momentum", ctypes.c_float), ("eps", ctypes.c_float), ("training", ctypes.c_int),
        ("a0", ctypes.c_void_p),
        ("xhat", ctypes.c_void_p), ("invstd", ctypes.c_void_p), ("a1", ctypes.c_void_p), ("a2", ctypes.c_void_p),
        ("logits", ctypes.c_void_p),
        ("B", ctypes.c_int), ("N0", ctypes.c_int), ("N1", ctypes.c_int), ("N2", ctypes.c_int), ("N3", ctypes.c_int),
        ("drop_out", ctypes.c_void_p), ("drop_p", ctypes.c_float), ("drop_seed", ctypes.c_ulonglong),
        ("drop_state", ctypes.c_void_p), ("num_batches_tracked", ctypes.c_void_p),
    ]


class HeadBwdDesc(ctypes.Structure):
    _fields_ = [
        ("dlogits", ctypes.c_void_p),
        ("a0", ctypes.c_void_p), ("drop", ctypes.c_void_p), ("xhat", ctypes.c_void_p), ("invstd", ctypes.c_void_p),
        ("a1", ctypes.c_void_p), ("a2", ctypes.c_void_p),
        ("W1", ctypes.c_void_p), ("W2", ctypes.c_void_p), ("W3", ctypes.c_void_p),
        ("bn_w", ctypes.c_void_p),
        ("dW1", ctypes.c_void_p), ("db1", ctypes.c_void_p), ("dbn_w", ctypes.c_void_p), ("dbn_b", ctypes.c_void_p),
        ("dW2", ctypes.c_void_p), ("db2", ctypes.c_void_p), ("dW3", ctypes.c_void_p), ("db3", ctypes.c_void_p),
        ("dx", ctypes.c_void_p), ("dz1", ctypes.c_void_p),
        ("training", ctypes.c_int), ("accumulate", ctypes.c_int),
        ("B", ctypes.c_int), ("N0", ctypes.c_int), ("N1", ctypes.c_int), ("N2", ctypes.c_int), ("N3", ctypes.c_int),
        ("dz2", ctypes.c_void_p),
    ]


def vp(t):
    """c_void_p of a tensor's storage (None -> NULL)."""
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def ll(v):
    return ctypes.c_longlong(int(v))


# ---- rankDAD device kernels (csrc/dad.cu) ---------------------------------------------------------------
DAD_R = 16
DAD_MAX_LAYERS = 8
DAD_MAX_OUT = 1152
DAD_MAX_SPANS = 32
DAD_MAX_PACK = 8


class DadLayerDesc(ctypes.Structure):
    _fields_ = [
        ("A", ctypes.c_void_p), ("AT", ctypes.c_void_p), ("D", ctypes.c_void_p), ("DT", ctypes.c_void_p),
        ("lda", ctypes.c_longlong), ("ldat", ctypes.c_longlong), ("ldd", ctypes.c_longlong), ("lddt", ctypes.c_longlong),
        ("N", ctypes.c_int), ("n_in", ctypes.c_int), ("n_out", ctypes.c_int), ("pad_", ctypes.c_int),
        ("ZN", ctypes.c_void_p), ("ZI", ctypes.c_void_p), ("ZO", ctypes.c_void_p), ("V0", ctypes.c_void_p),
        ("ldzn", ctypes.c_longlong), ("ldzi", ctypes.c_longlong), ("ldzo", ctypes.c_longlong),
        ("W", ctypes.c_void_p), ("U", ctypes.c_void_p), ("V", ctypes.c_void_p),
        ("fact_off", ctypes.c_longlong),
        ("Uc", ctypes.c_void_p), ("Vc", ctypes.c_void_p),
        ("G", ctypes.c_void_p), ("ldg", ctypes.c_longlong),
    ]


class DadDesc(ctypes.Structure):
    _fields_ = [
        ("layer", DadLayerDesc * DAD_MAX_LAYERS),
        ("fact", ctypes.c_void_p * MAX_RANKS),
        ("flags", ctypes.c_void_p * MAX_RANKS),
        ("slot_floats", ctypes.c_longlong), ("dense_off", ctypes.c_longlong),
        ("span_off", ctypes.c_longlong * DAD_MAX_SPANS),
        ("span_len", ctypes.c_int * DAD_MAX_SPANS),
        ("flat_grad", ctypes.c_void_p),
        ("dev_state", ctypes.c_void_p),
        ("grad_scale", ctypes.c_float),
        ("n_layers", ctypes.c_int), ("mode", ctypes.c_int), ("iters", ctypes.c_int), ("cluster", ctypes.c_int),
        ("nsplit", ctypes.c_int), ("nranks", ctypes.c_int), ("rank", ctypes.c_int), ("use_flags", ctypes.c_int),
        ("n_spans", ctypes.c_int), ("rank_cols", ctypes.c_int),
    ]


class DadPackDesc(ctypes.Structure):
    _fields_ = [
        ("src", ctypes.c_void_p * DAD_MAX_PACK), ("rows", ctypes.c_void_p * DAD_MAX_PACK),
        ("colsT", ctypes.c_void_p * DAD_MAX_PACK),
        ("lds", ctypes.c_longlong * DAD_MAX_PACK), ("ldr", ctypes.c_longlong * DAD_MAX_PACK),
        ("ldt", ctypes.c_longlong * DAD_MAX_PACK),
        ("N", ctypes.c_int * DAD_MAX_PACK), ("d", ctypes.c_int * DAD_MAX_PACK),
        ("count", ctypes.c_int),
    ]


class PsgdDesc(ctypes.Structure):
    _fields_ = [("grad", ctypes.c_void_p), ("E", ctypes.c_void_p), ("Q", ctypes.c_void_p), ("P", ctypes.c_void_p),
                ("n", ctypes.c_int), ("m", ctypes.c_int), ("r", ctypes.c_int), ("pass_", ctypes.c_int)]


def memset_async(t, value=0):
    """``cudaMemsetAsync`` over a contiguous tensor on the current stream (no kernel launch)."""
    l = require_gpu()
    l.dnb_memset_async.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t, ctypes.c_void_p]
    assert t.is_contiguous()
    check(l.dnb_memset_async(t.data_ptr(), int(value), t.numel() * t.element_size(), stream()), "dnb_memset_async")


def seg_copy(pairs):
    """``dst.copy_(src)`` for a list of ``(src, dst)`` contiguous fp32 tensor pairs in one launch (per 32 pairs)."""
    if not pairs:
        return
    l = require_gpu()
    k = len(pairs)
    src = (ctypes.c_void_p * k)(*[s.data_ptr() for s, _ in pairs])
    dst = (ctypes.c_void_p * k)(*[d.data_ptr() for _, d in pairs])
    ln = (ctypes.c_int * k)(*[s.numel() for s, _ in pairs])
    for s_, d_ in pairs:
        assert s_.dtype == d_.dtype and s_.element_size() == 4 and s_.is_contiguous() and d_.is_contiguous() \
            and s_.numel() == d_.numel()
    check(l.dnb_seg_copy(src, dst, ln, ctypes.c_int(k), stream()), "dnb_seg_copy")


MLP_SMALL_MAX_WIDTH, MLP_SMALL_MAX_BATCH, MLP_SMALL_MAX_LAYERS, MLP_SMALL_MAX_CLASSES = 256, 64, 8, 16


class MlpSmallDesc(ctypes.Structure):
    _L = MLP_SMALL_MAX_LAYERS
    _fields_ = [
        ("x", ctypes.c_void_p),
        ("W", ctypes.c_void_p * _L), ("gamma", ctypes.c_void_p * _L), ("beta", ctypes.c_void_p * _L),
        ("Wfc", ctypes.c_void_p), ("bfc", ctypes.c_void_p),
        ("act", ctypes.c_void_p * _L), ("xhat", ctypes.c_void_p * _L), ("invstd", ctypes.c_void_p * _L),
        ("logits", ctypes.c_void_p), ("dlogits", ctypes.c_void_p),
        ("dW", ctypes.c_void_p * _L), ("dgamma", ctypes.c_void_p * _L), ("dbeta", ctypes.c_void_p * _L),
        ("dWfc", ctypes.c_void_p), ("dbfc", ctypes.c_void_p), ("dx", ctypes.c_void_p),
        ("width", ctypes.c_int * _L),
        ("B", ctypes.c_int), ("in0", ctypes.c_int), ("L", ctypes.c_int), ("C", ctypes.c_int), ("accumulate", ctypes.c_int),
        ("eps", ctypes.c_float),
    ]
