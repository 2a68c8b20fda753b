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
