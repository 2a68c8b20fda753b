"""File transport for the COINSTAC compatibility mode (SURVEY §2.6 "data plane").

Payloads are numpy arrays written to a node's ``transferDirectory``; the COINSTAC runtime
(or our in-repo simulator) copies them to the peer's ``baseDirectory`` — for the remote
under ``baseDirectory/<site_id>/``.  ``precision_bits`` (``compspec.json:161-176``) selects
the wire dtype.  The cross-site mean is done by the native threaded reducer when the
library is built, else by an ``mp.Pool`` over memory-mapped chunks (the reference's
``num_reducers`` pool, ``remote.py:20-21``), else by numpy.
"""
from __future__ import annotations

import os

import numpy as np


def wire_dtype(precision_bits):
    return np.float16 if str(precision_bits) == "16" else np.float32


def save_arrays(path: str, arrays: dict, precision_bits="32"):
    """``np.savez`` of named arrays cast to the wire dtype.  Returns the file name."""
    dt = wire_dtype(precision_bits)
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    with open(path, "wb") as f:
        np.savez(f, **{k: np.ascontiguousarray(v, dtype=dt) for k, v in arrays.items()})
    return os.path.basename(path)


def load_arrays(path: str) -> dict:
    with np.load(path) as z:
        return {k: z[k].astype(np.float32) for k in z.files}


def save_flat(path: str, flat: np.ndarray, precision_bits="32"):
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    np.save(path, np.ascontiguousarray(flat, dtype=wire_dtype(precision_bits)))
    return os.path.basename(path)


def load_flat(path: str) -> np.ndarray:
    return np.load(path).astype(np.float32, copy=False)


def _mean_chunk(paths, lo, hi):
    acc = None
    for p in paths:
        a = np.load(p, mmap_mode="r")[lo:hi].astype(np.float32)
        acc = a if acc is None else acc + a
    return acc / float(len(paths))


def mean_flat_files(paths, mp_pool=None, num_chunks=None):
    """Element-wise mean over same-length flat ``.npy`` files → fp32 array."""
    from ..ops import native
    if native.available(cpu_only=True):
        out = native.reduce_mean_npy(paths)
        if out is not None:
            return out
    n = int(np.load(paths[0], mmap_mode="r").shape[0])
    if mp_pool is not None and n >= (1 << 16):
        k = int(num_chunks or getattr(mp_pool, "_processes", 2) or 2)
        bounds = np.linspace(0, n, k + 1).astype(np.int64)
        parts = mp_pool.starmap(_mean_chunk, [(list(paths), int(bounds[i]), int(bounds[i + 1]))
                                              for i in range(k)])
        return np.concatenate(parts)
    return _mean_chunk(list(paths), 0, n)
