"""Python face of the classifier-head kernels (``csrc/head.cu``)."""
from __future__ import annotations

import ctypes

import torch

from . import native


def head_forward(x, drop, W1, b1, bn_w, bn_b, run_mean, run_var, W2, b2, W3, b3, training, momentum=0.1, eps=1e-5,
                 gen_drop=None, num_batches_tracked=None):
    """Returns ``(logits, saved)``; ``saved`` feeds ``head_backward``.  Weights bf16, rest fp32.

    Dropout: either pass the scale mask ``drop`` ``[B,N0]``, or ``gen_drop=(p, seed, state)`` with ``state`` a
    device int64 scalar — the kernel then draws the mask itself (no RNG launches) and advances ``state``.
    ``num_batches_tracked`` (device int64 scalar) is incremented by the kernel in training mode."""
    lib = native.require_gpu()
    B, N0 = x.shape
    N1, N2, N3 = W1.shape[0], W2.shape[0], W3.shape[0]
    dev = x.device
    gen = drop is None and gen_drop is not None and float(gen_drop[0]) > 0
    if gen:
        drop = torch.empty(B, N0, device=dev)          # written by the kernel
    a0 = torch.empty(B, N0, device=dev) if drop is not None else None
    xhat = torch.empty(B, N1, device=dev)
    invstd = torch.empty(N1, device=dev)
    a1 = torch.empty(B, N1, device=dev)
    a2 = torch.empty(B, N2, device=dev)
    logits = torch.empty(B, N3, device=dev)
    d = native.HeadFwdDesc()
    d.x, d.drop = x.data_ptr(), (drop.data_ptr() if (drop is not None and not gen) else None)
    if gen:
        p_, seed_, state_ = gen_drop
        assert state_.dtype == torch.int64 and state_.is_cuda
        d.drop_out, d.drop_p, d.drop_seed, d.drop_state = drop.data_ptr(), float(p_), int(seed_) & (2 ** 64 - 1), state_.data_ptr()
    if num_batches_tracked is not None:
        assert num_batches_tracked.dtype == torch.int64 and num_batches_tracked.is_cuda
        d.num_batches_tracked = num_batches_tracked.data_ptr()
    d.W1, d.W2, d.W3 = W1.data_ptr(), W2.data_ptr(), W3.data_ptr()
    d.b1, d.b2, d.b3 = b1.data_ptr(), b2.data_ptr(), b3.data_ptr()
    d.bn_w, d.bn_b = bn_w.data_ptr(), bn_b.data_ptr()
    d.run_mean = run_mean.data_ptr() if run_mean is not None else None
    d.run_var = run_var.data_ptr() if run_var is not None else None
    d.momentum, d.eps, d.training = float(momentum), float(eps), int(training)
    d.a0 = a0.data_ptr() if a0 is not None else None
    d.xhat, d.invstd, d.a1, d.a2, d.logits = (xhat.data_ptr(), invstd.data_ptr(), a1.data_ptr(), a2.data_ptr(),
                                              logits.data_ptr())
    d.B, d.N0, d.N1, d.N2, d.N3 = B, N0, N1, N2, N3
    native.check(lib.dnb_head_forward(ctypes.byref(d), native.stream()), "dnb_head_forward")
    saved = dict(a0=a0 if a0 is not None else x, drop=drop, xhat=xhat, invstd=invstd, a1=a1, a2=a2, W1=W1, W2=W2,
                 W3=W3, bn_w=bn_w, training=bool(training), dims=(B, N0, N1, N2, N3))
    return logits, saved


def head_backward(dlogits, saved, grads, accumulate=True, want_dz2=False):
    """``grads``: dict with fp32 tensors ``dW1 db1 dbn_w dbn_b dW2 db2 dW3 db3``.  Returns ``dx``."""
    lib = native.require_gpu()
    B, N0, N1, N2, N3 = saved["dims"]
    dlogits = dlogits.contiguous().float()
    dx = torch.empty(B, N0, device=dlogits.device)
    d = native.HeadBwdDesc()
    d.dlogits = dlogits.data_ptr()
    d.a0, d.drop = saved["a0"].data_ptr(), (saved["drop"].data_ptr() if saved["drop"] is not None else None)
    d.xhat, d.invstd, d.a1, d.a2 = (saved["xhat"].data_ptr(), saved["invstd"].data_ptr(), saved["a1"].data_ptr(),
                                    saved["a2"].data_ptr())
    d.W1, d.W2, d.W3, d.bn_w = saved["W1"].data_ptr(), saved["W2"].data_ptr(), saved["W3"].data_ptr(), saved["bn_w"].data_ptr()
    for k in ("dW1", "db1", "dbn_w", "dbn_b", "dW2", "db2", "dW3", "db3"):
        setattr(d, k, grads[k].data_ptr())
    d.dx = dx.data_ptr()
    dz1 = torch.empty(B, N1, device=dlogits.device)
    d.dz1 = dz1.data_ptr()
    if want_dz2:
        saved["dz2"] = torch.empty(B, N2, device=dlogits.device)
        d.dz2 = saved["dz2"].data_ptr()
    d.training, d.accumulate = int(saved["training"]), int(accumulate)
    d.B, d.N0, d.N1, d.N2, d.N3 = B, N0, N1, N2, N3
    native.check(lib.dnb_head_backward(ctypes.byref(d), native.stream()), "dnb_head_backward")
    saved["dz1"] = dz1          # gradient at the BatchNorm input (rankDAD capture reads it)
    saved["dlogits32"] = dlogits
    return dx
