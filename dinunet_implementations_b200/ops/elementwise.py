"""Python faces of the layout/cast kernels, the window gather (K10) and softmax-CE (K6)."""
from __future__ import annotations

import ctypes

import torch

from . import native
from .native import vp, ll


def _r8(n):
    return (int(n) + 7) // 8 * 8


def cast_time_major(x: torch.Tensor, want_transpose=True):
    """x fp32 ``[B,S,...]`` → ``rows`` bf16 ``[S*B, F]`` (row = t*B+b; storage ld = F rounded to 8)
    and ``colsT`` bf16 ``[F, S*B]`` (storage ld = S*B rounded to 8)."""
    lib = native.require_gpu()
    B, S = x.shape[0], x.shape[1]
    F = x[0, 0].numel()
    x = x.contiguous()
    rows = torch.empty(S * B, _r8(F), device=x.device, dtype=torch.bfloat16)
    if _r8(F) != F:
        rows.zero_()
    colsT = torch.empty(F, _r8(S * B), device=x.device, dtype=torch.bfloat16) if want_transpose else None
    native.check(lib.dnb_cast_time_major(vp(x), vp(rows), ll(rows.stride(0)), vp(colsT),
                                         ll(colsT.stride(0) if colsT is not None else 0), ctypes.c_int(B),
                                         ctypes.c_int(S), ctypes.c_int(F), native.stream()), "dnb_cast_time_major")
    return rows[:, :F], (colsT[:, :S * B] if colsT is not None else None)


def transpose_bf16(x: torch.Tensor):
    """``[R,C]`` bf16 (unit inner stride) → ``[C,R]`` view of a buffer with ld rounded up to 8."""
    lib = native.require_gpu()
    R, C = x.shape
    assert x.dtype == torch.bfloat16 and x.stride(1) == 1
    out = torch.empty(C, _r8(R), device=x.device, dtype=torch.bfloat16)
    native.check(lib.dnb_transpose_bf16(vp(x), ll(x.stride(0)), vp(out), ll(out.stride(0)), ctypes.c_int(R),
                                        ctypes.c_int(C), native.stream()), "dnb_transpose_bf16")
    return out[:, :R]


def cast_bf16(x: torch.Tensor, out: torch.Tensor | None = None):
    lib = native.require_gpu()
    x = x.contiguous()
    if out is None:
        out = torch.empty(x.shape, device=x.device, dtype=torch.bfloat16)
    native.check(lib.dnb_cast_f32_bf16(vp(x), vp(out), ll(x.numel()), native.stream()), "dnb_cast_f32_bf16")
    return out


def window_gather(data: torch.Tensor, window_size: int, window_stride: int, temporal_size: int):
    """Device version of ``tasks.icalstm.window_view``: ``[N,C,T]`` → ``[N,S,C,W]``."""
    lib = native.require_gpu()
    data = data.contiguous().float()
    N, C, T = data.shape
    S = int(temporal_size // window_size)
    out = torch.empty(N, S, C, window_size, device=data.device, dtype=torch.float32)
    native.check(lib.dnb_window_gather(vp(data), vp(out), *[ctypes.c_int(int(v)) for v in
                                                            (N, C, T, S, window_size, window_stride)],
                                       native.stream()), "dnb_window_gather")
    return out


# data pointers of tensors known to hold the constant 1.0 (the engine's persistent seed gradient): when the
# incoming gradient of the loss IS one of them, the saved dlogits are returned as they are — no multiply launch
UNIT_GRADS = set()


class _SoftmaxCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, log_softmax_out):
        lib = native.require_gpu()
        logits = logits.contiguous().float()
        labels = labels.contiguous().long()
        B, ncls = logits.shape
        prob = torch.empty_like(logits)
        dlogits = torch.empty_like(logits)
        loss = torch.empty((), device=logits.device, dtype=torch.float32)
        native.check(lib.dnb_softmax_ce(vp(logits), vp(labels), vp(prob), vp(dlogits), vp(loss), ctypes.c_int(B),
                                        ctypes.c_int(ncls), ctypes.c_int(int(log_softmax_out)), native.stream()),
                     "dnb_softmax_ce")
        ctx.save_for_backward(dlogits)
        ctx.mark_non_differentiable(prob)
        ctx.set_materialize_grads(False)       # no zero tensor for the gradient of prob
        return loss, prob

    @staticmethod
    def backward(ctx, gloss, gprob):
        (dlogits,) = ctx.saved_tensors
        if gloss is None:
            return None, None, None
        if gloss.data_ptr() in UNIT_GRADS:
            return dlogits, None, None
        return dlogits * gloss, None, None


def softmax_cross_entropy(logits, labels, log_softmax_out=False):
    """``(mean CE loss, softmax probabilities)`` in one kernel; CPU tensors fall back to torch."""
    if not logits.is_cuda:
        lp = torch.log_softmax(logits.float(), 1)
        return torch.nn.functional.nll_loss(lp, labels.long()), (lp if log_softmax_out else lp.exp()).detach()
    return _SoftmaxCE.apply(logits, labels, log_softmax_out)
