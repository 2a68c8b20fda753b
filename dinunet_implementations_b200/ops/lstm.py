"""Python face of the persistent LSTM recurrence kernels (``csrc/lstm.cu``)."""
from __future__ import annotations

import ctypes

import torch

from . import native


def lstm_supported(H: int, D: int) -> bool:
    return H % 32 == 0 and 32 <= H <= 256 and D in (1, 2)


def lstm_forward(px, whh, bhh, B, S, H, D, dbg=None):
    """px ``[S*B, D*4H]`` fp32 (time-major rows, overwritten with the full pre-activations);
    whh: list of D bf16 ``[4H,H]``; bhh: list of D fp32 ``[4H]``.
    Returns ``(Hmean [B,D*H], hT, cT, C [D,S,B,H], Hprev bf16 [S*B, D*H])``."""
    lib = native.require_gpu()
    dev = px.device
    C = torch.empty(D, S, B, H, device=dev, dtype=torch.float32)
    Hprev = torch.empty(S * B, D * H, device=dev, dtype=torch.bfloat16)
    Hmean = torch.empty(B, D * H, device=dev, dtype=torch.float32)
    hT = torch.empty_like(Hmean)
    cT = torch.empty_like(Hmean)
    d = native.LstmFwdDesc()
    d.PX = px.data_ptr()
    for i in range(D):
        assert whh[i].dtype == torch.bfloat16 and whh[i].is_contiguous() and whh[i].shape == (4 * H, H)
        assert bhh[i].dtype == torch.float32 and bhh[i].is_contiguous()
        d.Whh[i], d.bhh[i] = whh[i].data_ptr(), bhh[i].data_ptr()
    d.C, d.Hprev, d.Hmean, d.hT, d.cT = C.data_ptr(), Hprev.data_ptr(), Hmean.data_ptr(), hT.data_ptr(), cT.data_ptr()
    d.dbg = dbg.data_ptr() if dbg is not None else None
    d.B, d.S, d.H, d.D = B, S, H, D
    native.check(lib.dnb_lstm_forward(ctypes.byref(d), native.stream()), "dnb_lstm_forward")
    return Hmean, hT, cT, C, Hprev


def lstm_backward(P, C, whh, dHmean, B, S, H, D, db_ih, db_hh=None, accumulate=False, dbg=None, whhT=None):
    """Returns ``dP`` bf16 ``[S*B, D*4H]``; bias gradients are written into ``db_ih`` / ``db_hh``
    (lists of D fp32 ``[4H]`` tensors).  ``whhT`` (optional list of D contiguous bf16 ``[H,4H]``): the
    recurrent weights already transposed — saves the in-kernel transposition (~6 us per launch)."""
    lib = native.require_gpu()
    dP = torch.empty(S * B, D * 4 * H, device=P.device, dtype=torch.bfloat16)
    d = native.LstmBwdDesc()
    d.P, d.C, d.dHmean, d.dP = P.data_ptr(), C.data_ptr(), dHmean.data_ptr(), dP.data_ptr()
    assert dHmean.dtype == torch.float32 and dHmean.is_contiguous()
    for i in range(D):
        d.Whh[i] = whh[i].data_ptr()
        if whhT is not None:
            assert whhT[i].dtype == torch.bfloat16 and whhT[i].is_contiguous() and whhT[i].shape == (H, 4 * H)
            d.WhhT[i] = whhT[i].data_ptr()
        d.db_ih[i] = db_ih[i].data_ptr()
        d.db_hh[i] = db_hh[i].data_ptr() if db_hh is not None else None
    d.accumulate = int(accumulate)
    d.dbg = dbg.data_ptr() if dbg is not None else None
    d.B, d.S, d.H, d.D = B, S, H, D
    native.check(lib.dnb_lstm_backward(ctypes.byref(d), native.stream()), "dnb_lstm_backward")
    return dP
