"""Python face of the tcgen05 + TMA GEMM (``csrc/gemm.cu``)."""
from __future__ import annotations

import ctypes

import torch

from . import native


def _check_bf16_2d(t, name):
    if t.dtype != torch.bfloat16 or t.dim() != 2 or t.stride(1) != 1:
        raise ValueError(f"{name} must be a 2-D bf16 tensor with unit inner stride, got {t.dtype} {tuple(t.shape)} "
                         f"strides {t.stride()}")
    if t.stride(0) % 8 != 0 or t.data_ptr() % 16 != 0:
        raise ValueError(f"{name}: leading dimension must be a multiple of 8 and the base 16-byte aligned "
                         f"(ld={t.stride(0)})")


def gemm_bf16(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor | None = None, *, out_dtype=torch.float32,
              out_t: torch.Tensor | None = None, bias_n=None, bias_m=None, relu=False, mask=None, colsum=None,
              alpha=1.0, accumulate=False, split_k=1, block_n=0, want_out=True, bn=None):
    """``out[M,N] = alpha * a[M,K] @ b[N,K].T`` with the fused epilogue of ``DnbGemmDesc``.

    ``a``/``b`` bf16 row-major (possibly strided views with ld % 8 == 0).  Returns ``out``
    (allocated if None and ``want_out``); ``out_t`` receives a transposed bf16 copy.
    """
    lib = native.require_gpu()
    _check_bf16_2d(a, "a")
    _check_bf16_2d(b, "b")
    M, K = a.shape
    N, Kb = b.shape
    if K != Kb:
        raise ValueError(f"K mismatch: a is {tuple(a.shape)}, b is {tuple(b.shape)}")
    if out is None and want_out:
        out = (torch.zeros if (split_k != 1) else torch.empty)((M, N), device=a.device, dtype=out_dtype)
    d = native.GemmDesc()
    d.A, d.lda, d.B, d.ldb = a.data_ptr(), a.stride(0), b.data_ptr(), b.stride(0)
    d.M, d.N, d.K = M, N, K
    if out is not None:
        if out.stride(1) != 1 or out.shape != (M, N) or out.dtype not in (torch.float32, torch.bfloat16):
            raise ValueError("out must be [M,N] fp32/bf16 with unit inner stride")
        d.out, d.ldo, d.out_bf16 = out.data_ptr(), out.stride(0), int(out.dtype == torch.bfloat16)
    if out_t is not None:
        if out_t.dtype != torch.bfloat16 or out_t.shape != (N, M) or out_t.stride(1) != 1:
            raise ValueError("out_t must be [N,M] bf16 with unit inner stride")
        d.outT, d.ldt = out_t.data_ptr(), out_t.stride(0)
    for nm, t, ln in (("bias_n", bias_n, N), ("bias_m", bias_m, M), ("colsum", colsum, N)):
        if t is not None:
            if t.dtype != torch.float32 or t.numel() != ln or not t.is_contiguous():
                raise ValueError(f"{nm} must be contiguous fp32 of length {ln}")
            setattr(d, nm, t.data_ptr())
    if mask is not None:
        if mask.dtype != torch.bfloat16 or mask.shape != (M, N) or mask.stride(1) != 1:
            raise ValueError("mask must be [M,N] bf16")
        d.mask, d.ldmask = mask.data_ptr(), mask.stride(0)
    if bn is not None:
        # bn = dict(gamma [M], beta [M], eps, xhat [M,N] fp32 or None, invstd [M] or None)
        d.bn_gamma, d.bn_beta, d.bn_eps = bn["gamma"].data_ptr(), bn["beta"].data_ptr(), float(bn.get("eps", 1e-5))
        if bn.get("xhat") is not None:
            d.bn_xhat, d.ldxh = bn["xhat"].data_ptr(), bn["xhat"].stride(0)
        if bn.get("invstd") is not None:
            d.bn_invstd = bn["invstd"].data_ptr()
    d.alpha, d.relu, d.accumulate, d.split_k, d.block_n = float(alpha), int(relu), int(accumulate), int(split_k), int(block_n)
    native.check(lib.dnb_gemm_bf16(ctypes.byref(d), native.stream()), "dnb_gemm_bf16")
    return out
