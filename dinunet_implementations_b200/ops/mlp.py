"""Fused sm_100a forward/backward of the FreeSurfer MLP hidden stack (K4).

Every hidden layer ``Linear(bias=False) → BatchNorm1d(batch statistics) → ReLU`` (reference
``comps/fs/models.py:13-23``) is one tcgen05 GEMM whose epilogue does the BatchNorm and the ReLU:
the weight matrix is the M operand and the batch the N tile, so one TMEM lane holds one feature
for the whole batch.  The same kernel serves the default 66→256→128→64→32 net and the widened
66→4096×6 net of BASELINE.json config 4.  ``fc_out`` (tiny, with bias) stays a torch op.

Backward per layer: ``dnb_bn_relu_backward`` (ReLU + BatchNorm) → weight gradient GEMM
(K = batch) accumulated into ``param.grad`` → data gradient GEMM against the transposed weights.
"""
from __future__ import annotations

import ctypes

import torch

from . import native
from .native import vp, ll
from .gemm import gemm_bf16
from .elementwise import cast_time_major, transpose_bf16, cast_bf16


def _r8(n):
    return (int(n) + 7) // 8 * 8


def supported(model, x):
    if not native.available() or x.dim() != 2:
        return False
    B = x.shape[0]
    return (2 <= B <= 256 and all(h % 8 == 0 for h in model.hidden_sizes) and not model.dropout_in
            and all(isinstance(blk[1], torch.nn.BatchNorm1d) and not blk[1].track_running_stats for blk in model.layers))


def _w16(model, p):
    sh = getattr(model, "_bf16_shadow", None)
    w = sh[id(p)] if (sh is not None and id(p) in sh) else cast_bf16(p.detach())
    if w.shape[1] % 8 != 0:                      # TMA needs a 16-byte row pitch: pad K once per call
        wp = torch.zeros(w.shape[0], _r8(w.shape[1]), device=w.device, dtype=torch.bfloat16)
        wp[:, :w.shape[1]].copy_(w)
        w = wp[:, :w.shape[1]]
    return w


def _grad_buf(p):
    if p.grad is None:
        p.grad = torch.zeros_like(p)
    return p.grad


def bn_relu_backward(daT, xhatT, gamma, beta, invstd, B, dgamma, dbeta, accumulate=True):
    lib = native.require_gpu()
    N = daT.shape[0]
    dev = daT.device
    dzT = torch.empty(N, _r8(B), device=dev, dtype=torch.bfloat16)
    dz = torch.empty(B, _r8(N), device=dev, dtype=torch.bfloat16)
    native.check(lib.dnb_bn_relu_backward(vp(daT), ll(daT.stride(0)), vp(xhatT), ll(xhatT.stride(0)), vp(gamma), vp(beta),
                                          vp(invstd), ctypes.c_int(N), ctypes.c_int(B), vp(dzT), ll(dzT.stride(0)),
                                          vp(dz), ll(dz.stride(0)), vp(dgamma), vp(dbeta), ctypes.c_int(int(accumulate)),
                                          native.stream()), "dnb_bn_relu_backward")
    return dzT[:, :B], dz[:, :N]


class _MSANHidden(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, model, *params):
        B, K0 = x.shape
        Bp = _r8(B)
        a, aT = cast_time_major(x.float().view(B, 1, K0), want_transpose=True)     # [B,K0] bf16, [K0,B] bf16
        saved = []
        for blk in model.layers:
            lin, bnm = blk[0], blk[1]
            N = lin.out_features
            W = _w16(model, lin.weight)
            xhatT = torch.empty(N, Bp, device=x.device, dtype=torch.float32)[:, :B]
            invstd = torch.empty(N, device=x.device, dtype=torch.float32)
            outT_feat = torch.empty(N, Bp, device=x.device, dtype=torch.bfloat16)[:, :B]    # feature-major
            out_batch = torch.empty(B, N, device=x.device, dtype=torch.bfloat16)            # batch-major
            gemm_bf16(W, a, out=outT_feat, out_t=out_batch, relu=True,
                      bn=dict(gamma=bnm.weight.detach(), beta=bnm.bias.detach(), eps=bnm.eps, xhat=xhatT, invstd=invstd))
            saved.append((aT, xhatT, invstd, W))
            a, aT = out_batch, outT_feat
        ctx.model, ctx.saved, ctx.B = model, saved, B
        return a.float()

    @staticmethod
    def backward(ctx, d_out):
        model, saved, B = ctx.model, ctx.saved, ctx.B
        ctx.saved = None
        Bp = _r8(B)
        daT = torch.zeros(d_out.shape[1], Bp, device=d_out.device, dtype=torch.float32)
        daT[:, :B].copy_(d_out.t())
        daT = daT[:, :B]
        # engine-driven steps with one micro-batch per aggregation let the kernels STORE the parameter gradients
        # (model._grad_overwrite): no zero-fill of the flat gradient buffer and no atomic read-modify-write of the
        # [4096, 4096] weight gradients — with `accumulate` the weight-gradient GEMM was 108 us per wide layer
        acc = not getattr(model, "_grad_overwrite", False)
        for li in range(len(model.layers) - 1, -1, -1):
            blk = model.layers[li]
            lin, bnm = blk[0], blk[1]
            aT_prev, xhatT, invstd, W = saved[li]
            dzT, dz = bn_relu_backward(daT, xhatT, bnm.weight.detach(), bnm.bias.detach(), invstd, B,
                                       _grad_buf(bnm.weight), _grad_buf(bnm.bias), accumulate=acc)
            gemm_bf16(dzT, aT_prev, out=_grad_buf(lin.weight), accumulate=acc, split_k=1)
            if li > 0:
                WT = transpose_bf16(W)                                                  # [K, N]
                daT = torch.empty(W.shape[1], Bp, device=d_out.device, dtype=torch.float32)[:, :B]
                gemm_bf16(WT, dz, out=daT)
        return (None, None) + (None,) * (len(ctx.needs_input_grad) - 2)


def msannet_forward(model, x):
    params = [p for blk in model.layers for p in blk.parameters() if p.requires_grad]
    h = _MSANHidden.apply(x, model, *params)
    return model.fc_out(h)
