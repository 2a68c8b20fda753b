"""Whole-network kernels for the default-size FreeSurfer MLP (``csrc/mlp_small.cu``): one launch
forward (all hidden layers with BatchNorm batch statistics + ReLU, then ``fc_out``), one launch backward (every
parameter gradient, written straight into ``param.grad`` — in device mode a slice of the flat gradient buffer).
Reference math: ``comps/fs/models.py:13-31``."""
from __future__ import annotations

import ctypes

import torch

from . import native
from .elementwise import cast_bf16


def supported(model, x):
    if not native.available() or x.dim() != 2 or not x.is_cuda:
        return False
    B = x.shape[0]
    ws = list(model.hidden_sizes)
    # shared-memory budget of the single-CTA kernels (mirrors launch_bwd in csrc/mlp_small.cu): da, dz and one staged
    # a_prev plane, the logits gradient and one padded weight buffer for the largest layer must fit 227 KB
    r4 = lambda n: (n + 3) // 4 * 4
    dims = [model.in_size] + ws
    plane = r4(B) * (r4(max(dims)) + 4) * 4
    wbytes = (max(r4(o) * ((i + 1) // 2 * 2 + 2) for i, o in zip(dims[:-1], dims[1:])) + 7) // 8 * 8 * 2 if ws else 0
    if 3 * plane + 4096 + wbytes > 227 * 1024:
        return False
    return (2 <= B <= native.MLP_SMALL_MAX_BATCH and 1 <= len(ws) <= native.MLP_SMALL_MAX_LAYERS
            and max(ws + [model.in_size]) <= native.MLP_SMALL_MAX_WIDTH and model.out_size <= native.MLP_SMALL_MAX_CLASSES
            and not model.dropout_in
            and all(isinstance(blk[1], torch.nn.BatchNorm1d) and not blk[1].track_running_stats for blk in model.layers))


def _w16(model, p):
    sh = getattr(model, "_bf16_shadow", None)
    if sh is not None and id(p) in sh:
        return sh[id(p)]
    return cast_bf16(p.detach())


def _grad_buf(p):
    if p.grad is None:
        p.grad = torch.zeros_like(p)
    return p.grad


def _desc(model, x, saved, logits):
    d = native.MlpSmallDesc()
    d.x = x.data_ptr()
    keep = []
    for l, blk in enumerate(model.layers):
        lin, bn = blk[0], blk[1]
        w = _w16(model, lin.weight)
        keep.append(w)
        d.W[l], d.gamma[l], d.beta[l] = w.data_ptr(), bn.weight.data_ptr(), bn.bias.data_ptr()
        d.act[l], d.xhat[l], d.invstd[l] = saved["act"][l].data_ptr(), saved["xhat"][l].data_ptr(), saved["invstd"][l].data_ptr()
        d.width[l] = lin.out_features
    wfc = _w16(model, model.fc_out.weight)
    keep.append(wfc)
    d.Wfc, d.bfc = wfc.data_ptr(), model.fc_out.bias.data_ptr()
    d.logits = logits.data_ptr()
    d.B, d.in0, d.L, d.C = x.shape[0], model.in_size, len(model.layers), model.out_size
    d.eps = float(model.layers[0][1].eps)
    return d, keep


class _MSANSmall(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, model, *params):
        lib = native.require_gpu()
        x = x.detach().float().contiguous()
        B, dev = x.shape[0], x.device
        saved = {"act": [torch.empty(B, h, device=dev) for h in model.hidden_sizes],
                 "xhat": [torch.empty(B, h, device=dev) for h in model.hidden_sizes],
                 "invstd": [torch.empty(h, device=dev) for h in model.hidden_sizes]}
        logits = torch.empty(B, model.out_size, device=dev)
        d, keep = _desc(model, x, saved, logits)
        native.check(lib.dnb_mlp_small_forward(ctypes.byref(d), native.stream()), "dnb_mlp_small_forward")
        ctx.model, ctx.x, ctx.saved, ctx.keep = model, x, saved, keep
        ctx.want_dx = bool(ctx.needs_input_grad[0])
        return logits

    @staticmethod
    def backward(ctx, dlogits):
        lib = native.require_gpu()
        model, x, saved = ctx.model, ctx.x, ctx.saved
        dlogits = dlogits.contiguous().float()
        d, keep = _desc(model, x, saved, dlogits)          # logits slot unused by the backward
        d.dlogits = dlogits.data_ptr()
        for l, blk in enumerate(model.layers):
            d.dW[l] = _grad_buf(blk[0].weight).data_ptr()
            d.dgamma[l], d.dbeta[l] = _grad_buf(blk[1].weight).data_ptr(), _grad_buf(blk[1].bias).data_ptr()
        d.dWfc, d.dbfc = _grad_buf(model.fc_out.weight).data_ptr(), _grad_buf(model.fc_out.bias).data_ptr()
        dx = torch.empty_like(x) if ctx.want_dx else None
        d.dx = dx.data_ptr() if dx is not None else None
        # bit 0: hidden layers add to their gradients, bit 1: fc_out adds.  Engine-driven single-micro-batch steps let the
        # hidden layers store (model._grad_overwrite; their spans of the flat buffer are then not zero-filled either)
        d.accumulate = 2 | (0 if getattr(model, "_grad_overwrite", False) else 1)
        native.check(lib.dnb_mlp_small_backward(ctypes.byref(d), native.stream()), "dnb_mlp_small_backward")
        ctx.saved = ctx.keep = None
        return (dx, None) + (None,) * (len(ctx.needs_input_grad) - 2)


def msannet_small_forward(model, x):
    params = [p for p in model.parameters() if p.requires_grad]
    return _MSANSmall.apply(x, model, *params)
