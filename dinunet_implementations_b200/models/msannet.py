"""FreeSurfer-volume MLP classifier.

Math and parameter names follow the reference's ``MSANNet`` (``comps/fs/models.py:4-31``):
per hidden size ``Linear(bias=False) → BatchNorm1d(track_running_stats=False) → ReLU
[→ Dropout(0.5) if the layer index is in dropout_in]``, then ``fc_out = Linear(last, out)``.
BatchNorm always normalises with *batch* statistics, also in eval (no running stats).
``state_dict`` keys are identical (``layers.<i>.0.weight``, ``layers.<i>.1.weight|bias``,
``fc_out.weight|bias``), so checkpoints interchange with the reference class.

On a CUDA device with ``fused=True`` the whole network runs in the hand-written sm_100a
kernels of ``ops.mlp`` (single-CTA whole-net kernel for small widths, tcgen05 GEMM + fused
BN/ReLU for wide ones); the eager path below is the oracle.
"""
from __future__ import annotations

import torch
import torch.nn as nn


class MSANNet(nn.Module):
    def __init__(self, in_size, hidden_sizes, out_size, dropout_in=(), fused=True):
        super().__init__()
        self.in_size, self.out_size = int(in_size), int(out_size)
        self.hidden_sizes = [int(h) for h in hidden_sizes]
        self.dropout_in = tuple(dropout_in)
        self.fused = fused
        self.layers = nn.ModuleList()
        d = self.in_size
        for i, h in enumerate(self.hidden_sizes):
            block = [nn.Linear(d, h, bias=False), nn.BatchNorm1d(h, track_running_stats=False), nn.ReLU()]
            if i in self.dropout_in:
                block.append(nn.Dropout(p=0.5))
            self.layers.append(nn.Sequential(*block))
            d = h
        self.fc_out = nn.Linear(d, self.out_size)

    def grad_overwrite_params(self):
        """Parameters whose gradients the fused backward can STORE instead of accumulate (every one is written by
        exactly one kernel per backward); ``fc_out`` goes through autograd and keeps accumulating."""
        return [p for blk in self.layers for p in blk.parameters()]

    def forward_eager(self, x):
        for block in self.layers:
            x = block(x)
        return self.fc_out(x)

    def forward(self, x):
        if self.fused and x.is_cuda and not self.dropout_in:
            from ..ops import mlp as _mlp
            if _mlp.supported(self, x):
                return _mlp.msannet_forward(self, x)
        if getattr(self, "_grad_overwrite", False) and self.training and torch.is_grad_enabled():
            # the engine skips the zero-fill of these gradients because the fused backward overwrites them; the eager
            # path (shape outside the fused envelope) accumulates, so clear them here
            for p in self.grad_overwrite_params():
                if p.grad is not None:
                    p.grad.zero_()
        return self.forward_eager(x)
