"""FreeSurfer-volume MLP classifier.

Math and parameter names follow the reference's ``MSANNet`` (``comps/fs/models.py:4-31``):
per hidden size ``Linear(bias=False) → BatchNorm1d(track_running_stats=False) → ReLU
[→ Dropout(0.5) if the layer index is in dropout_in]``, then ``fc_out = Linear(last, out)``.
BatchNorm always normalises with *batch* statistics, also in eval (no running stats).
``state_dict`` keys are identical (``layers.<i>.0.weight``, ``layers.<i>.1.weight|bias``,
``fc_out.weight|bias``), so checkpoints interchange with the reference class.

On a CUDA device with ``fused=True`` the network runs in hand-written sm_100a kernels: widths up to 256 at
batch <= 64 (the reference's default 66-256-128-64-32-2) in ONE launch forward and ONE backward
(``ops.mlp_small`` / ``csrc/mlp_small.cu``), wider nets per layer on the tcgen05 GEMM with a fused
BatchNorm/ReLU epilogue (``ops.mlp``); other shapes fall back to the eager path below (logged once), which is also the oracle.
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
            from ..ops import mlp as _mlp, mlp_small as _small
            if _small.supported(self, x):        # default-size net: the whole network in one launch each way
                return _small.msannet_small_forward(self, x)
            if _mlp.supported(self, x):          # wide layers: tcgen05 GEMM with BatchNorm/ReLU epilogue per layer
                return _mlp.msannet_forward(self, x)
            if not getattr(self, "_warned_eager", False):
                self._warned_eager = True
                from ..utils import logger
                logger.warn(f"MSANNet: shape {tuple(x.shape)} / widths {self.hidden_sizes} are outside the fused kernels' "
                            "envelope (batch 2..256, widths multiple of 8); running the eager PyTorch path", True)
        if getattr(self, "_grad_overwrite", False) and self.training and torch.is_grad_enabled():
            # the engine skips the zero-fill of these gradients because the fused backward overwrites them; the eager
            # path (shape outside the fused envelope) accumulates, so clear them here
            for p in self.grad_overwrite_params():
                if p.grad is not None:
                    p.grad.zero_()
        return self.forward_eager(x)
