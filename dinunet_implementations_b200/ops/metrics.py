"""K9 — device-side metric accumulation for the in-process engine.

``DeviceMetricBuffer.append(scores, labels, loss)`` is one tiny launch (CUDA-graph capturable) that
stores the batch's positive-class scores / labels, bumps confusion counts and the loss sum, all in
device memory.  ``collect()`` does the single host read of an epoch and returns the same metric
objects the file mode uses (``COINNAverages``, ``AUCROCMetrics``), so cross-site reduction is
unchanged.
"""
from __future__ import annotations

import ctypes

import torch

from . import native
from .native import vp, ll
from ..metrics import COINNAverages, AUCROCMetrics, Prf1a


class DeviceMetricBuffer:
    def __init__(self, device, capacity: int = 1 << 20, threshold: float = 0.5):
        self.capacity = int(capacity)
        self.threshold = float(threshold)
        self.scores = torch.zeros(self.capacity, device=device, dtype=torch.float32)
        self.labels = torch.zeros(self.capacity, device=device, dtype=torch.int32)
        self.state = torch.zeros(5, device=device, dtype=torch.int64)     # cursor, tp, fp, tn, fn
        self.lossacc = torch.zeros(2, device=device, dtype=torch.float32)

    def append(self, scores: torch.Tensor, labels: torch.Tensor, loss: torch.Tensor | None = None, col: int | None = None):
        """scores ``[B]`` or ``[B, C]`` (column ``col``, default last); labels int64 ``[B]``; loss 0-d."""
        lib = native.require_gpu()
        scores = scores.detach()
        if scores.dim() == 1:
            scores = scores.unsqueeze(1)
        if scores.dtype != torch.float32 or scores.stride(1) != 1:
            scores = scores.float().contiguous()
        col = scores.shape[1] - 1 if col is None else int(col)
        labels = labels.detach().long().contiguous()
        lp = loss.detach().float().reshape(1) if loss is not None else None
        native.check(lib.dnb_metrics_append(vp(scores), ll(scores.stride(0)), ctypes.c_int(col), vp(labels), vp(lp),
                                            ctypes.c_int(scores.shape[0]), vp(self.scores), vp(self.labels),
                                            ctypes.c_int(self.capacity), vp(self.state), vp(self.lossacc),
                                            ctypes.c_float(self.threshold), native.stream()), "dnb_metrics_append")

    def collect(self, reset: bool = True):
        """One host sync: ``(COINNAverages, AUCROCMetrics, Prf1a)`` of everything appended so far."""
        st = self.state.cpu().tolist()
        n = min(int(st[0]), self.capacity)
        la = self.lossacc.cpu().tolist()
        avg = COINNAverages()
        if la[1] > 0:
            avg.add(la[0] / la[1], la[1])
        auc = AUCROCMetrics(self.threshold)
        if n:
            auc.update(probabilities=self.scores[:n].cpu().double().tolist(), labels=self.labels[:n].cpu().tolist())
        prf = Prf1a(self.threshold).update(tp=st[1], fp=st[2], tn=st[3], fn=st[4])
        if reset:
            self.reset()
        return avg, auc, prf

    def reset(self):
        self.state.zero_()
        self.lossacc.zero_()
