"""Classification metrics: precision/recall/F1/accuracy, confusion matrix, ROC-AUC.

Anchors in the reference: hard predictions go in for the FreeSurfer task
(``comps/fs/__init__.py:57-59``), positive-class probabilities for the ICA task
(``comps/icalstm/__init__.py:65``); ``monitor_metric='auc'`` and
``log_header="loss|auc"`` (``local.py:36``) select what is tracked; the notebooks read
``test_metrics`` as ``[[loss, auc]]`` (``nnlogs.ipynb:85``).
"""
from __future__ import annotations

import numpy as np
import torch

_ROUND = 5
_EPS = 1e-12


def _np(x):
    if isinstance(x, torch.Tensor):
        return x.detach().reshape(-1).cpu().numpy()
    return np.asarray(x).reshape(-1)


def roc_auc(scores, labels) -> float:
    """Area under the ROC curve by the rank-sum (Mann-Whitney) identity, ties averaged.

    Returns 0.5 when only one class is present (undefined AUC), so early epochs on tiny
    validation splits do not crash the round machine.
    """
    s = np.asarray(scores, dtype=np.float64).reshape(-1)
    y = np.asarray(labels).reshape(-1).astype(np.int64)
    n_pos = int((y == 1).sum())
    n_neg = int(y.size - n_pos)
    if n_pos == 0 or n_neg == 0:
        return 0.5
    order = np.argsort(s, kind="mergesort")
    ranks = np.empty(s.size, dtype=np.float64)
    sorted_s = s[order]
    # average ranks over runs of equal scores
    boundaries = np.flatnonzero(np.diff(sorted_s)) + 1
    starts = np.concatenate(([0], boundaries))
    ends = np.concatenate((boundaries, [s.size]))
    avg = (starts + ends - 1) / 2.0 + 1.0
    ranks[order] = np.repeat(avg, ends - starts)
    return float((ranks[y == 1].sum() - n_pos * (n_pos + 1) / 2.0) / (n_pos * n_neg))


class COINNMetrics:
    """Base protocol; subclasses override the five methods below."""

    def add(self, *args, **kw):
        raise NotImplementedError

    def accumulate(self, other):
        raise NotImplementedError

    def reset(self):
        raise NotImplementedError

    def get(self):
        raise NotImplementedError

    def serialize(self):
        raise NotImplementedError

    def update(self, **kw):
        raise NotImplementedError

    def reduce_sites(self, payloads):
        for p in payloads:
            if p:
                self.update(**p) if isinstance(p, dict) else self.update(*p)
        return self

    def extract(self, name):
        v = getattr(self, str(name))
        return v() if callable(v) else v

    @property
    def time(self):  # kept for log symmetry with averages
        return 0.0


class Prf1a(COINNMetrics):
    """Binary precision / recall / F1 / accuracy from TP/FP/TN/FN counts."""

    def __init__(self, threshold: float = 0.5, **kw):
        self.threshold = threshold
        self.tn = self.fp = self.fn = self.tp = 0.0

    def add(self, pred, true):
        p, t = _np(pred), _np(true)
        if p.dtype.kind == "f":
            p = (p >= self.threshold)
        p = p.astype(np.int64)
        t = t.astype(np.int64)
        self.tp += float(((p == 1) & (t == 1)).sum())
        self.fp += float(((p == 1) & (t == 0)).sum())
        self.fn += float(((p == 0) & (t == 1)).sum())
        self.tn += float(((p == 0) & (t == 0)).sum())

    def accumulate(self, other):
        self.tp += other.tp
        self.fp += other.fp
        self.fn += other.fn
        self.tn += other.tn
        return self

    def reset(self):
        self.tn = self.fp = self.fn = self.tp = 0.0
        return self

    def update(self, tn=0, fp=0, fn=0, tp=0, **kw):
        self.tn += tn
        self.fp += fp
        self.fn += fn
        self.tp += tp
        return self

    def serialize(self):
        return {"tn": self.tn, "fp": self.fp, "fn": self.fn, "tp": self.tp}

    @property
    def precision(self):
        return self.tp / max(self.tp + self.fp, _EPS)

    @property
    def recall(self):
        return self.tp / max(self.tp + self.fn, _EPS)

    @property
    def accuracy(self):
        return (self.tp + self.tn) / max(self.tp + self.fp + self.fn + self.tn, _EPS)

    @property
    def f1(self):
        p, r = self.precision, self.recall
        return 2 * p * r / max(p + r, _EPS)

    @property
    def overlap(self):
        return self.tp / max(self.tp + self.fp + self.fn, _EPS)

    def get(self):
        return [round(v, _ROUND) for v in (self.precision, self.recall, self.f1, self.accuracy)]


class ConfusionMatrix(COINNMetrics):
    """``num_classes``-way confusion matrix; macro precision/recall/F1 and accuracy."""

    def __init__(self, num_classes: int = 2, **kw):
        self.num_classes = int(num_classes)
        self.matrix = np.zeros((self.num_classes, self.num_classes), dtype=np.float64)

    def add(self, pred, true):
        p, t = _np(pred).astype(np.int64), _np(true).astype(np.int64)
        idx = t * self.num_classes + p
        self.matrix += np.bincount(idx, minlength=self.num_classes ** 2).reshape(
            self.num_classes, self.num_classes)

    def accumulate(self, other):
        self.matrix += other.matrix
        return self

    def reset(self):
        self.matrix[:] = 0
        return self

    def update(self, matrix=None, **kw):
        self.matrix += np.asarray(matrix, dtype=np.float64)
        return self

    def serialize(self):
        return {"matrix": self.matrix.tolist()}

    def _per_class(self):
        tp = np.diag(self.matrix)
        prec = tp / np.maximum(self.matrix.sum(0), _EPS)
        rec = tp / np.maximum(self.matrix.sum(1), _EPS)
        return prec, rec

    @property
    def precision(self):
        return float(self._per_class()[0].mean())

    @property
    def recall(self):
        return float(self._per_class()[1].mean())

    @property
    def f1(self):
        p, r = self._per_class()
        return float((2 * p * r / np.maximum(p + r, _EPS)).mean())

    @property
    def accuracy(self):
        return float(np.trace(self.matrix) / max(self.matrix.sum(), _EPS))

    def get(self):
        return [round(v, _ROUND) for v in (self.precision, self.recall, self.f1, self.accuracy)]


class AUCROCMetrics(COINNMetrics):
    """ROC-AUC over every (score, label) pair seen, plus thresholded P/R/F1/accuracy.

    AUC is not decomposable into per-site scalars, so sites ship their raw
    (score, label) lists and the aggregator recomputes it on the union (SURVEY §2.3 E8:
    "metrics are aggregated as sufficient statistics, not averaged as scalars").
    Scores stay on their device until read; see ``ops.metrics`` for the device-side
    accumulator used by the in-process GPU engine.
    """

    def __init__(self, threshold: float = 0.5, **kw):
        self.threshold = threshold
        self.probabilities = []
        self.labels = []
        self._pending = []

    def add(self, pred, true):
        if isinstance(pred, torch.Tensor) and pred.is_cuda:
            self._pending.append((pred.detach().reshape(-1).float(), true.detach().reshape(-1)))
            return
        self.probabilities += _np(pred).astype(np.float64).tolist()
        self.labels += _np(true).astype(np.int64).tolist()

    def _flush(self):
        if self._pending:
            p = torch.cat([a for a, _ in self._pending]).cpu().numpy()
            t = torch.cat([b for _, b in self._pending]).cpu().numpy()
            self.probabilities += p.astype(np.float64).tolist()
            self.labels += t.astype(np.int64).tolist()
            self._pending = []

    def accumulate(self, other):
        self._flush()
        other._flush()
        self.probabilities += other.probabilities
        self.labels += other.labels
        return self

    def reset(self):
        self.probabilities, self.labels, self._pending = [], [], []
        return self

    def update(self, probabilities=None, labels=None, **kw):
        self._flush()
        self.probabilities += list(probabilities or [])
        self.labels += list(labels or [])
        return self

    def serialize(self):
        self._flush()
        return {"probabilities": self.probabilities, "labels": self.labels}

    @property
    def auc(self):
        self._flush()
        return roc_auc(self.probabilities, self.labels)

    def _prf(self):
        self._flush()
        m = Prf1a(self.threshold)
        if self.labels:
            m.add(np.asarray(self.probabilities, dtype=np.float64), np.asarray(self.labels))
        return m

    @property
    def accuracy(self):
        return self._prf().accuracy

    @property
    def f1(self):
        return self._prf().f1

    @property
    def precision(self):
        return self._prf().precision

    @property
    def recall(self):
        return self._prf().recall

    def get(self):
        return [round(self.auc, _ROUND)]
