"""Running weighted means (``new_averages()`` in the trainer hooks)."""
from __future__ import annotations

import torch

_ROUND = 5


def _to_float(v):
    if isinstance(v, torch.Tensor):
        return float(v.detach().float().sum().item()) if v.numel() != 1 else float(v.item())
    return float(v)


class COINNAverages:
    """``num_averages`` independent weighted running means.

    ``add(value, n)`` adds ``value * n`` to the numerator and ``n`` to the denominator of
    slot ``index`` (reference call site: ``val.add(loss.item(), len(inputs))``,
    ``comps/fs/__init__.py:61``).  ``value`` may be a 0-d device tensor: it is parked and
    only read back when a number is requested, so the hot loop stays asynchronous.
    """

    def __init__(self, num_averages: int = 1, **kw):
        self.num_averages = int(num_averages)
        self.values = [0.0] * self.num_averages
        self.counts = [0.0] * self.num_averages
        self._pending = []  # (index, device_tensor, n)

    # -- accumulation -----------------------------------------------------------------
    def add(self, val=0, n=1, index=0):
        if isinstance(val, torch.Tensor) and val.is_cuda:
            self._pending.append((index, val.detach(), float(n)))
        else:
            self.values[index] += _to_float(val) * float(n)
            self.counts[index] += float(n)

    def _flush(self):
        if self._pending:
            stacked = torch.stack([p[1].float().reshape(()) for p in self._pending]).cpu().tolist()
            for (ix, _, n), v in zip(self._pending, stacked):
                self.values[ix] += v * n
                self.counts[ix] += n
            self._pending = []

    def accumulate(self, other: "COINNAverages"):
        self._flush()
        other._flush()
        for i in range(min(self.num_averages, other.num_averages)):
            self.values[i] += other.values[i]
            self.counts[i] += other.counts[i]
        return self

    def reset(self):
        self.values = [0.0] * self.num_averages
        self.counts = [0.0] * self.num_averages
        self._pending = []
        return self

    # -- read-out ---------------------------------------------------------------------
    @property
    def averages(self):
        self._flush()
        return [v / max(c, 1e-12) if c else 0.0 for v, c in zip(self.values, self.counts)]

    @property
    def average(self):
        return self.averages[0]

    def get(self):
        return [round(a, _ROUND) for a in self.averages]

    def extract(self, name=None):
        return self.average

    # -- cross-site -------------------------------------------------------------------
    def serialize(self):
        self._flush()
        return [list(map(float, self.values)), list(map(float, self.counts))]

    def update(self, values=None, counts=None, **kw):
        self._flush()
        for i in range(min(self.num_averages, len(values))):
            self.values[i] += float(values[i])
            self.counts[i] += float(counts[i])
        return self

    def reduce_sites(self, payloads):
        for p in payloads:
            if p:
                self.update(values=p[0], counts=p[1])
        return self

    def __repr__(self):
        return f"COINNAverages({self.get()})"
