"""Phase timing.

The reference measures wall clock per round only (``duration`` lists in ``logs.json``,
``local.py:51-52``).  Device mode adds:

* ``PhaseTimer`` — CUDA events around named phases of a step (zero-grad / forward / backward /
  aggregate+update), recorded on the launching stream and read after ONE synchronise, so the timer
  does not serialise the step it measures; ``summary()`` reduces over ranks with MAX (the step time
  of a synchronous job is the slowest site's).  Each phase is also an NVTX range, so a timeline tool
  attached to the process shows the same names.
* ``timed(key)`` — decorator for host-side methods of objects that own a ``cache`` dict: appends
  wall-clock seconds to ``cache['profile'][key]`` when ``cache['profile_enabled']`` is set
  (dumped with the rest of the cache into ``logs.json``).

On CPU ``PhaseTimer`` falls back to ``time.perf_counter``.
"""
from __future__ import annotations

import functools
import time
from contextlib import contextmanager

import torch


class PhaseTimer:
    def __init__(self, device=None, enabled=True, nvtx=True):
        self.device = torch.device(device) if device is not None else None
        self.cuda = bool(self.device is not None and self.device.type == "cuda" and torch.cuda.is_available())
        self.enabled = bool(enabled)
        self.nvtx = bool(nvtx) and self.cuda
        self._pending = []          # (name, start, stop) events or perf_counter floats
        self._acc = {}              # name -> [total_ms, max_ms, calls]
        self._order = []

    @contextmanager
    def phase(self, name):
        if not self.enabled:
            yield
            return
        if self.nvtx:
            torch.cuda.nvtx.range_push(name)
        if self.cuda:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(torch.cuda.current_stream(self.device))
        else:
            a = time.perf_counter()
        try:
            yield
        finally:
            if self.cuda:
                b.record(torch.cuda.current_stream(self.device))
                self._pending.append((name, a, b))
            else:
                self._pending.append((name, a, time.perf_counter()))
            if self.nvtx:
                torch.cuda.nvtx.range_pop()

    def flush(self):
        """Resolve recorded events into milliseconds (one device synchronise)."""
        if not self._pending:
            return
        if self.cuda:
            torch.cuda.synchronize(self.device)
        for name, a, b in self._pending:
            ms = a.elapsed_time(b) if self.cuda else (b - a) * 1e3
            if name not in self._acc:
                self._acc[name] = [0.0, 0.0, 0]
                self._order.append(name)
            acc = self._acc[name]
            acc[0] += ms
            acc[1] = max(acc[1], ms)
            acc[2] += 1
        self._pending = []

    def reset(self):
        self._pending, self._acc, self._order = [], {}, []

    def summary(self, reduce=True):
        """``{phase: {'ms_mean', 'ms_max', 'calls'}}``; with ``reduce`` the mean and max are the MAX
        over ranks (needs an initialised process group; otherwise local values)."""
        self.flush()
        names = list(self._order)
        mean = [self._acc[n][0] / max(self._acc[n][2], 1) for n in names]
        mx = [self._acc[n][1] for n in names]
        if reduce and names:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
                dev = self.device if self.cuda else torch.device("cpu")
                t = torch.tensor([mean, mx], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                mean, mx = t[0].tolist(), t[1].tolist()
        return {n: {"ms_mean": mean[i], "ms_max": mx[i], "calls": self._acc[n][2]} for i, n in enumerate(names)}


def timed(key):
    """Decorator: ``cache['profile'][key].append(seconds)`` when ``cache['profile_enabled']``."""
    def deco(fn):
        @functools.wraps(fn)
        def wrap(self, *a, **kw):
            cache = getattr(self, "cache", None)
            if not (isinstance(cache, dict) and cache.get("profile_enabled")):
                return fn(self, *a, **kw)
            t0 = time.perf_counter()
            try:
                return fn(self, *a, **kw)
            finally:
                cache.setdefault("profile", {}).setdefault(key, []).append(time.perf_counter() - t0)
        return wrap
    return deco
