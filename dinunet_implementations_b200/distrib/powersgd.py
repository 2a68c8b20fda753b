"""powerSGD in file mode (engine name ``comps/__init__.py:16``).

Two aggregator round-trips per optimiser step:

  phase ``P``: backward; per ≥2-D gradient ``M = grad.view(n,-1) + E`` (local error
               feedback), send ``P = M·Q`` and every 1-D gradient dense  → mean over sites
  phase ``Q``: ``P̂ = orth(mean P)``, send ``Q = Mᵀ·P̂``                   → mean over sites
  finish     : ``M̂ = P̂·Q̄ᵀ``, ``E = M − M̂``, ``grad = M̂``, ``optimizer.step()``; ``Q̄``
               warm-starts the next step.  ``Q`` is seeded identically on every site.

The aggregator only ever averages flat payloads, so it reuses the dSGD reducer.
"""
from __future__ import annotations

import numpy as np
import torch

from . import COINNLearner, fileio
from . import lowrank as lr
from .dsgd import DSGDReducer
from ..utils.profiler import timed

P_FILE = "psgd_p.npy"
Q_FILE = "psgd_q.npy"


class PowerSGDLearner(COINNLearner):
    def __init__(self, **kw):
        super().__init__(**kw)
        self.rank = int(self.cache.get("powersgd_rank", 4))
        self.mats = [p for p in self.params if p.dim() >= 2]
        self.vecs = [p for p in self.params if p.dim() < 2]
        seed = int(self.cache.get("seed", 0))
        self.Q = [lr.psgd_init_q((p.shape[0], p.numel() // p.shape[0]), self._r(p), seed + 17 * i, p.device)
                  for i, p in enumerate(self.mats)]
        self.E = [torch.zeros(p.shape[0], p.numel() // p.shape[0], device=p.device) for p in self.mats]
        self.M = None
        self.P_hat = None
        self.vec_grads = None

    def _r(self, p):
        n, m = p.shape[0], p.numel() // p.shape[0]
        return max(1, min(self.rank, n, m))

    # -- helpers ----------------------------------------------------------------------
    def _pack(self, tensors):
        if not tensors:
            return np.zeros(0, dtype=np.float32)
        return torch.cat([t.detach().reshape(-1).float().cpu() for t in tensors]).numpy()

    def _unpack(self, flat, shapes, device):
        out, off = [], 0
        t = torch.from_numpy(np.ascontiguousarray(flat, dtype=np.float32))
        for s in shapes:
            n = int(np.prod(s))
            out.append(t[off:off + n].view(*s).to(device))
            off += n
        return out, off

    # -- protocol ---------------------------------------------------------------------
    @timed("learner_step")
    def step(self):
        phase = self.input.get("psgd_phase")
        name = self.input.get("avg_grads_file")
        if not name or phase is None or self.M is None:
            return {}
        flat = fileio.load_flat(self.incoming_path(name))
        if phase == "P":
            shapes = [(m.shape[0], self._r(p)) for m, p in zip(self.M, self.mats)]
            Ps, off = self._unpack(flat, shapes, self.mats[0].device if self.mats else "cpu")
            self.P_hat = [lr._orth_fast(P) for P in Ps]
            vec_shapes = [tuple(v.shape) for v in self.vecs]
            self.vec_grads, _ = self._unpack(flat[off:], vec_shapes, self.vecs[0].device if self.vecs else "cpu")
        elif phase == "Q":
            shapes = [(m.shape[1], self._r(p)) for m, p in zip(self.M, self.mats)]
            Qs, _ = self._unpack(flat, shapes, self.mats[0].device if self.mats else "cpu")
            for i, (p, M, Ph, Qm) in enumerate(zip(self.mats, self.M, self.P_hat, Qs)):
                M_hat, self.E[i] = lr.psgd_finish(M, Ph, Qm)
                p.grad = M_hat.view_as(p).to(p.dtype).clone()
                self.Q[i] = Qm
            for v, g in zip(self.vecs, self.vec_grads or []):
                v.grad = g.view_as(v).to(v.dtype).clone()
            self.optimizer_step()
            self.M = self.P_hat = self.vec_grads = None
        return {}

    @timed("learner_to_reduce")
    def to_reduce(self):
        bits = self.cache.get("precision_bits", "32")
        if self.P_hat is not None and self.M is not None:
            # second half of the step: no new batches
            Qs = [lr.psgd_q(M, Ph) for M, Ph in zip(self.M, self.P_hat)]
            name = fileio.save_flat(self.transfer_path(Q_FILE), self._pack(Qs), bits)
            return {"grads_file": name, "reduce": True, "psgd_phase": "Q"}, []
        its = self.backward()
        if not its:
            return {}, its
        self.M = []
        for p, E in zip(self.mats, self.E):
            g = p.grad if p.grad is not None else torch.zeros_like(p)
            self.M.append(g.detach().reshape(p.shape[0], -1).float() + E)
        Ps = [lr.psgd_p(M, Q) for M, Q in zip(self.M, self.Q)]
        dense = [v.grad if v.grad is not None else torch.zeros_like(v) for v in self.vecs]
        name = fileio.save_flat(self.transfer_path(P_FILE), self._pack(Ps + dense), bits)
        return {"grads_file": name, "reduce": True, "psgd_phase": "P"}, its


class PowerSGDReducer(DSGDReducer):
    @timed("reducer_reduce")
    def reduce(self):
        out = super().reduce()
        phases = {self.input[s].get("psgd_phase") for s in self.input}
        assert len(phases) == 1, f"sites out of phase: {phases}"
        out["psgd_phase"] = phases.pop()
        return out
