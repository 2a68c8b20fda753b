"""Aggregation engines: a site-side *Learner* and an aggregator-side *Reducer* per engine.

Engine names are the reference's (``comps/__init__.py:13-16``): ``"dSGD"``, ``"rankDAD"``,
``"powerSGD"`` — all three are accepted even though the GUI only lists two
(``compspec.json:58-61``).

Two data planes implement the same math:

* **file mode** (this package): payloads are numpy files relayed by COINSTAC / the
  simulator; the reducer runs on the CPU.  This is the compatibility mode and also the
  ``recon-file`` baseline.
* **device mode** (``parallel/``): one process per GPU, payloads never leave device memory
  and the reduce is fused with the optimiser step in sm_100a kernels over NVLink peer
  memory.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from ..config import AggEngine
from ..utils.profiler import timed


def trainable_params(nn: dict):
    ps = []
    for m in nn.values():
        ps += [p for p in m.parameters() if p.requires_grad]
    return ps


def flat_grads(params) -> np.ndarray:
    chunks = []
    for p in params:
        g = p.grad if p.grad is not None else torch.zeros_like(p)
        chunks.append(g.detach().reshape(-1).float().cpu())
    return torch.cat(chunks).numpy() if chunks else np.zeros(0, dtype=np.float32)


def assign_flat_grads(params, flat: np.ndarray):
    t = torch.from_numpy(np.ascontiguousarray(flat, dtype=np.float32))
    off = 0
    for p in params:
        n = p.numel()
        g = t[off:off + n].view_as(p).to(device=p.device, dtype=p.dtype)
        p.grad = g.clone()
        off += n
    assert off == t.numel(), f"payload has {t.numel()} values, model needs {off}"


class COINNLearner:
    """Site side of an engine.

    ``batch_source()`` yields the next training batch (the node owns the loader cursor and
    the epoch boundary); ``to_reduce()`` runs ``local_iterations`` micro-batches and writes
    the payload; ``step()`` applies what the aggregator sent back.
    """

    def __init__(self, trainer=None, mp_pool=None, batch_source=None, **kw):
        self.trainer = trainer
        self.cache = trainer.cache
        self.state = trainer.state
        self.input = trainer.input
        self.mp_pool = mp_pool
        self.batch_source = batch_source
        self.params = trainable_params(trainer.nn)

    def rebind(self, input):
        self.input = input
        self.trainer.input = input

    def zero_grad(self):
        for p in self.params:
            p.grad = None

    @timed("learner_backward")
    def backward(self):
        self.trainer.train_mode()
        self.zero_grad()
        its = []
        for i in range(max(int(self.cache.get("local_iterations", 1)), 1)):
            batch = self.batch_source()
            if batch is None:
                break
            its.append(self.trainer.training_iteration_local(i, batch))
        return its

    def optimizer_step(self):
        for opt in self.trainer.optimizer.values():
            opt.step()

    def transfer_path(self, name):
        d = self.state.get("transferDirectory", ".")
        os.makedirs(d, exist_ok=True)
        return os.path.join(d, name)

    def incoming_path(self, name):
        return os.path.join(self.state.get("baseDirectory", "."), name)

    def step(self) -> dict:
        raise NotImplementedError

    def to_reduce(self):
        raise NotImplementedError


class COINNReducer:
    """Aggregator side of an engine."""

    def __init__(self, trainer=None, mp_pool=None, cache=None, input=None, state=None, **kw):
        self.trainer = trainer
        self.cache = cache if cache is not None else trainer.cache
        self.input = input if input is not None else trainer.input
        self.state = state if state is not None else trainer.state
        self.mp_pool = mp_pool

    def site_file(self, site, name):
        return os.path.join(self.state.get("baseDirectory", "."), str(site), name)

    def transfer_path(self, name):
        d = self.state.get("transferDirectory", ".")
        os.makedirs(d, exist_ok=True)
        return os.path.join(d, name)

    def reduce(self) -> dict:
        raise NotImplementedError


def get_engine(name):
    """``(LearnerCls, ReducerCls)`` for an engine string / enum."""
    from .dsgd import DSGDLearner, DSGDReducer
    from .rankdad import DADLearner, DADReducer
    from .powersgd import PowerSGDLearner, PowerSGDReducer
    key = getattr(name, "value", name)
    table = {
        AggEngine.DECENTRALIZED_SGD.value: (DSGDLearner, DSGDReducer),
        AggEngine.RANK_DAD.value: (DADLearner, DADReducer),
        AggEngine.POWER_SGD.value: (PowerSGDLearner, PowerSGDReducer),
    }
    if key not in table:
        raise ValueError(f"unknown agg_engine {name!r}; expected one of {list(table)}")
    return table[key]


__all__ = ["COINNLearner", "COINNReducer", "get_engine", "trainable_params", "flat_grads",
           "assign_flat_grads"]
