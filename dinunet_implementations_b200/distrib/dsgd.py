"""dSGD in file mode: average the raw gradients across sites.

Site: ``local_iterations`` micro-batches of accumulated grads → flat array → cast to
``precision_bits`` → ``.npy`` in the transfer dir.  Aggregator: unweighted mean over sites
(parallel over chunks) → ``avg_grads.npy``.  Site, next round: assign to ``p.grad`` and
``optimizer.step()`` (Adam, ``lr = learning_rate``).

The device-mode counterpart fuses mean + unscale + cast + Adam into one sm_100a kernel
over NVLink peer memory (``csrc/dsgd.cu``).
"""
from __future__ import annotations

from . import COINNLearner, COINNReducer, flat_grads, assign_flat_grads
from . import fileio
from ..utils.profiler import timed

GRADS_FILE = "grads.npy"
AVG_GRADS_FILE = "avg_grads.npy"


class DSGDLearner(COINNLearner):
    @timed("learner_step")
    def step(self):
        out = {}
        name = self.input.get("avg_grads_file")
        if name:
            assign_flat_grads(self.params, fileio.load_flat(self.incoming_path(name)))
            self.optimizer_step()
        return out

    @timed("learner_to_reduce")
    def to_reduce(self):
        its = self.backward()
        out = {}
        if its:
            out["grads_file"] = fileio.save_flat(self.transfer_path(GRADS_FILE), flat_grads(self.params),
                                                 self.cache.get("precision_bits", "32"))
            out["reduce"] = True
        return out, its


class DSGDReducer(COINNReducer):
    @timed("reducer_reduce")
    def reduce(self):
        sites = sorted(self.input.keys())
        paths = [self.site_file(s, self.input[s]["grads_file"]) for s in sites]
        avg = fileio.mean_flat_files(paths, self.mp_pool, self.cache.get("num_reducers"))
        name = fileio.save_flat(self.transfer_path(AVG_GRADS_FILE), avg,
                                self.cache.get("precision_bits", "32"))
        return {"avg_grads_file": name, "update": True}
