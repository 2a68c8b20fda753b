"""CPU stand-in for ``DeviceEngine``: same interface, plain PyTorch + ``torch.distributed`` (gloo).

It exists so that the *host-side* logic of a device-mode run — ``DeviceFederation``'s folds, epoch
boundaries with unequal site sizes, metric merging, early stopping, checkpoints, log layout — can be
exercised where there is no GPU (multi-process, ``world_size > 1``, gloo), and so that a machine without a
B200 can still run a federated job end to end.  It is not a performance path: one site per process, eager
model, gradients averaged with ``all_reduce`` and a ``torch.optim`` step — the same update rule the fused
dSGD kernel applies (unweighted mean over sites, then Adam/SGD; ``precision_bits 16`` rounds the
gradients to bf16 before they are summed, like the 16-bit wire format).
"""
from __future__ import annotations

import torch


def _dist():
    import torch.distributed as dist
    return dist if (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1) else None


class HostEngine:
    def __init__(self, trainer, use_graph: bool = False, track_metrics: bool = False, **_):
        self.trainer = trainer
        self.cache = trainer.cache
        self.device = trainer.device["gpu"]
        self.agg_engine = str(getattr(self.cache.get("agg_engine", "dSGD"), "value", self.cache.get("agg_engine", "dSGD")))
        if self.agg_engine != "dSGD":
            raise NotImplementedError(f"HostEngine implements dSGD only (got {self.agg_engine!r}); the low-rank engines "
                                      "run in device mode (CUDA) or in COINSTAC file mode (simulator)")
        self.params = [p for m in trainer.nn.values() for p in m.parameters() if p.requires_grad]
        lr = float(self.cache.get("learning_rate", 1e-3))
        if str(self.cache.get("optimizer", "adam")).lower() == "sgd":
            self.opt = torch.optim.SGD(self.params, lr=lr, momentum=float(self.cache.get("momentum", 0.0)))
        else:
            self.opt = torch.optim.Adam(self.params, lr=lr)
        self.half_wire = str(self.cache.get("precision_bits", "32")) == "16"
        self.steps = 0
        self.use_graph = False
        self._graph = None
        self.kernels_per_step = 0
        self.track = bool(track_metrics)
        self._avg, self._met = trainer.new_averages(), trainer.new_metrics()

    # ------------------------------------------------------------------ training
    def to_device(self, batch):
        return batch

    def _aggregate(self):
        d = _dist()
        flat = torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1) for p in self.params])
        if self.half_wire:
            flat = flat.to(torch.bfloat16).float()
        if d is not None:
            d.all_reduce(flat)
            flat /= d.get_world_size()
        off = 0
        for p in self.params:
            n = p.numel()
            p.grad = flat[off:off + n].view_as(p).clone()
            off += n

    def train_step(self, batch, micro_batches=None):
        self.trainer.train_mode()
        self.opt.zero_grad(set_to_none=True)
        mbs = micro_batches or [batch]
        its = []
        for mb in mbs:
            it = self.trainer.iteration(mb)
            (it["loss"] / len(mbs)).backward()
            its.append(it)
            if self.track:
                self._avg.accumulate(it["averages"])
                self._met.accumulate(it["metrics"])
        self._aggregate()
        self.opt.step()
        self.steps += 1
        return its[-1]

    def epoch_metrics(self, reset=True):
        out = (self._avg, self._met)
        if reset:
            self._avg, self._met = self.trainer.new_averages(), self.trainer.new_metrics()
        return out

    # ------------------------------------------------------------------ parameters / evaluation
    def sync_shadow(self):
        pass

    def load_checkpoint(self, name="best.pt", full_path=None):
        return self.trainer.load_checkpoint(name=name, full_path=full_path, load_optimizer_state=False)

    def evaluate(self, dataset, mode="validation"):
        avg, met = self.trainer.evaluation(mode, [dataset])
        d = _dist()
        if d is not None:
            payloads = [None] * d.get_world_size()
            d.all_gather_object(payloads, {"a": avg.serialize(), "m": met.serialize()})
            avg, met = self.trainer.new_averages(), self.trainer.new_metrics()
            avg.reduce_sites([p["a"] for p in payloads])
            met.reduce_sites([p["m"] for p in payloads])
        return avg, met

    def close(self):
        pass
