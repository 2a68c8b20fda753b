"""CPU stand-in for ``DeviceEngine``: same interface, plain PyTorch + ``torch.distributed`` (gloo).

It exists so that the *host-side* logic of a device-mode run — ``DeviceFederation``'s folds, epoch
boundaries with unequal site sizes, metric merging, early stopping, checkpoints, log layout — can be
exercised where there is no GPU (multi-process, ``world_size > 1``, gloo), and so that a machine without a
B200 can still run a federated job end to end.  It is not a performance path: one site per process, eager
model, gradients averaged with ``all_reduce`` and a ``torch.optim`` step — the same update rule the fused
dSGD kernel applies (unweighted mean over sites, then Adam/SGD; ``precision_bits 16`` rounds the
gradients to bf16 before they are summed, like the 16-bit wire format).  rankDAD and powerSGD use the
reference math of ``distrib/lowrank.py`` with ``all_gather`` / ``all_reduce`` in place of the aggregator:
rankDAD gathers every site's rank-r factors and re-compresses them redundantly on every rank; powerSGD
averages P and Q with error feedback and a warm-started, identically seeded Q.
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
        if self.agg_engine not in ("dSGD", "rankDAD", "powerSGD"):
            raise ValueError(f"unknown agg_engine {self.agg_engine!r}")
        self.params = [p for m in trainer.nn.values() for p in m.parameters() if p.requires_grad]
        self.seed = int(self.cache.get("seed", 0))
        if self.agg_engine == "rankDAD":
            from ..distrib.rankdad import _LinearTap
            for m in trainer.nn.values():
                if hasattr(m, "fused"):
                    m.fused = False                    # hooks need the eager nn.Linear calls
            self.taps = [_LinearTap(mod) for m in trainer.nn.values() for mod in m.modules()
                         if isinstance(mod, torch.nn.Linear) and mod.weight.requires_grad]
            covered = {id(p) for t in self.taps for p in (t.module.weight, t.module.bias) if p is not None}
            self.dense = [p for p in self.params if id(p) not in covered]
        elif self.agg_engine == "powerSGD":
            from ..distrib import lowrank
            r = int(self.cache.get("powersgd_rank", 4))
            self.mats = [p for p in self.params if p.dim() >= 2]
            self.vecs = [p for p in self.params if p.dim() < 2]
            self.Q = [lowrank.psgd_init_q((p.shape[0], p.numel() // p.shape[0]), max(1, min(r, p.shape[0], p.numel() // p.shape[0])),
                                          self.seed + 17 * i, p.device) for i, p in enumerate(self.mats)]
            self.E = [torch.zeros(p.shape[0], p.numel() // p.shape[0], device=p.device) for p in self.mats]
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

    def _mean_(self, t):
        d = _dist()
        if d is not None:
            d.all_reduce(t)
            t /= d.get_world_size()
        return t

    def _aggregate_rankdad(self):
        from ..distrib import lowrank
        d = _dist()
        world = d.get_world_size() if d is not None else 1
        r, iters = int(self.cache.get("dad_reduction_rank", 10)), int(self.cache.get("dad_num_pow_iters", 5))
        tol = float(self.cache.get("dad_tol", 1e-3))
        for k, t in enumerate(self.taps):
            D, A = t.pop()
            w = t.module.weight
            n_in = w.shape[1] + (1 if t.module.bias is not None else 0)
            if D is None:
                D, A = torch.zeros(1, w.shape[0]), torch.zeros(1, n_in)
            dk, ak, _ = lowrank.dad_compress(D, A, r, iters, tol, self.seed + k)
            if d is not None:
                ds, as_ = [torch.empty_like(dk) for _ in range(world)], [torch.empty_like(ak) for _ in range(world)]
                d.all_gather(ds, dk.contiguous())
                d.all_gather(as_, ak.contiguous())
            else:
                ds, as_ = [dk], [ak]
            dr, ar, _ = lowrank.dad_reduce(ds, as_, r, iters, tol, self.seed + 101 * k, average=True)
            G = lowrank.dad_reconstruct(dr, ar)
            if t.module.bias is not None:
                t.module.bias.grad = G[:, -1].contiguous()
                G = G[:, :-1]
            w.grad = G.contiguous().view_as(w)
        for p in self.dense:
            if p.grad is None:
                p.grad = torch.zeros_like(p)
            self._mean_(p.grad)

    def _aggregate_powersgd(self):
        from ..distrib import lowrank
        for i, p in enumerate(self.mats):
            g = p.grad if p.grad is not None else torch.zeros_like(p)
            M = g.view(p.shape[0], -1) + self.E[i]
            P = lowrank._orth_fast(self._mean_(lowrank.psgd_p(M, self.Q[i])))
            Qn = self._mean_(lowrank.psgd_q(M, P))
            Mh = P @ Qn.t()
            self.E[i] = M - Mh
            self.Q[i] = Qn
            p.grad = Mh.view_as(p).contiguous()
        for v in self.vecs:
            if v.grad is None:
                v.grad = torch.zeros_like(v)
            self._mean_(v.grad)

    def _aggregate(self):
        if self.agg_engine == "rankDAD":
            return self._aggregate_rankdad()
        if self.agg_engine == "powerSGD":
            return self._aggregate_powersgd()
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
