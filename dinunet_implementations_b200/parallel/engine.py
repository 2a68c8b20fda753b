"""In-process device mode: every GPU of the box is a site (one process per GPU).

``DeviceEngine`` wraps a task's ``COINNTrainer`` and replaces the per-round
JSON + file + aggregator round trip of the COINSTAC path (2 network hops,
4 file I/Os and 2 PCIe copies per optimiser step) with device buffers: the model's parameters
and gradients live in a ``SiteGroup`` arena that all sites have mapped, and one fused kernel
per step does barrier → cross-site mean → unscale → optimiser → parameter push.

The trainer hooks are the same ones the file mode uses (``iteration(batch)`` etc.), so a task
written for COINSTAC runs here unchanged.  Public API::

    engine = DeviceEngine(trainer)             # after trainer.init_nn()
    it = engine.train_step(batch)              # batch: dict of host (pinned) or device tensors
    avg, met = engine.evaluate(dataset)        # metrics merged across sites

Launch with ``torchrun --nproc-per-node N``; with a single process it runs one site (or S
simulated sites on one GPU through ``SiteGroup.local`` for tests).
"""
from __future__ import annotations

import os

import torch

from ..distrib import trainable_params
from .comm import SiteGroup
from .dsgd_fused import FusedDSGD, wire_dtype_from_bits


def init_distributed_from_env():
    """``(rank, world, local_rank)``; initialises NCCL when launched by torchrun."""
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if torch.cuda.is_available():
        torch.cuda.set_device(local_rank)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29512")
        dist.init_process_group(backend="nccl" if torch.cuda.is_available() else "gloo",
                                device_id=torch.device(f"cuda:{local_rank}") if torch.cuda.is_available() else None)
    return rank, world, local_rank


class DeviceEngine:
    def __init__(self, trainer, group: SiteGroup | None = None, arena_bytes: int | None = None, site: int | None = None,
                 shadow_bf16: bool = False):
        self.trainer = trainer
        self.cache = trainer.cache
        self.device = trainer.device["gpu"]
        if self.device.type != "cuda":
            raise RuntimeError("DeviceEngine needs the trainer on a CUDA device (cache['gpus'] = [i])")
        self.params = trainable_params(trainer.nn)
        numel = sum(p.numel() for p in self.params)
        wire = wire_dtype_from_bits(self.cache.get("precision_bits", "32"))
        need = numel * (4 + wire.itemsize + (2 if shadow_bf16 else 0)) + (1 << 20)
        if group is None:
            import torch.distributed as dist
            ab = int(arena_bytes or need)
            if dist.is_available() and dist.is_initialized():
                group = SiteGroup.from_torch_distributed(ab, self.device)
            else:
                group = SiteGroup.local(1, ab, self.device)
        self.group = group
        self.site = group.rank if site is None else site
        opt = str(self.cache.get("optimizer", "adam")).lower()
        self.fused = FusedDSGD(group, numel, wire_dtype=wire, lr=float(self.cache.get("learning_rate", 1e-3)),
                               optimizer=opt, momentum=float(self.cache.get("momentum", 0.0)), shadow_bf16=shadow_bf16)
        holder = torch.nn.ModuleList(list(trainer.nn.values()))
        self.fused.bind(holder, site=self.site)
        self.local_iterations = max(int(self.cache.get("local_iterations", 1)), 1)
        self.steps = 0

    # ------------------------------------------------------------------ training
    def to_device(self, batch):
        return {k: (v.to(self.device, non_blocking=True) if isinstance(v, torch.Tensor) else v)
                for k, v in batch.items()}

    def train_step(self, batch, micro_batches=None):
        """One optimiser step: ``local_iterations`` micro-batches, then the fused aggregation."""
        self.trainer.train_mode()
        self.fused.zero_grad(site=self.site)
        its = []
        for mb in (micro_batches or [batch]):
            it = self.trainer.iteration(self.to_device(mb))
            (it["loss"] / len(micro_batches or [batch])).backward()
            its.append(it)
        self.fused.step()
        self.steps += 1
        return its[-1] if len(its) == 1 else self.trainer.reduce_iteration(its) | {"loss": its[-1]["loss"]}

    # ------------------------------------------------------------------ evaluation
    def evaluate(self, dataset, mode="validation"):
        """Local evaluation, then merge the sufficient statistics across sites."""
        avg, met = self.trainer.evaluation(mode, [dataset])
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            payloads = [None] * dist.get_world_size()
            dist.all_gather_object(payloads, {"a": avg.serialize(), "m": met.serialize()})
            avg, met = self.trainer.new_averages(), self.trainer.new_metrics()
            avg.reduce_sites([p["a"] for p in payloads])
            met.reduce_sites([p["m"] for p in payloads])
        return avg, met


def make_ica_trainer(device_index=0, batch_size=16, hidden_size=384, input_size=256, num_components=100,
                     window_size=10, temporal_size=980, precision_bits="32", fused=True, seed=1, lr=1e-3):
    """ICA-LSTM trainer with the BASELINE.json config-2 shapes, ready for ``DeviceEngine``."""
    from ..tasks.icalstm import ICATrainer
    cache = {"task_id": "ICA-Classification", "window_size": window_size, "window_stride": window_size,
             "temporal_size": temporal_size, "num_components": num_components, "input_size": input_size,
             "hidden_size": hidden_size, "num_class": 2, "monitor_metric": "auc", "log_header": "loss|auc",
             "batch_size": batch_size, "learning_rate": lr, "gpus": [device_index], "seed": seed,
             "precision_bits": precision_bits, "fused": fused, "agg_engine": "dSGD"}
    tr = ICATrainer(cache=cache, input={}, state={})
    tr.init_nn(init_optim=False)
    return tr


def smoke_step():
    """Tiny ICA-LSTM forward + backward + fused dSGD on cuda:0; returns the loss."""
    tr = make_ica_trainer(0, batch_size=4, hidden_size=64, input_size=32, num_components=8, window_size=5,
                          temporal_size=50)
    eng = DeviceEngine(tr)
    g = torch.Generator().manual_seed(0)
    batch = {"inputs": torch.randn(4, 10, 8, 5, generator=g), "labels": torch.randint(0, 2, (4,), generator=g)}
    loss = None
    for _ in range(2):
        loss = float(eng.train_step(batch)["loss"].item())
    eng.group.close()
    return loss
