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

import contextlib
import ctypes
import os

import torch

from ..distrib import trainable_params
from .comm import SiteGroup
from .dsgd_fused import FusedDSGD, wire_dtype_from_bits, flat_layout, ordered_params


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
    """See module docstring.  ``use_graph=True`` captures the whole step (zero-grad, forward,
    backward, fused aggregation) in one CUDA graph after ``graph_warmup`` eager steps; every later
    step is a single graph launch.  Requires an ``iteration`` without host syncs (no ``.item()``),
    otherwise the engine silently keeps stepping eagerly."""

    def __init__(self, trainer, group: SiteGroup | None = None, arena_bytes: int | None = None, site: int | None = None,
                 shadow_bf16: bool = True, use_graph: bool = False, graph_warmup: int = 3,
                 track_metrics: bool = False, metric_capacity: int = 1 << 20):
        self.trainer = trainer
        self.cache = trainer.cache
        self.device = trainer.device["gpu"]
        if self.device.type != "cuda":
            raise RuntimeError("DeviceEngine needs the trainer on a CUDA device (cache['gpus'] = [i])")
        self.params = []
        for m in trainer.nn.values():
            self.params += ordered_params(m)
        numel = flat_layout(self.params)[1]
        wire = wire_dtype_from_bits(self.cache.get("precision_bits", "32"))
        need = numel * (4 + wire.itemsize + (2 if shadow_bf16 else 0)) * 3 + (8 << 20)
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
        lr_ = float(self.cache.get("learning_rate", 1e-3))
        mom = float(self.cache.get("momentum", 0.0))
        self.agg_engine = str(getattr(self.cache.get("agg_engine", "dSGD"), "value", self.cache.get("agg_engine", "dSGD")))
        holder = torch.nn.ModuleList(list(trainer.nn.values()))
        if self.agg_engine == "dSGD":
            self.fused = FusedDSGD(group, numel, wire_dtype=wire, lr=lr_, optimizer=opt, momentum=mom,
                                   shadow_bf16=shadow_bf16, device_state=True)
            self.fused.bind(holder, site=self.site, params=self.params)
        elif self.agg_engine == "powerSGD":
            from .lowrank_fused import DevicePowerSGD
            self.fused = DevicePowerSGD(group, [self.params], rank=int(self.cache.get("powersgd_rank", 4)), lr=lr_,
                                        optimizer=opt, momentum=mom, seed=int(self.cache.get("seed", 0)))
            # both exchange phases are device-side (reduce kernel with device-resident epochs) and all state
            # (error feedback, warm-started Q) is updated in place: the step can be captured
        elif self.agg_engine == "rankDAD":
            from ..distrib.rankdad import make_taps, CapturedTap
            from .lowrank_fused import DeviceRankDAD
            # device mode reconstructs the weights only; every bias travels dense (exact, a few KB)
            taps, _ = make_taps(trainer.nn.values())
            covered = {id(t.module.weight) for t in taps}
            dense = [p for p in self.params if id(p) not in covered]
            self.fused = DeviceRankDAD(group, [self.params], [taps], [dense],
                                       rank=int(self.cache.get("dad_reduction_rank", 10)),
                                       num_pow_iters=int(self.cache.get("dad_num_pow_iters", 5)),
                                       tol=float(self.cache.get("dad_tol", 1e-3)), lr=lr_, optimizer=opt, momentum=mom,
                                       seed=int(self.cache.get("seed", 0)),
                                       nsplit=1 if str(self.cache.get("precision_bits", "32")) == "16" else 3)
            # the device rankDAD step has no host syncs and no per-step host state (fixed iteration count,
            # cached seeded start matrices, device-resident exchange epochs): it can live in a CUDA graph
            # when every tapped Linear hands its pair over from the fused backward (module hooks cannot)
            use_graph = use_graph and all(isinstance(t, CapturedTap) for t in taps)
        else:
            raise ValueError(f"unknown agg_engine {self.agg_engine!r}")
        if self.agg_engine != "dSGD":
            shadow = getattr(self.fused.opt[0].holder, "_bf16_shadow", None)
            if shadow is not None:
                for m in trainer.nn.values():
                    for sub in m.modules():
                        sub._bf16_shadow = shadow
        # models whose fused backward can store (not accumulate) their parameter gradients: skip the zero-fill of those
        # spans and the atomic read-modify-write in the kernels.  Only with one micro-batch per aggregation.
        self._overwrite = []
        if int(self.cache.get("local_iterations", 1) or 1) == 1 and self.agg_engine in ("dSGD", "powerSGD"):
            for m in trainer.nn.values():
                if getattr(m, "fused", False) and hasattr(m, "grad_overwrite_params"):
                    m._grad_overwrite = True
                    self._overwrite += list(m.grad_overwrite_params())
            if self._overwrite:
                for f in ([self.fused] if self.agg_engine == "dSGD" else [o.f for o in self.fused.opt]):
                    f.set_overwritten(self._overwrite)
        self.steps = 0
        self._one = None
        self.timer = None
        if self.cache.get("profile_enabled"):
            from ..utils.profiler import PhaseTimer
            self.timer = PhaseTimer(self.device)
        self.train_metrics = None
        if track_metrics:
            from ..ops.metrics import DeviceMetricBuffer
            self.train_metrics = DeviceMetricBuffer(self.device, capacity=int(metric_capacity))
        # CUDA graph state
        self.use_graph = bool(use_graph)
        self.graph_warmup = int(graph_warmup)
        self._graph = None
        self.graph_replays = 0           # counted, not derived: bench.py reports launches from it
        self.kernels_per_step = None
        self._static_in = None
        self._static_out = None
        # async pipeline state (train_step_async / drain)
        self._copy_stream = None
        self._staging = None
        self._loss_ring = None
        self._fast = None
        self._pending = []
        self._batcher = self._rgraph = self._rout = None

    # ------------------------------------------------------------------ training
    def to_device(self, batch):
        return {k: (v.to(self.device, non_blocking=True) if isinstance(v, torch.Tensor) else v)
                for k, v in batch.items()}

    def _ph(self, name):
        """Phase timer context (utils/profiler.py); a no-op unless profiling, and never inside a capture."""
        t = self.timer
        if t is None or not t.enabled or torch.cuda.is_current_stream_capturing():
            return contextlib.nullcontext()
        return t.phase(name)

    def _step_body(self, batch):
        self.trainer.train_mode()
        with self._ph("zero_grad"):
            self.fused.zero_grad(site=self.site)
        with self._ph("forward"):
            it = self.trainer.iteration(batch)
        with self._ph("backward"):
            loss = it["loss"]
            if self._one is None or self._one.dtype != loss.dtype or self._one.shape != loss.shape:
                self._one = torch.ones_like(loss)          # persistent seed gradient: no fill launch per step
                from ..ops.elementwise import UNIT_GRADS
                UNIT_GRADS.add(self._one.data_ptr())
            loss.backward(gradient=self._one)
        with self._ph("aggregate+update"):
            self.fused.step()
        if self.train_metrics is not None and it.get("out") is not None and "labels" in batch:
            # K9: scores / labels / loss of this step go to device accumulators (inside the graph too)
            self.train_metrics.append(it["out"], batch["labels"], it["loss"])
        return it

    def profile_steps(self, batches, warmup=3):
        """Eager (kernel-by-kernel) steps over ``batches`` with CUDA-event phase timing; returns
        ``{phase: {'ms_mean','ms_max','calls'}}`` with MAX over ranks.  Advances training like
        ``train_step``; the captured graph, if any, is left untouched."""
        from ..utils.profiler import PhaseTimer
        keep = self.timer
        self.timer = PhaseTimer(self.device, enabled=False)
        for i, b in enumerate(batches):
            if i == warmup:
                self.timer.enabled = True
            with self._ph("step"):
                self._step_body(self.to_device(b))
            self.steps += 1
        out = self.timer.summary(reduce=True)
        self.timer = keep
        return out

    def epoch_metrics(self, reset=True):
        """``(averages, metrics)`` of the training steps since the last call — one host sync."""
        if self.train_metrics is None:
            return self.trainer.new_averages(), self.trainer.new_metrics()
        avg, auc, prf = self.train_metrics.collect(reset)
        want_auc = str(self.cache.get("monitor_metric", "auc")).lower() == "auc"
        return avg, (auc if want_auc else prf)

    def _capture(self, dev_batch):
        self._static_in = {k: (v.clone() if isinstance(v, torch.Tensor) else v) for k, v in dev_batch.items()}
        try:
            from ..ops import native
            g = torch.cuda.CUDAGraph()
            n0 = native.launch_count()
            with torch.cuda.graph(g):
                it = self._step_body(self._static_in)
            self.kernels_per_step = native.launch_count() - n0     # our kernels inside one replay
            self._graph, self._static_out = g, it
        except Exception as e:   # host sync inside iteration(), unsupported op, ...
            from ..utils import logger
            logger.warn(f"CUDA graph capture failed ({type(e).__name__}: {e}); stepping eagerly", True)
            torch.cuda.synchronize()
            self.use_graph = False
            self._graph = None
            return False
        return True

    def _load_static(self, dev_batch):
        for k, v in dev_batch.items():
            if isinstance(v, torch.Tensor):
                self._static_in[k].copy_(v, non_blocking=True)

    def train_step(self, batch, micro_batches=None):
        """One optimiser step on ``batch`` (dict of host or device tensors): forward, backward,
        cross-site aggregation, update.  Returns the trainer's ``iteration`` dict; with a captured
        graph its tensors are static buffers that the next step overwrites."""
        if micro_batches:
            return self._accumulate_step(micro_batches)
        dev_batch = self.to_device(batch)
        self.steps += 1
        if self.use_graph and self._graph is None and self.steps > self.graph_warmup:
            if self._capture(dev_batch):
                self._graph.replay()      # capture records but does not execute: run the captured step once
                self.graph_replays += 1
                return self._static_out
        if self._graph is not None:
            self._load_static(dev_batch)
            self._graph.replay()
            self.graph_replays += 1
            return self._static_out
        return self._step_body(dev_batch)

    # ------------------------------------------------------------------ device-resident stepping
    def attach_batcher(self, batcher):
        """``batcher``: ``data.device.DeviceBatcher`` over this site's training split.  ``train_step_resident`` then
        steps without touching the host's data path: the batch gather is one more kernel in front of the forward
        pass (inside the captured graph), and no bytes cross PCIe per step."""
        self._batcher = batcher
        self._rgraph = self._rout = None

    def train_step_resident(self):
        b = self._batcher
        b.tick()                                  # host bookkeeping only: reshuffles at pass boundaries
        self.steps += 1
        if self.use_graph and self._rgraph is None and self.steps > self.graph_warmup:
            try:
                from ..ops import native
                g = torch.cuda.CUDAGraph()
                n0 = native.launch_count()
                with torch.cuda.graph(g):
                    it = self._step_body(b.launch())
                self.kernels_per_step = native.launch_count() - n0
                self._rgraph, self._rout = g, it
            except Exception as e:      # noqa: BLE001
                from ..utils import logger
                logger.warn(f"CUDA graph capture of the device-resident step failed ({type(e).__name__}: {e}); "
                            "stepping kernel by kernel", True)
                torch.cuda.synchronize()
                self.use_graph = False
        if self._rgraph is not None:
            self._rgraph.replay()
            self.graph_replays += 1
            return self._rout
        return self._step_body(b.launch())

    def _accumulate_step(self, micro_batches):
        """``local_iterations`` > 1: accumulate the micro-batch gradients, then aggregate once."""
        if self._overwrite:                      # accumulation needs accumulating kernels and a full zero-fill
            for m in self.trainer.nn.values():
                m._grad_overwrite = False
            for f in ([self.fused] if self.agg_engine == "dSGD" else [o.f for o in self.fused.opt]):
                f._zero_spans = None
            self._overwrite = []
        self.trainer.train_mode()
        self.fused.zero_grad(site=self.site)
        its = []
        for mb in micro_batches:
            it = self.trainer.iteration(self.to_device(mb))
            (it["loss"] / len(micro_batches)).backward()
            its.append(it)
        self.fused.step()
        self.steps += 1
        red = self.trainer.reduce_iteration(its)
        red["loss"] = its[-1]["loss"]
        return red

    # ------------------------------------------------------------------ pipelined stepping
    def train_step_async(self, host_batch):
        """Pipelined variant for host-resident (pinned) batches: the H2D copy of this batch runs on
        a copy stream while the previous step still computes, and the loss comes back through a
        pinned ring without stalling the launch thread.  Call ``drain()`` to collect the losses.

        Host cost matters here: one GPU step is ~0.3 ms, and the obvious implementation (``.to()`` under a
        stream context, fresh events, ``copy_`` per tensor) spent 0.34 ms of host time per step — the
        pipeline was launch-bound.  Once the graph exists, a step is therefore a handful of raw async copies
        through the native library into preallocated ping-pong buffers, two event records/waits on
        persistent events, and the graph replay."""
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
            self._staging = [None, None]
            self._loss_ring = torch.zeros(64, dtype=torch.float32).pin_memory()
        if self._graph is not None and self._fast_ok(host_batch):
            return self._fast_async(host_batch)
        slot = self.steps & 1
        cur = torch.cuda.current_stream(self.device)
        with torch.cuda.stream(self._copy_stream):
            if self._staging[slot] is not None:
                self._copy_stream.wait_event(self._staging[slot][2])     # the step that used this slot consumed it
            dev = {k: (v.to(self.device, non_blocking=True) if isinstance(v, torch.Tensor) else v)
                   for k, v in host_batch.items()}
            ready = torch.cuda.Event()
            ready.record(self._copy_stream)
        cur.wait_event(ready)
        it = self.train_step(dev)
        consumed = torch.cuda.Event()
        consumed.record(cur)
        self._staging[slot] = (dev, ready, consumed)
        ring_ix = len(self._pending) % self._loss_ring.numel()
        self._loss_ring[ring_ix:ring_ix + 1].copy_(it["loss"].detach().reshape(1), non_blocking=True)
        self._pending.append(ring_ix)
        if len(self._pending) == self._loss_ring.numel():
            return self.drain()
        return None

    def _fast_ok(self, host_batch):
        si = self._static_in
        for k, v in host_batch.items():
            if isinstance(v, torch.Tensor):
                t = si.get(k)
                if (not isinstance(t, torch.Tensor) or v.is_cuda or not v.is_pinned() or not v.is_contiguous()
                        or v.dtype != t.dtype or v.shape != t.shape):
                    return False
        return self._static_out is not None and self._static_out["loss"].dtype == torch.float32

    def _fast_async(self, host_batch):
        from ..ops import native
        lib = native.lib()
        fs = self._fast
        if fs is None:
            keys = [k for k, v in host_batch.items() if isinstance(v, torch.Tensor)]
            fs = self._fast = {
                "keys": keys,
                "buf": [{k: torch.empty_like(self._static_in[k]) for k in keys} for _ in range(2)],
                "ready": [torch.cuda.Event(), torch.cuda.Event()],
                "consumed": [torch.cuda.Event(), torch.cuda.Event()],
                "used": [False, False],
                "nbytes": {k: self._static_in[k].numel() * self._static_in[k].element_size() for k in keys},
                "static_ptr": {k: self._static_in[k].data_ptr() for k in keys},
                "loss_ptr": self._static_out["loss"].data_ptr(),
                "ring_ptr": self._loss_ring.data_ptr(),
                "copy": lib.dnb_copy_async,
            }
            fs["copy"].argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
        self.steps += 1
        slot = self.steps & 1
        cs = self._copy_stream
        cur = torch.cuda.current_stream(self.device)
        cs_h, cur_h = cs.cuda_stream, cur.cuda_stream
        copy, buf = fs["copy"], fs["buf"][slot]
        if fs["used"][slot]:
            cs.wait_event(fs["consumed"][slot])          # the step that read this slot has consumed it
        for k in fs["keys"]:
            copy(buf[k].data_ptr(), host_batch[k].data_ptr(), fs["nbytes"][k], 1, cs_h)
        fs["ready"][slot].record(cs)
        cur.wait_event(fs["ready"][slot])
        for k in fs["keys"]:
            copy(fs["static_ptr"][k], buf[k].data_ptr(), fs["nbytes"][k], 3, cur_h)
        fs["consumed"][slot].record(cur)                   # staging slot free again once these copies have run
        self._graph.replay()
        self.graph_replays += 1
        fs["used"][slot] = True
        fs.setdefault("host_ref", [None, None])[slot] = host_batch   # keep the pinned source alive while its DMA may run
        ring_ix = len(self._pending) % 64
        copy(fs["ring_ptr"] + 4 * ring_ix, fs["loss_ptr"], 4, 2, cur_h)
        self._pending.append(ring_ix)
        if len(self._pending) == 64:
            return self.drain()
        return None

    def drain(self):
        """Wait for the in-flight steps and return their losses (floats, in submission order)."""
        torch.cuda.current_stream(self.device).synchronize()
        out = [float(self._loss_ring[i]) for i in self._pending]
        self._pending = []
        return out

    # ------------------------------------------------------------------ parameter I/O
    def sync_shadow(self):
        """Call after writing parameters in place (``load_state_dict``, checkpoint load, broadcast):
        the kernels read bf16 copies of the weights, which are otherwise refreshed only by the fused
        optimiser step."""
        if self.agg_engine == "dSGD":
            self.fused.sync_shadow(self.site)
        else:
            for o in self.fused.opt:
                o.f.sync_shadow()

    def load_checkpoint(self, name="best.pt", full_path=None):
        """Load model weights (reference checkpoint format) into the live flat buffers."""
        out = self.trainer.load_checkpoint(name=name, full_path=full_path, load_optimizer_state=False)
        self.sync_shadow()
        return out

    def close(self):
        """Release the graph and the peer-mapped arenas (collective in multi-process runs)."""
        import torch.distributed as dist
        self._graph = self._static_in = self._static_out = None
        self._rgraph = self._rout = None
        torch.cuda.synchronize(self.device)
        multi = self.group.backend in ("ipc", "symm") and dist.is_available() and dist.is_initialized()
        self.group.close(barrier=dist.barrier if multi else None)
        for o in getattr(self.fused, "opt", []) or []:
            if hasattr(o, "group"):
                o.group.close()

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
