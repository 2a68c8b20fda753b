"""NN runtime: the ``COINNTrainer`` base class the task definitions subclass.

Hook contract from the reference's call sites: ``self.nn`` (dict of named
modules, ``comps/fs/__init__.py:47``), ``self.cache``, ``self.device['gpu']``
(``comps/fs/__init__.py:51``), ``_init_nn_model()``, ``iteration(batch) -> dict`` with a
``'loss'`` tensor plus ``'averages'`` / ``'metrics'`` objects, and the factories
``new_metrics()`` / ``new_averages()``.  Everything else in this file (seeded init,
optimiser, evaluation loop, best/early-stop, checkpoint I/O, the local training loop used
for pre-training and ``SiteRunner``) is capability E5/E7/E10 of SURVEY §2.3.
"""
from __future__ import annotations

import math
import os

import torch

from .. import config as cfg
from ..config import Key
from ..metrics import COINNAverages, Prf1a, ConfusionMatrix, AUCROCMetrics
from ..utils import set_seed, logger
from ..utils import checkpoint as ckpt


class NNTrainer:
    def __init__(self, data_handle=None, cache=None, input=None, state=None, **kw):
        self.cache = cache if cache is not None else {}
        self.input = input if input is not None else {}
        self.state = state if state is not None else {}
        self.data_handle = data_handle
        self.nn = {}
        self.optimizer = {}
        self.device = {"gpu": torch.device("cpu")}

    # ------------------------------------------------------------------ construction
    def init_nn(self, init_model=True, init_optim=True, set_devices=True, init_weights=True):
        if init_model:
            # identical starting point on every site: same seed right before construction
            set_seed(self.cache.get("seed", cfg.DEFAULT_ARGS["seed"]))
            self._init_nn_model()
        if init_weights:
            self._init_nn_weights()
        if set_devices:
            self._set_gpus()
            self._set_cpu_threads()
        if init_optim:
            self._init_optimizer()
        return self

    def _set_cpu_threads(self):
        """CPU sites with a small model: cap PyTorch's intra-op threads.  With the default (all cores) the
        61 k-parameter FreeSurfer net spends more time waking worker threads for 256-element kernels than
        computing (file-mode dSGD on an 8-core box: 35 rounds/s with 8 threads, 60 with 1-4).  ``cpu_threads``
        in the cache / inputspec: ``"auto"`` (default), an integer, or ``0`` to leave PyTorch alone; an explicit
        ``OMP_NUM_THREADS`` always wins."""
        if self.device["gpu"].type != "cpu" or os.environ.get("OMP_NUM_THREADS"):
            return
        want = self.cache.get("cpu_threads", "auto")
        if want in (0, "0", None, False):
            return
        if want == "auto":
            n_params = sum(p.numel() for m in self.nn.values() for p in m.parameters())
            if n_params >= 5_000_000 or torch.get_num_threads() <= 4:
                return
            want = 4
        try:
            torch.set_num_threads(max(1, int(want)))
        except (TypeError, ValueError, RuntimeError):
            pass

    def _init_nn_model(self):
        raise NotImplementedError("assign modules into self.nn['name'] here")

    def _init_nn_weights(self, **kw):
        """Default keeps torch's seeded default init; loads ``pretrained_path`` if given."""
        p = self.cache.get("pretrained_path")
        if p:
            if not os.path.isabs(p):
                p = os.path.join(self.state.get("baseDirectory", "."), p)
            if os.path.exists(p):
                self.load_checkpoint(full_path=p, load_optimizer_state=False)
                logger.info(f"loaded pretrained weights from {p}", self.cache.get("verbose"))

    def _set_gpus(self):
        gpus = self.cache.get("gpus") or []
        if torch.cuda.is_available() and len(gpus) > 0:
            self.device["gpu"] = torch.device(f"cuda:{int(gpus[0])}")
        else:
            self.device["gpu"] = torch.device("cpu")
        for name in list(self.nn.keys()):
            self.nn[name] = self.nn[name].to(self.device["gpu"])
            if self.device["gpu"].type == "cuda" and len(gpus) >= 2:
                # intra-site data parallelism, as in the reference engine.  DataParallel replicates
                # the module per device and reduces autograd gradients back; the fused kernels accumulate into
                # ``param.grad`` of whatever module object they are called on, so replicas take the eager path.
                for sub in self.nn[name].modules():
                    if hasattr(sub, "fused"):
                        sub.fused = False
                self.nn[name] = torch.nn.DataParallel(self.nn[name], device_ids=[int(g) for g in gpus])

    def _init_optimizer(self):
        # one optimiser over the parameters of EVERY module in self.nn: the engines flatten, exchange and assign
        # gradients for all of them (distrib.trainable_params), so all of them have to be stepped
        params = [p for m in self.nn.values() for p in m.parameters()]
        lr = float(self.cache.get("learning_rate", cfg.DEFAULT_ARGS["learning_rate"]))
        kind = str(self.cache.get("optimizer", "adam")).lower()
        if kind == "sgd":
            self.optimizer["sgd"] = torch.optim.SGD(params, lr=lr,
                                                    momentum=float(self.cache.get("momentum", 0.0)))
        else:
            self.optimizer["adam"] = torch.optim.Adam(params, lr=lr)

    # ------------------------------------------------------------------ hooks
    def iteration(self, batch) -> dict:
        raise NotImplementedError

    def new_metrics(self):
        mm = str(self.cache.get("monitor_metric", "f1")).lower()
        if mm == "auc":
            return AUCROCMetrics()
        if int(self.cache.get("num_class", 2)) > 2:
            return ConfusionMatrix(num_classes=int(self.cache["num_class"]))
        return Prf1a()

    def new_averages(self):
        return COINNAverages(num_averages=1)

    def save_predictions(self, dataset, its):
        pass

    def _on_iteration_end(self, **kw):
        pass

    def _on_epoch_end(self, **kw):
        pass

    # ------------------------------------------------------------------ train / eval steps
    def train_mode(self):
        for m in self.nn.values():
            m.train()

    def eval_mode(self):
        for m in self.nn.values():
            m.eval()

    def training_iteration_local(self, i, batch):
        """One micro-batch: forward + backward with the loss scaled for accumulation."""
        it = self.iteration(batch)
        k = max(int(self.cache.get("local_iterations", 1)), 1)
        (it["loss"] / k).backward()
        return it

    def reduce_iteration(self, its):
        out = {"averages": self.new_averages(), "metrics": self.new_metrics()}
        for it in its:
            if it.get("averages") is not None:
                out["averages"].accumulate(it["averages"])
            if it.get("metrics") is not None:
                out["metrics"].accumulate(it["metrics"])
        return out

    def evaluation(self, mode="validation", dataset_list=None, save_pred=False):
        self.eval_mode()
        averages, metrics = self.new_averages(), self.new_metrics()
        with torch.no_grad():
            for ds in dataset_list or []:
                if ds is None or len(ds) == 0:
                    continue
                tbl = getattr(ds, "_device_table", None)
                if tbl is not None:
                    # device-resident split: consecutive slices of the table, same batching as the loader below
                    bs, n = int(self.cache.get("batch_size", 16)), len(tbl)
                    if n > 1 and n % bs == 1:
                        bs += 1
                    loader = (tbl.slice(lo, min(lo + bs, n)) for lo in range(0, n, bs))
                else:
                    loader = self.data_handle.get_loader(handle_key=mode, dataset=ds, shuffle=False)
                its = []
                for batch in loader:
                    it = self.iteration(batch)
                    averages.accumulate(it["averages"])
                    metrics.accumulate(it["metrics"])
                    if save_pred:
                        its.append(it)
                if save_pred:
                    self.save_predictions(ds, its)
        return averages, metrics

    # ------------------------------------------------------------------ scores / logs
    def score_of(self, averages, metrics):
        mm = str(self.cache.get("monitor_metric", "f1")).lower()
        if mm in ("loss", "average"):
            return averages.average
        try:
            return float(metrics.extract(mm))
        except AttributeError:
            return float(metrics.get()[0])

    def log_row(self, averages, metrics):
        return list(averages.get()) + list(metrics.get())

    # ------------------------------------------------------------------ checkpoints
    def log_dir(self):
        d = os.path.join(self.state.get("outputDirectory", "."), str(self.cache.get("task_id", "task")),
                         f"fold_{self.cache.get('fold', 0)}")
        os.makedirs(d, exist_ok=True)
        return d

    def save_checkpoint(self, name="latest.pt", full_path=None, **extra):
        path = full_path or os.path.join(self.log_dir(), name)
        return ckpt.save_checkpoint(path, self.nn, self.optimizer, **extra)

    def load_checkpoint(self, name="best.pt", full_path=None, load_optimizer_state=True):
        path = full_path or os.path.join(self.log_dir(), name)
        return ckpt.load_checkpoint(path, self.nn, self.optimizer, map_location="cpu",
                                    load_optimizer_state=load_optimizer_state)

    # ------------------------------------------------------------------ plain local loop
    def train_local(self, train_dataset, val_dataset, epochs=None, patience=None, lr=None,
                    validation_epochs=1, local_iterations=None, tag="local"):
        """Single-site training without a remote (pre-training, ``SiteRunner``).

        Returns ``{'best_val_epoch', 'best_val_score', 'train_log', 'validation_log'}``.
        """
        epochs = int(epochs if epochs is not None else self.cache.get("epochs", 1))
        patience = int(patience if patience is not None else self.cache.get("patience", epochs))
        k = max(int(local_iterations or self.cache.get("local_iterations", 1)), 1)
        if lr is not None:
            for opt in self.optimizer.values():
                for g in opt.param_groups:
                    g["lr"] = float(lr)
        direction = self.cache.get("metric_direction", "maximize")
        best_score, best_epoch = None, 0
        train_log, val_log = [], []
        opt = next(iter(self.optimizer.values()))
        for ep in range(1, epochs + 1):
            self.cache["epoch"] = ep
            self.train_mode()
            loader = self.data_handle.get_loader(handle_key="train", dataset=train_dataset)
            ep_avg, ep_met = self.new_averages(), self.new_metrics()
            opt.zero_grad(set_to_none=True)
            for i, batch in enumerate(loader, 1):
                it = self.iteration(batch)
                (it["loss"] / k).backward()
                ep_avg.accumulate(it["averages"])
                ep_met.accumulate(it["metrics"])
                if i % k == 0:
                    opt.step()
                    opt.zero_grad(set_to_none=True)
                self._on_iteration_end(i=i, epoch=ep, it=it)
            train_log.append(self.log_row(ep_avg, ep_met))
            if ep % max(int(validation_epochs), 1) == 0 and val_dataset is not None and len(val_dataset):
                va, vm = self.evaluation("validation", [val_dataset])
                val_log.append(self.log_row(va, vm))
                score = self.score_of(va, vm)
                if cfg.metric_is_better(score, best_score, direction):
                    best_score, best_epoch = score, ep
                    self.save_checkpoint(name=f"{tag}_best.pt")
                logger.info(f"[{tag}] epoch {ep}: train {train_log[-1]} val {val_log[-1]}",
                            self.cache.get("verbose"))
                if ep - best_epoch >= patience:
                    break
            self._on_epoch_end(epoch=ep)
        if best_epoch == 0:
            self.save_checkpoint(name=f"{tag}_best.pt")
        return {"best_val_epoch": best_epoch, "best_val_score": best_score,
                Key.TRAIN_LOG: train_log, Key.VALIDATION_LOG: val_log}


class COINNTrainer(NNTrainer):
    """Name used by the reference's task definitions (``comps/fs/__init__.py:42``)."""
    pass
