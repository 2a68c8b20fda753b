"""``COINNLocal`` — the site side of the round state machine.

Call-site contract (``local.py:29-49``): constructed every round with ``task_id,
agg_engine, cache, input, state`` plus arbitrary default kwargs (``batch_size, epochs,
patience, split_ratio, pretrain_args, dataloader_args, num_class, monitor_metric,
log_header`` …) that must land in the cache; ``local.cache['task_id']`` is readable right
after construction and may have been overridden by ``input`` (the GUI field); calling
``local(mp_pool, TrainerCls, DatasetCls, DataHandleCls)`` returns the dict that goes back
to COINSTAC; ``cache['num_reducers']`` is set after the first call (``local.py:26``).

Precedence, lowest first: built-in defaults < constructor kwargs <
``<task_id>_args`` object < first-round ``input`` < what the aggregator echoes back.

The Python process is long-lived (module-global ``CACHE`` in ``local.py:15``), so live
objects (trainer, datasets, loader iterator, learner) are parked in a process-level
registry keyed by a token stored in the cache; the cache itself stays JSON-serialisable.
If the registry entry is gone (process restarted) the site rebuilds from its ``latest``
checkpoint — a crash-resume path the reference does not have.
"""
from __future__ import annotations

import os
import shutil
import uuid

from .. import config as cfg
from ..config import Phase, Mode, Key
from ..distrib import get_engine
from ..utils import logger, save_json, stringify_enum

_LIVE = {}

_RESERVED = {"cache", "input", "state"}


class _Live:
    """Per-site, per-fold live objects."""

    def __init__(self):
        self.trainer = None
        self.handle = None
        self.learner = None
        self.train_ds = self.val_ds = self.test_ds = None
        self.loader_iter = None
        self.batches_per_epoch = 0


class COINNLocal:
    def __init__(self, task_id=None, agg_engine=None, cache=None, input=None, state=None, **kw):
        self.cache = cache if cache is not None else {}
        self.input = dict(input or {})
        self.state = state if state is not None else {}
        self.out = {}
        first = not self.cache.get(Key.ARGS_CACHED)
        if first:
            merged = dict(cfg.DEFAULT_ARGS)
            merged["task_id"] = stringify_enum(task_id)
            merged["agg_engine"] = stringify_enum(agg_engine) or cfg.DEFAULT_ARGS["agg_engine"]
            merged.update({k: stringify_enum(v) for k, v in kw.items() if k not in _RESERVED})
            inp = {k: v for k, v in self.input.items()}
            chosen_task = inp.get("task_id", merged["task_id"])
            for src in (kw, inp):  # "<task_id>_args" objects are flattened in (compspec.json:225-281)
                targs = src.get(f"{chosen_task}_args")
                if isinstance(targs, dict):
                    merged.update(targs)
            merged.update({k: v for k, v in inp.items() if not str(k).endswith("_args") or k == "pretrain_args"})
            if merged.get("num_folds"):
                merged["split_ratio"] = None
            elif not merged.get("split_ratio") and not merged.get("split_files"):
                merged["split_ratio"] = [0.8, 0.1, 0.1]
            self._dataloader_args = merged.pop("dataloader_args", None) or {}
            self.cache.update(merged)
            self.cache["dataloader_args"] = self._dataloader_args
            self.cache[Key.ARGS_CACHED] = True
            self.cache["live_token"] = uuid.uuid4().hex
        self._dataloader_args = self.cache.get("dataloader_args") or {}

    # ------------------------------------------------------------------ properties
    @property
    def site_id(self):
        return str(self.state.get("clientId", "local0"))

    @property
    def live(self) -> _Live:
        tok = self.cache.setdefault("live_token", uuid.uuid4().hex)
        return _LIVE.setdefault(tok, _Live())

    def _verbose(self):
        return bool(self.cache.get("verbose"))

    # ------------------------------------------------------------------ entry
    def __call__(self, mp_pool=None, trainer_cls=None, dataset_cls=None, datahandle_cls=None, **kw):
        self._classes = (trainer_cls, dataset_cls, datahandle_cls)
        self.mp_pool = mp_pool
        phase = self.input.get("phase", Phase.INIT_RUNS.value)
        self.out["phase"] = phase

        if phase == Phase.INIT_RUNS.value:
            self._init_runs()
        elif phase == Phase.NEXT_RUN.value:
            self._absorb_shared(self.input.get("shared_args"))
            self._next_run()
        elif phase == Phase.PRE_COMPUTATION.value:
            self._pre_computation()
            phase = Phase.COMPUTATION.value
        if phase == Phase.COMPUTATION.value:
            self._computation()
        elif phase == Phase.SUCCESS.value:
            self._finish()

        self.out["site_id"] = self.site_id
        return {"output": self.out}

    # ------------------------------------------------------------------ phases
    def _make_handle(self):
        trainer_cls, dataset_cls, datahandle_cls = self._classes
        return datahandle_cls(cache=self.cache, input=self.input, state=self.state,
                              dataloader_args=self._dataloader_args)

    def _init_runs(self):
        handle = self._make_handle()
        self.cache["data_size"] = len(handle.list_files())
        self.out["data_size"] = self.cache["data_size"]
        self.out["shared_args"] = {k: self.cache.get(k) for k in cfg.SHARED_ARGS}
        self.out["phase"] = Phase.INIT_RUNS.value

    def _absorb_shared(self, shared):
        if shared:
            self.cache.update(shared)

    def _next_run(self):
        """Aggregator assigned ``fold`` and ``seed``: build splits, model, datasets."""
        trainer_cls, dataset_cls, datahandle_cls = self._classes
        self.cache["fold"] = int(self.input.get("fold", 0))
        self.cache["seed"] = int(self.input.get("seed", self.cache.get("seed", 0)))
        self.cache.update(epoch=0, cursor=0, best_val_epoch=0, best_val_score=None)
        self.cache[Key.TRAIN_LOG], self.cache[Key.VALIDATION_LOG], self.cache[Key.TEST_METRICS] = [], [], []

        tok = self.cache["live_token"]
        _LIVE[tok] = lv = _Live()
        lv.handle = self._make_handle()
        if self.cache["fold"] == 0 or not self.cache.get("splits"):
            lv.handle.prepare_data()
        lv.trainer = trainer_cls(data_handle=lv.handle, cache=self.cache, input=self.input, state=self.state)
        lv.trainer.init_nn()
        mode = str(self.cache.get("mode", "train")).lower()
        if mode == "train":
            lv.train_ds = lv.handle.get_train_dataset(dataset_cls)
            lv.val_ds = lv.handle.get_validation_dataset(dataset_cls)
        lv.test_ds = lv.handle.get_test_dataset(dataset_cls)
        self.out["phase"] = Phase.PRE_COMPUTATION.value
        self.out["fold"] = self.cache["fold"]

        pre = self.cache.get("pretrain_args") or {}
        if (mode == "train" and self.cache.get("pretrain") and int(pre.get("epochs", 0)) > 0
                and self.input.get("pretrain_site") == self.site_id):
            self._pretrain(lv, pre)

    def _pretrain(self, lv, pre):
        """This site holds the most data: train alone first, ship the weights (E10)."""
        logger.info(f"{self.site_id}: pre-training for {pre.get('epochs')} epochs", self._verbose())
        saved = {k: self.cache.get(k) for k in ("batch_size", "local_iterations", "learning_rate")}
        self.cache.update({k: pre[k] for k in ("batch_size", "local_iterations") if k in pre})
        res = lv.trainer.train_local(lv.train_ds, lv.val_ds, epochs=pre.get("epochs"),
                                     patience=pre.get("patience"), lr=pre.get("learning_rate"),
                                     validation_epochs=pre.get("validation_epochs", 1), tag="pretrain")
        self.cache.update(saved)
        lv.trainer.load_checkpoint(name="pretrain_best.pt", load_optimizer_state=False)
        name = f"pretrained_{self.site_id}.pt"
        tdir = self.state.get("transferDirectory", ".")
        os.makedirs(tdir, exist_ok=True)
        lv.trainer.save_checkpoint(full_path=os.path.join(tdir, name))
        self.out["pretrained_weights"] = name
        self.cache["pretrain_log"] = {k: res[k] for k in ("best_val_epoch", "best_val_score")}
        self.cache["epoch"] = 0

    def _pre_computation(self):
        lv = self.live
        name = self.input.get("pretrained_weights")
        if name and lv.trainer is not None:
            lv.trainer.load_checkpoint(full_path=os.path.join(self.state.get("baseDirectory", "."), name),
                                       load_optimizer_state=False)
            lv.trainer._init_optimizer()  # fresh optimiser state for the federated phase
        learner_cls, _ = get_engine(self.cache["agg_engine"])
        lv.learner = learner_cls(trainer=lv.trainer, mp_pool=self.mp_pool, batch_source=self._next_batch)
        self.cache["mode"] = str(self.cache.get("mode", "train")).lower()

    # ------------------------------------------------------------------ batches
    def _next_batch(self):
        lv = self.live
        if lv.train_ds is None or len(lv.train_ds) == 0:
            return None
        if lv.loader_iter is None:
            loader = lv.handle.get_loader(handle_key="train", dataset=lv.train_ds)
            lv.batches_per_epoch = max(len(loader), 1)
            lv.loader_iter = iter(loader)
        try:
            batch = next(lv.loader_iter)
        except StopIteration:
            # wrap around: a site that is waiting for slower sites keeps contributing
            self.cache["pass"] = int(self.cache.get("pass", 0)) + 1
            loader = lv.handle.get_loader(handle_key="train", dataset=lv.train_ds)
            lv.loader_iter = iter(loader)
            batch = next(lv.loader_iter)
        self.cache["cursor"] = int(self.cache.get("cursor", 0)) + 1
        return batch

    # ------------------------------------------------------------------ computation
    def _computation(self):
        lv = self.live
        if lv.trainer is None:
            self._resume()
            lv = self.live
        lv.learner.rebind(self.input)
        self.out["phase"] = Phase.COMPUTATION.value
        modes = self.input.get("global_modes") or {}
        default_mode = Mode.TEST.value if self.cache.get("mode") == "test" else Mode.TRAIN.value
        mode = modes.get(self.site_id, default_mode)

        if self.input.get("save_current_as_best"):
            lv.trainer.save_checkpoint(name="best.pt")
            self.cache["best_val_epoch"] = self.input.get("best_val_epoch", self.cache.get("epoch"))
            self.cache["best_val_score"] = self.input.get("best_val_score")

        if self.input.get("update"):
            lv.learner.step()

        if mode in (Mode.TRAIN.value, Mode.VALIDATION_WAITING.value):
            self._train_round(lv, mode)
        elif mode == Mode.VALIDATION.value:
            self._validation_round(lv)
        elif mode == Mode.TEST.value:
            self._test_round(lv)
        else:
            self.out["mode"] = mode

    def _train_round(self, lv, mode):
        out, its = lv.learner.to_reduce()
        self.out.update(out)
        if its:
            red = lv.trainer.reduce_iteration(its)
            self.out[Key.TRAIN_SERIALIZABLE] = {"averages": red["averages"].serialize(),
                                                "metrics": red["metrics"].serialize()}
        self.out["mode"] = mode
        if mode == Mode.TRAIN.value and self.cache["cursor"] >= lv.batches_per_epoch:
            self.cache["cursor"] = 0
            self.cache["epoch"] = int(self.cache.get("epoch", 0)) + 1
            lv.trainer._on_epoch_end(epoch=self.cache["epoch"])
            if self.cache["epoch"] % max(int(self.cache.get("validation_epochs", 1)), 1) == 0:
                self.out["mode"] = Mode.VALIDATION_WAITING.value
        if int(self.state.get("iteration", 0)) % 50 == 0:
            lv.trainer.save_checkpoint(name="latest.pt")

    def _validation_round(self, lv):
        avg, met = lv.trainer.evaluation("validation", [lv.val_ds])
        self.out[Key.VALIDATION_SERIALIZABLE] = {"averages": avg.serialize(), "metrics": met.serialize()}
        self.cache[Key.VALIDATION_LOG].append(lv.trainer.log_row(avg, met))
        self.out["mode"] = Mode.TRAIN_WAITING.value
        lv.trainer.save_checkpoint(name="latest.pt")
        lv.loader_iter = None  # next epoch reshuffles
        self.cache["cursor"] = 0

    def _test_round(self, lv):
        best = os.path.join(lv.trainer.log_dir(), "best.pt")
        if os.path.exists(best):
            lv.trainer.load_checkpoint(name="best.pt", load_optimizer_state=False)
        avg, met = lv.trainer.evaluation("test", [lv.test_ds], save_pred=True)
        self.out[Key.TEST_SERIALIZABLE] = {"averages": avg.serialize(), "metrics": met.serialize()}
        self.cache[Key.TEST_METRICS] = [lv.trainer.log_row(avg, met)]
        self.out["mode"] = Mode.TEST_WAITING.value
        self.out["phase"] = Phase.NEXT_RUN_WAITING.value
        self._save_logs(lv)

    def _save_logs(self, lv):
        if not self.cache.get("persist_logs", True):
            return
        keep = {k: v for k, v in self.cache.items() if k not in ("dataloader_args",)}
        save_json(os.path.join(lv.trainer.log_dir(), "logs.json"), keep)
        try:
            from ..utils.plotter import plot_progress
            plot_progress(self.cache, lv.trainer.log_dir(), name=f"{self.site_id}")
        except Exception as e:  # plotting must never break a run
            logger.warn(f"plot skipped: {e}", self._verbose())

    def _finish(self):
        """SUCCESS: pull the aggregator's global-results zip into our output dir."""
        name = self.input.get("results_zip")
        if name:
            src = os.path.join(self.state.get("baseDirectory", "."), name)
            if os.path.exists(src):
                os.makedirs(self.state.get("outputDirectory", "."), exist_ok=True)
                shutil.copy(src, os.path.join(self.state.get("outputDirectory", "."), name))
        lv = self.live
        if lv.trainer is not None:
            self._save_logs(lv)
        self.out["phase"] = Phase.SUCCESS.value
        _LIVE.pop(self.cache.get("live_token"), None)

    # ------------------------------------------------------------------ crash-resume
    def _resume(self):
        """Live objects lost (new process): rebuild the fold from cache + ``latest.pt``."""
        trainer_cls, dataset_cls, datahandle_cls = self._classes
        lv = self.live
        lv.handle = self._make_handle()
        lv.trainer = trainer_cls(data_handle=lv.handle, cache=self.cache, input=self.input, state=self.state)
        lv.trainer.init_nn()
        latest = os.path.join(lv.trainer.log_dir(), "latest.pt")
        if os.path.exists(latest):
            lv.trainer.load_checkpoint(name="latest.pt")
        lv.train_ds = lv.handle.get_train_dataset(dataset_cls)
        lv.val_ds = lv.handle.get_validation_dataset(dataset_cls)
        lv.test_ds = lv.handle.get_test_dataset(dataset_cls)
        learner_cls, _ = get_engine(self.cache["agg_engine"])
        lv.learner = learner_cls(trainer=lv.trainer, mp_pool=self.mp_pool, batch_source=self._next_batch)
        logger.warn(f"{self.site_id}: resumed fold {self.cache.get('fold')} from checkpoint", True)
