"""``SiteRunner`` — train one site's data with no federation.

Call sites: ``SiteRunner(taks_id='FSL', data_path='../../datasets/test_fsl', mode='Train',
split_ratio=[0.8, 0.1, 0.1])`` (``comps/fs/site_run.py:5``) and ``SiteRunner(taks_id='ICA',
data_path=…, mode='train', seed=10, site_index=0, split_ratio=…, monitor_metric='auc',
log_header='Loss|Auc', batch_size=32)`` (``comps/icalstm/site_run.py:6-8``), then
``.run(Trainer, Dataset, DataHandle)``.  The ``taks_id`` spelling is the reference's and is
accepted alongside ``task_id``.

``data_path`` is a simulator-layout directory: ``inputspec.json`` (one entry per site) and
``input/local<site_index>/simulatorRun`` with that site's files.  Outputs go to
``output/local<site_index>/simulatorRun`` under ``data_path`` when writable, else under
``out_dir`` / a temp dir.
"""
from __future__ import annotations

import json
import os
import tempfile

from . import config as cfg
from .config import Key
from .utils import unwrap_inputspec, save_json, logger, stringify_enum


class SiteRunner:
    def __init__(self, taks_id=None, data_path=".", mode="train", site_index=0, seed=None,
                 split_ratio=None, out_dir=None, task_id=None, **kw):
        self.task_id = stringify_enum(task_id if task_id is not None else taks_id)
        self.data_path = os.path.abspath(data_path)
        self.site_index = int(site_index)
        self.out_dir = out_dir
        self.cache = dict(cfg.DEFAULT_ARGS)
        self.cache.update({k: stringify_enum(v) for k, v in kw.items()})
        self.cache.update(task_id=self.task_id, mode=str(mode).lower(), fold=0)
        if seed is not None:
            self.cache["seed"] = int(seed)
        self._split_ratio = split_ratio
        self.dataloader_args = self.cache.pop("dataloader_args", None) or {}
        self.input, self.state = {}, {}

    def _load_spec(self):
        spec_path = os.path.join(self.data_path, "inputspec.json")
        if os.path.exists(spec_path):
            with open(spec_path) as f:
                spec = json.load(f)
            entry = spec[self.site_index] if isinstance(spec, list) else spec
            self.input = unwrap_inputspec(entry)
        targs = self.input.get(f"{self.input.get('task_id', self.task_id)}_args")
        if isinstance(targs, dict):
            self.cache.update(targs)
        self.cache.update({k: v for k, v in self.input.items() if not str(k).endswith("_args")})
        self.cache["task_id"] = self.task_id or self.cache.get("task_id")
        if self._split_ratio is not None:
            self.cache["split_ratio"], self.cache["num_folds"] = list(self._split_ratio), None
        elif not self.cache.get("num_folds") and not self.cache.get("split_ratio"):
            self.cache["split_ratio"] = [0.8, 0.1, 0.1]

    def _make_state(self):
        site = f"local{self.site_index}"
        base = os.path.join(self.data_path, "input", site, "simulatorRun")
        out = self.out_dir or os.path.join(self.data_path, "output", site, "simulatorRun")
        try:
            os.makedirs(out, exist_ok=True)
            if not os.access(out, os.W_OK):
                raise PermissionError(out)
        except OSError:
            out = tempfile.mkdtemp(prefix=f"dinunet_{site}_")
        self.state = {"baseDirectory": base, "outputDirectory": out, "clientId": site,
                      "cacheDirectory": out, "transferDirectory": out, "iteration": 0}

    def run(self, trainer_cls, dataset_cls, datahandle_cls):
        self._load_spec()
        self._make_state()
        handle = datahandle_cls(cache=self.cache, input=self.input, state=self.state,
                                dataloader_args=self.dataloader_args)
        handle.prepare_data()
        results = []
        for fold in range(len(self.cache["splits"])):
            self.cache["fold"] = fold
            trainer = trainer_cls(data_handle=handle, cache=self.cache, input=self.input, state=self.state)
            trainer.init_nn()
            res = {}
            if self.cache["mode"] == "train":
                train_ds = handle.get_train_dataset(dataset_cls)
                val_ds = handle.get_validation_dataset(dataset_cls)
                res = trainer.train_local(train_ds, val_ds,
                                          validation_epochs=self.cache.get("validation_epochs", 1), tag="site")
                best = os.path.join(trainer.log_dir(), "site_best.pt")
                if os.path.exists(best):
                    trainer.load_checkpoint(name="site_best.pt", load_optimizer_state=False)
                    trainer.save_checkpoint(name="best.pt")
            test_ds = handle.get_test_dataset(dataset_cls)
            avg, met = trainer.evaluation("test", [test_ds], save_pred=True)
            res[Key.TEST_METRICS] = [trainer.log_row(avg, met)]
            self.cache.update({k: v for k, v in res.items()})
            save_json(os.path.join(trainer.log_dir(), "logs.json"),
                      {k: v for k, v in self.cache.items() if not str(k).startswith("_")})
            logger.success(f"[site {self.site_index} fold {fold}] test {res[Key.TEST_METRICS]} "
                           f"best epoch {res.get('best_val_epoch')}", True)
            results.append(res)
        return results
