"""``COINNRemote`` — the aggregator side of the round state machine.

Call-site contract (``remote.py:23-37``): ``COINNRemote(cache, input, state)`` where
``input`` maps ``site_id → that site's output dict``; ``remote.cache['task_id']`` is
populated from what the sites sent (the remote has no constructor argument for it,
``remote.py:28``); ``remote(mp_pool, TrainerCls)`` returns ``{'output': …, 'success': …}``.
The Trainer class is only needed to build the metrics/averages objects that merge the
sites' serialised scores (it never sees a dataset).

Control decisions made here: fold / seed assignment, which site pre-trains, when every
site switches train → validation → train/test, best-checkpoint flag, early stop on
``patience`` (``compspec.json:193-204``), next fold or SUCCESS, result files
(``logs.json``, ``test_metrics.csv``, results zip — layout in SURVEY §2.2).
"""
from __future__ import annotations

import os
import shutil
import zipfile

from .. import config as cfg
from ..config import Phase, Mode, Key
from ..data import splits as _splits
from ..distrib import get_engine
from ..utils import logger, save_json


def write_test_metrics_csv(path, fold, cols, row, met):
    """One header line + one row per fold; columns = the ``log_header`` names, then whichever of
    accuracy / f1 / precision / recall the metrics object also offers."""
    cols, extra = list(cols), []
    for nm in ("accuracy", "f1", "precision", "recall"):
        if nm not in [c.lower() for c in cols] and hasattr(met, nm):
            cols.append(nm)
            extra.append(round(float(getattr(met, nm)), 5))
    with open(path, "w") as f:
        f.write(",".join(["fold"] + cols[:len(row) + len(extra)]) + "\n")
        f.write(",".join([f"fold_{fold}"] + [str(v) for v in list(row) + extra]) + "\n")


class COINNRemote:
    def __init__(self, cache=None, input=None, state=None, **kw):
        self.cache = cache if cache is not None else {}
        self.input = dict(input or {})
        self.state = state if state is not None else {}
        self.out = {}
        if not self.cache.get(Key.ARGS_CACHED):
            shared = None
            for site in sorted(self.input.keys()):
                shared = (self.input[site] or {}).get("shared_args")
                if shared:
                    break
            if shared:
                self.cache.update(dict(cfg.DEFAULT_ARGS))
                self.cache.update(shared)
                self.cache.update({k: v for k, v in kw.items()})
                self.cache[Key.ARGS_CACHED] = True
                self.cache["all_sites"] = sorted(self.input.keys())

    # ------------------------------------------------------------------ helpers
    @property
    def sites(self):
        return sorted(self.input.keys())

    def _all(self, key, value):
        return all((self.input[s] or {}).get(key) == value for s in self.sites)

    def _log_dir(self):
        d = os.path.join(self.state.get("outputDirectory", "."), str(self.cache.get("task_id", "task")),
                         f"fold_{self.cache.get('fold', 0)}")
        os.makedirs(d, exist_ok=True)
        return d

    def _trainer(self, trainer_cls):
        return trainer_cls(cache=self.cache, input=self.input, state=self.state)

    # ------------------------------------------------------------------ entry
    def __call__(self, mp_pool=None, trainer_cls=None, **kw):
        self.mp_pool = mp_pool
        self.trainer_cls = trainer_cls
        phases = {(self.input[s] or {}).get("phase") for s in self.sites}
        success = False

        if phases == {Phase.INIT_RUNS.value}:
            self._init_runs()
        elif phases == {Phase.PRE_COMPUTATION.value}:
            self._pre_computation()
        elif phases == {Phase.COMPUTATION.value}:
            self._computation()
        elif phases == {Phase.NEXT_RUN_WAITING.value}:
            success = self._next_run_or_finish()
        elif phases == {Phase.SUCCESS.value}:
            success = True
            self.out["phase"] = Phase.SUCCESS.value
        else:
            raise RuntimeError(f"sites are in inconsistent phases: {phases}")
        return {"output": self.out, "success": success}

    # ------------------------------------------------------------------ phases
    def _init_runs(self):
        sizes = {s: int(self.input[s].get("data_size", 0)) for s in self.sites}
        self.cache["data_sizes"] = sizes
        self.cache["num_runs"] = _splits.num_runs(self.cache.get("split_ratio"), self.cache.get("num_folds"),
                                                  self.cache.get("split_files"))
        self.cache["fold"] = 0
        self.out["shared_args"] = {k: self.cache.get(k) for k in cfg.SHARED_ARGS}
        self._start_fold()

    def _start_fold(self):
        self.cache.update(epoch=0, best_val_epoch=0, best_val_score=None)
        self.cache[Key.TRAIN_LOG], self.cache[Key.VALIDATION_LOG] = [], []
        self.cache[Key.TEST_METRICS] = []
        self.cache["_train_scores"] = None
        self.out["phase"] = Phase.NEXT_RUN.value
        self.out["fold"] = self.cache["fold"]
        self.out["seed"] = int(self.cache.get("seed", 0)) + int(self.cache["fold"])
        pre = self.cache.get("pretrain_args") or {}
        if self.cache.get("pretrain") and int(pre.get("epochs", 0)) > 0:
            sizes = self.cache["data_sizes"]
            self.out["pretrain_site"] = max(sorted(sizes), key=lambda s: sizes[s])

    def _pre_computation(self):
        self.out["phase"] = Phase.PRE_COMPUTATION.value
        for s in self.sites:
            name = self.input[s].get("pretrained_weights")
            if name:
                src = os.path.join(self.state.get("baseDirectory", "."), s, name)
                tdir = self.state.get("transferDirectory", ".")
                os.makedirs(tdir, exist_ok=True)
                shutil.copy(src, os.path.join(tdir, name))
                self.out["pretrained_weights"] = name
                break
        mode = Mode.TEST.value if str(self.cache.get("mode")).lower() == "test" else Mode.TRAIN.value
        self.out["global_modes"] = {s: mode for s in self.sites}

    def _computation(self):
        self.out["phase"] = Phase.COMPUTATION.value
        modes = {s: self.input[s].get("mode") for s in self.sites}

        # 1. gradient / factor aggregation
        if all(self.input[s].get("reduce") for s in self.sites):
            _, reducer_cls = get_engine(self.cache["agg_engine"])
            reducer = reducer_cls(trainer=None, mp_pool=self.mp_pool, cache=self.cache, input=self.input,
                                  state=self.state)
            self.out.update(reducer.reduce())

        # 2. running train scores
        self._accumulate_train_scores()

        # 3. mode transitions
        vals = set(modes.values())
        if vals == {Mode.VALIDATION_WAITING.value}:
            self.out["global_modes"] = {s: Mode.VALIDATION.value for s in self.sites}
        elif vals == {Mode.TRAIN_WAITING.value}:
            self._on_validation_scores()
        elif vals == {Mode.TEST_WAITING.value}:
            self.out["global_modes"] = {s: Mode.TEST_WAITING.value for s in self.sites}
        else:
            self.out["global_modes"] = {s: (m if m in (Mode.TRAIN.value, Mode.VALIDATION_WAITING.value)
                                            else Mode.TRAIN.value) for s, m in modes.items()}

    def _accumulate_train_scores(self):
        payloads = [self.input[s].get(Key.TRAIN_SERIALIZABLE) for s in self.sites]
        payloads = [p for p in payloads if p]
        if not payloads:
            return
        tr = self._trainer(self.trainer_cls)
        avg, met = tr.new_averages(), tr.new_metrics()
        prev = self.cache.get("_train_scores")
        if prev:
            avg.reduce_sites([prev["averages"]])
            met.reduce_sites([prev["metrics"]])
        avg.reduce_sites([p["averages"] for p in payloads])
        met.reduce_sites([p["metrics"] for p in payloads])
        self.cache["_train_scores"] = {"averages": avg.serialize(), "metrics": met.serialize()}

    def _on_validation_scores(self):
        tr = self._trainer(self.trainer_cls)
        avg, met = tr.new_averages(), tr.new_metrics()
        payloads = [self.input[s][Key.VALIDATION_SERIALIZABLE] for s in self.sites]
        avg.reduce_sites([p["averages"] for p in payloads])
        met.reduce_sites([p["metrics"] for p in payloads])
        self.cache["epoch"] = int(self.cache.get("epoch", 0)) + max(int(self.cache.get("validation_epochs", 1)), 1)
        self.cache[Key.VALIDATION_LOG].append(tr.log_row(avg, met))

        prev = self.cache.get("_train_scores")
        if prev:
            tavg, tmet = tr.new_averages(), tr.new_metrics()
            tavg.reduce_sites([prev["averages"]])
            tmet.reduce_sites([prev["metrics"]])
            self.cache[Key.TRAIN_LOG].append(tr.log_row(tavg, tmet))
            self.cache["_train_scores"] = None

        score = tr.score_of(avg, met)
        if cfg.metric_is_better(score, self.cache.get("best_val_score"),
                                self.cache.get("metric_direction", "maximize")):
            self.cache["best_val_score"] = score
            self.cache["best_val_epoch"] = self.cache["epoch"]
            self.out["save_current_as_best"] = True
            self.out["best_val_epoch"] = self.cache["epoch"]
            self.out["best_val_score"] = score
        logger.info(f"epoch {self.cache['epoch']}: val {self.cache[Key.VALIDATION_LOG][-1]} "
                    f"best@{self.cache['best_val_epoch']}", bool(self.cache.get("verbose")))

        done = self.cache["epoch"] >= int(self.cache.get("epochs", 1))
        stale = self.cache["epoch"] - int(self.cache.get("best_val_epoch", 0)) >= int(self.cache.get("patience", 1 << 30))
        nxt = Mode.TEST.value if (done or stale) else Mode.TRAIN.value
        self.out["global_modes"] = {s: nxt for s in self.sites}

    def _next_run_or_finish(self):
        tr = self._trainer(self.trainer_cls)
        avg, met = tr.new_averages(), tr.new_metrics()
        payloads = [self.input[s][Key.TEST_SERIALIZABLE] for s in self.sites]
        avg.reduce_sites([p["averages"] for p in payloads])
        met.reduce_sites([p["metrics"] for p in payloads])
        self.cache[Key.TEST_METRICS] = [tr.log_row(avg, met)]
        self.cache.setdefault(Key.GLOBAL_TEST_SERIALIZABLE, []).append(
            {"averages": avg.serialize(), "metrics": met.serialize()})
        self._save_fold_results(tr, avg, met)

        if int(self.cache["fold"]) + 1 < int(self.cache.get("num_runs", 1)):
            self.cache["fold"] = int(self.cache["fold"]) + 1
            self._start_fold()
            return False
        return self._finish(tr)

    # ------------------------------------------------------------------ results
    def _header_cols(self):
        return [c.strip() for grp in str(self.cache.get("log_header", "loss")).split("|") for c in grp.split(",")]

    def _save_fold_results(self, tr, avg, met):
        d = self._log_dir()
        keep = {k: v for k, v in self.cache.items() if not str(k).startswith("_")}
        save_json(os.path.join(d, "logs.json"), keep)
        write_test_metrics_csv(os.path.join(d, "test_metrics.csv"), self.cache["fold"], self._header_cols(),
                               tr.log_row(avg, met), met)
        try:
            from ..utils.plotter import plot_progress
            plot_progress(self.cache, d, name="global")
        except Exception:
            pass

    def _finish(self, tr):
        """Global score over all folds + the results zip the notebooks read."""
        gavg, gmet = tr.new_averages(), tr.new_metrics()
        ser = self.cache.get(Key.GLOBAL_TEST_SERIALIZABLE, [])
        gavg.reduce_sites([p["averages"] for p in ser])
        gmet.reduce_sites([p["metrics"] for p in ser])
        self.cache[Key.GLOBAL_TEST_METRICS] = [tr.log_row(gavg, gmet)]
        root = os.path.join(self.state.get("outputDirectory", "."), str(self.cache.get("task_id", "task")))
        os.makedirs(root, exist_ok=True)
        save_json(os.path.join(root, "global_test_metrics.json"),
                  {"log_header": self.cache.get("log_header"), Key.GLOBAL_TEST_METRICS: self.cache[Key.GLOBAL_TEST_METRICS]})
        zip_name = f"{self.cache.get('task_id', 'task')}_global_results.zip"
        zpath = os.path.join(self.state.get("outputDirectory", "."), zip_name)
        with zipfile.ZipFile(zpath, "w", zipfile.ZIP_DEFLATED) as z:
            for base, _, files in os.walk(root):
                for fn in files:
                    full = os.path.join(base, fn)
                    z.write(full, os.path.relpath(full, root))   # root of the zip holds fold_<k>/…
        tdir = self.state.get("transferDirectory", ".")
        os.makedirs(tdir, exist_ok=True)
        shutil.copy(zpath, os.path.join(tdir, zip_name))
        self.out["results_zip"] = zip_name
        self.out["phase"] = Phase.SUCCESS.value
        self.out[Key.GLOBAL_TEST_METRICS] = self.cache[Key.GLOBAL_TEST_METRICS]
        return False   # one more round so sites can pull the zip; they answer SUCCESS
