"""Device-mode federation: the whole COINSTAC run (folds, epochs, validation, early stopping, best
checkpoint, test, logs) with one process per GPU and every GPU acting as one site.

File mode (``nodes/local.py`` + ``nodes/remote.py``) moves a JSON dict and gradient files through
an aggregator once per optimiser step.  Here the same *decisions* are taken, but nothing leaves
the GPUs on the step path:

* one optimiser step = one replay of a CUDA graph that contains forward, backward and the fused
  cross-site aggregation + update kernel (``DeviceEngine``); peers' gradients are read over
  NVLink from inside that kernel;
* the control plane (what the remote decides in file mode: fold, epoch boundary, "save current as
  best", early stop, test) needs no messages at all: validation/test statistics are merged across
  sites as sufficient statistics (``engine.evaluate``), so every rank sees the same aggregated
  score and takes the same decision deterministically;
* a site with fewer batches than the largest site wraps around its loader so that every step has a
  gradient from every site (same rule as our file mode).

Outputs follow the reference layout: ``<work>/output/local<i>/simulatorRun/<task>/
fold_<k>/{logs.json,best.pt,latest.pt}``; rank 0 also writes the aggregator's view
``<work>/output/remote/simulatorRun/<task>/fold_<k>/{logs.json,test_metrics.csv}`` and the
``<task>_global_results.zip`` that every site receives.

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 -m dinunet_implementations_b200.parallel.federation \\
        --task ICA-Classification --data-path datasets/icalstm --work-dir /tmp/run [--agg-engine dSGD] [--set k=json]
"""
from __future__ import annotations

import json
import os
import shutil
import time
import zipfile

import torch

from .. import config as cfg
from ..config import Key
from ..utils import unwrap_inputspec, save_json, logger, stringify_enum, clean_json
from .engine import DeviceEngine, init_distributed_from_env

# what the reference's local.py passes to COINNLocal (local.py:29-37)
COMPUTATION_DEFAULTS = dict(batch_size=16, epochs=21, patience=31, split_ratio=[0.7, 0.15, 0.15], pretrain_args=None,
                            num_class=2, monitor_metric="auc", log_header="loss|auc")
DATALOADER_ARGS = {"train": {"drop_last": True}}


def _gather(obj):
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        out = [None] * dist.get_world_size()
        dist.all_gather_object(out, obj)
        return out
    return [obj]


def _barrier():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()


class _Cycle:
    """Endless batches from a dataset: a new (reshuffled) loader whenever one is exhausted."""

    def __init__(self, handle, dataset, cache):
        self.handle, self.dataset, self.cache = handle, dataset, cache
        self.passes = 0
        self._it = None

    def __len__(self):
        return len(self.handle.get_loader(handle_key="train", dataset=self.dataset))

    def next(self):
        for _ in range(2):
            if self._it is None:
                self.cache["_pass"] = self.passes
                g = torch.Generator()
                g.manual_seed(int(self.cache.get("seed", 0)) + 7919 * self.passes + 1)
                self._it = iter(self.handle.get_loader(handle_key="train", dataset=self.dataset, generator=g))
                self.passes += 1
            try:
                return next(self._it)
            except StopIteration:
                self._it = None
        raise RuntimeError("training split yields no batch (fewer samples than batch_size with drop_last?)")


class DeviceFederation:
    def __init__(self, task_id, data_path, work_dir, agg_engine=None, overrides=None, use_graph=True, verbose=False):
        self.rank, self.world, self.local_rank = init_distributed_from_env()
        self.data_path, self.work_dir = os.path.abspath(data_path), os.path.abspath(work_dir)
        self.use_graph, self.verbose = bool(use_graph), bool(verbose)
        self.t0 = time.time()
        cache = dict(cfg.DEFAULT_ARGS)
        cache.update(COMPUTATION_DEFAULTS)
        spec_path = os.path.join(self.data_path, "inputspec.json")
        self.input = {}
        if os.path.exists(spec_path):
            with open(spec_path) as f:
                spec = json.load(f)
            spec = spec if isinstance(spec, list) else [spec]
            self.input = unwrap_inputspec(spec[self.rank % len(spec)])
        self.input.update(overrides or {})
        task_id = stringify_enum(task_id or self.input.get("task_id"))
        targs = self.input.get(f"{task_id}_args")
        if isinstance(targs, dict):
            cache.update(targs)
        cache.update({k: stringify_enum(v) for k, v in self.input.items() if not str(k).endswith("_args")})
        cache["task_id"] = task_id
        if agg_engine is not None:
            cache["agg_engine"] = stringify_enum(agg_engine)
        # one site per GPU; without CUDA the same run goes through HostEngine (eager + gloo), see host_engine.py
        cache["gpus"] = [self.local_rank] if torch.cuda.is_available() else []
        if cache.get("num_folds"):
            cache["split_ratio"] = None
        self.cache = cache
        site = f"local{self.rank}"
        base = os.path.join(self.data_path, "input", site, "simulatorRun")
        out = os.path.join(self.work_dir, "output", site, "simulatorRun")
        os.makedirs(out, exist_ok=True)
        self.state = {"baseDirectory": base, "outputDirectory": out, "clientId": site, "cacheDirectory": out,
                      "transferDirectory": out, "iteration": 0}
        self.remote_out = os.path.join(self.work_dir, "output", "remote", "simulatorRun")

    # ------------------------------------------------------------------ pieces
    def _log(self, msg):
        if self.rank == 0:
            logger.info(msg, self.verbose)

    def _pretrain(self, trainer_cls, dataset_cls, handle):
        """E10: the site with the most data trains alone; its weights start everyone (compspec 120-148)."""
        pa = self.cache.get("pretrain_args") or {}
        if not (self.cache.get("pretrain") or pa) or self.cache.get("mode") != "train":
            return None
        sizes = _gather(int(self.cache.get("data_size", 0)))
        chosen = max(range(len(sizes)), key=lambda i: (sizes[i], -i))
        state = [None]
        if self.rank == chosen:
            tr = trainer_cls(data_handle=handle, cache=dict(self.cache, fused=False), input=self.input, state=self.state)
            tr.init_nn()
            tr.train_local(handle.get_train_dataset(dataset_cls), handle.get_validation_dataset(dataset_cls),
                           epochs=pa.get("epochs", 1), patience=pa.get("patience"), lr=pa.get("learning_rate"),
                           local_iterations=pa.get("local_iterations"), tag="pretrained")
            tr.load_checkpoint(name="pretrained_best.pt", load_optimizer_state=False)
            state[0] = {k: {n: t.detach().cpu() for n, t in m.state_dict().items()} for k, m in tr.nn.items()}
        import torch.distributed as dist
        if self.world > 1:
            dist.broadcast_object_list(state, src=chosen)
        self._log(f"pre-trained on site {chosen} ({sizes[chosen]} items)")
        return state[0]

    def _epoch_steps(self, cycle):
        k = max(int(self.cache.get("local_iterations", 1)), 1)
        return max(max(_gather((len(cycle) + k - 1) // k)), 1), k

    # ------------------------------------------------------------------ the run
    def run(self, trainer_cls=None, dataset_cls=None, datahandle_cls=None):
        if trainer_cls is None:
            from ..tasks import get_task
            trainer_cls, dataset_cls, datahandle_cls = get_task(self.cache["task_id"])
        cache = self.cache
        handle = datahandle_cls(cache=cache, input=self.input, state=self.state, dataloader_args=DATALOADER_ARGS)
        handle.prepare_data()
        folds = _gather(len(cache["splits"]))
        if len(set(folds)) != 1:
            raise RuntimeError(f"sites disagree on the number of folds: {folds}")
        results = []
        for fold in range(folds[0]):
            cache["fold"] = fold
            results.append(self._run_fold(fold, trainer_cls, dataset_cls, handle))
        if self.rank == 0:
            self._zip_results()
        _barrier()
        for r in range(self.world):            # every site receives the global results (nnlogs.ipynb:72-84)
            if self.rank == 0 and r != 0:
                z = os.path.join(self.remote_out, f"{cache['task_id']}_global_results.zip")
                dst = os.path.join(self.work_dir, "output", f"local{r}", "simulatorRun")
                if os.path.isdir(dst):
                    shutil.copy(z, dst)
        if self.rank == 0:
            shutil.copy(os.path.join(self.remote_out, f"{cache['task_id']}_global_results.zip"),
                        self.state["outputDirectory"])
        _barrier()
        return results

    def _run_fold(self, fold, trainer_cls, dataset_cls, handle):
        cache = self.cache
        t_fold = time.time()
        pre = self._pretrain(trainer_cls, dataset_cls, handle)
        trainer = trainer_cls(data_handle=handle, cache=cache, input=self.input, state=self.state)
        trainer.init_nn(init_optim=False)
        if pre is not None:
            for k, sd in pre.items():
                trainer.nn[k].load_state_dict(sd)
        if cache.get("pretrained_path"):
            p = cache["pretrained_path"]
            p = p if os.path.isabs(p) else os.path.join(self.state["baseDirectory"], p)
            trainer.load_checkpoint(full_path=p, load_optimizer_state=False)
        if torch.cuda.is_available():
            engine = DeviceEngine(trainer, use_graph=self.use_graph, track_metrics=True)
        else:
            from .host_engine import HostEngine
            engine = HostEngine(trainer, track_metrics=True)
        engine.sync_shadow()
        out = {"agg_engine": cache["agg_engine"], "fold": fold, Key.TRAIN_LOG: [], Key.VALIDATION_LOG: [],
               "best_val_epoch": 0, "time_spent_on_computation": [], "cumulative_total_duration": []}
        direction = cache.get("metric_direction", "maximize")
        best = None
        # device-resident data path (default on a GPU): every split is uploaded once, batches are cut by a kernel inside
        # the step's CUDA graph, evaluation walks slices of the table — no DataLoader, no per-step host-to-device copy
        resident = torch.cuda.is_available() and bool(cache.get("device_resident", True))

        def _table(ds):
            if resident and ds is not None and len(ds) > 0:
                from ..data.device import DeviceTable
                ds._device_table = DeviceTable.from_dataset(ds, engine.device)
            return ds
        if str(cache.get("mode", "train")).lower() == "train":
            train_ds = handle.get_train_dataset(dataset_cls)
            cycle = _Cycle(handle, train_ds, cache)
            val_ds = _table(handle.get_validation_dataset(dataset_cls))
            steps, k = self._epoch_steps(cycle)
            batcher = None
            if resident and k == 1 and len(train_ds) > 0:
                from ..data.device import DeviceBatcher, DeviceTable
                batcher = DeviceBatcher(DeviceTable.from_dataset(train_ds, engine.device), int(cache.get("batch_size", 16)),
                                        seed=int(cache.get("seed", 0)))
                engine.attach_batcher(batcher)
                steps = max(max(_gather(len(batcher))), 1)
                out["data_path"] = {"mode": "device_resident", "h2d_bytes_once": batcher.table.h2d_bytes,
                                    "h2d_bytes_per_step": 0, "h2d_bytes_per_pass": batcher.h2d_bytes_per_pass}
            vep = max(int(cache.get("validation_epochs", 1)), 1)
            for ep in range(1, int(cache["epochs"]) + 1):
                t_ep = time.time()
                cache["epoch"] = ep
                for _ in range(steps):
                    if batcher is not None:
                        engine.train_step_resident()
                    elif k == 1:
                        engine.train_step(cycle.next())
                    else:
                        engine.train_step(None, micro_batches=[cycle.next() for _ in range(k)])
                tavg, tmet = self._merge(trainer, *engine.epoch_metrics())
                out[Key.TRAIN_LOG].append(trainer.log_row(tavg, tmet))
                if ep % vep == 0:
                    vavg, vmet = engine.evaluate(val_ds, "validation")
                    out[Key.VALIDATION_LOG].append(trainer.log_row(vavg, vmet))
                    score = trainer.score_of(vavg, vmet)
                    if cfg.metric_is_better(score, best, direction):
                        best, out["best_val_epoch"] = score, ep
                        trainer.save_checkpoint(name="best.pt")
                    self._log(f"[fold {fold}] epoch {ep}: train {out[Key.TRAIN_LOG][-1]} val {out[Key.VALIDATION_LOG][-1]}")
                out["time_spent_on_computation"].append(time.time() - t_ep)
                out["cumulative_total_duration"].append(time.time() - self.t0)
                if ep - out["best_val_epoch"] >= int(cache.get("patience", cache["epochs"])):
                    break
            trainer.save_checkpoint(name="latest.pt")
            if out["best_val_epoch"] == 0:
                trainer.save_checkpoint(name="best.pt")
            engine.load_checkpoint(name="best.pt")
        test_ds = _table(handle.get_test_dataset(dataset_cls))
        lavg, lmet = trainer.evaluation("test", [test_ds], save_pred=True)       # this site alone
        gavg, gmet = self._merge(trainer, lavg, lmet)                            # all sites
        out[Key.TEST_METRICS] = [trainer.log_row(lavg, lmet)]
        out[Key.GLOBAL_TEST_METRICS] = [trainer.log_row(gavg, gmet)]
        out["fold_duration"] = time.time() - t_fold
        out["steps"] = engine.steps
        logs = {k: v for k, v in cache.items() if not str(k).startswith("_")}
        logs.update(out)
        save_json(os.path.join(trainer.log_dir(), "logs.json"), clean_json(logs))
        if self.rank == 0:
            d = os.path.join(self.remote_out, str(cache["task_id"]), f"fold_{fold}")
            os.makedirs(d, exist_ok=True)
            rlog = dict(logs)
            rlog[Key.TEST_METRICS] = out[Key.GLOBAL_TEST_METRICS]
            save_json(os.path.join(d, "logs.json"), clean_json(rlog))
            from ..nodes.remote import write_test_metrics_csv
            cols = [c.strip() for grp in str(cache.get("log_header", "loss")).split("|") for c in grp.split(",")]
            write_test_metrics_csv(os.path.join(d, "test_metrics.csv"), fold, cols, trainer.log_row(gavg, gmet), gmet)
        self._log(f"[fold {fold}] test (all sites) {out[Key.GLOBAL_TEST_METRICS]} best epoch {out['best_val_epoch']}")
        engine.close()
        return out

    def _merge(self, trainer, avg, met):
        if self.world == 1:
            return avg, met
        payloads = _gather({"a": avg.serialize(), "m": met.serialize()})
        a, m = trainer.new_averages(), trainer.new_metrics()
        a.reduce_sites([p["a"] for p in payloads])
        m.reduce_sites([p["m"] for p in payloads])
        return a, m

    def _zip_results(self):
        task = str(self.cache["task_id"])
        root = os.path.join(self.remote_out, task)
        z = os.path.join(self.remote_out, f"{task}_global_results.zip")
        with zipfile.ZipFile(z, "w", zipfile.ZIP_DEFLATED) as zf:
            for dirpath, _, files in os.walk(root):
                for f in files:
                    full = os.path.join(dirpath, f)
                    zf.write(full, os.path.relpath(full, root))
        return z


def main(argv=None):
    import argparse
    ap = argparse.ArgumentParser(description="device-mode federated run: one site per GPU (launch with torchrun)")
    ap.add_argument("--task", default=None, help="FS-Classification | ICA-Classification (default: from inputspec)")
    ap.add_argument("--data-path", required=True)
    ap.add_argument("--work-dir", required=True)
    ap.add_argument("--agg-engine", default=None, choices=["dSGD", "rankDAD", "powerSGD"])
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--set", action="append", default=[], help="override key=json_value for all sites")
    ap.add_argument("-v", "--verbose", action="store_true")
    a = ap.parse_args(argv)
    ov = {}
    for kv in a.set:
        k, v = kv.split("=", 1)
        try:
            ov[k] = json.loads(v)
        except json.JSONDecodeError:
            ov[k] = v
    fed = DeviceFederation(a.task, a.data_path, a.work_dir, a.agg_engine, ov, use_graph=not a.no_graph, verbose=a.verbose)
    res = fed.run()
    if fed.rank == 0:
        print(json.dumps(clean_json({"sites": fed.world, "folds": len(res),
                                     "test_metrics": [r[Key.GLOBAL_TEST_METRICS] for r in res],
                                     "best_val_epoch": [r["best_val_epoch"] for r in res],
                                     "steps": [r["steps"] for r in res],
                                     "seconds": time.time() - fed.t0})))


if __name__ == "__main__":
    main()
