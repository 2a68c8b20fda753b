#!/usr/bin/env python
"""Define a NEW federated task and run it three ways — the workflow a user of the reference follows when adding a
`comps/<task>` package (dataset + trainer + data handle, `comps/fs/__init__.py` is the template):

    python examples/custom_task.py file      # COINSTAC file mode, 3 simulated sites in this process
    python examples/custom_task.py site      # one site alone (SiteRunner)
    torchrun --nproc-per-node 2 --master-addr 127.0.0.1 examples/custom_task.py device
                                             # one site per process: GPUs when present (fused dSGD over NVLink),
                                             # otherwise CPU + gloo

The task: classify 2-D points drawn around class-dependent centres.  Each site's data is one CSV (`x0,x1,label`).
"""
import csv
import json
import os
import sys
import tempfile

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("DINUNET_QUIET", "1")

from coinstac_dinunet import COINNDataHandle, COINNDataset, COINNLocal, COINNRemote, COINNTrainer   # noqa: E402
from coinstac_dinunet.utils import duration                                                         # noqa: E402

TASK = "Blobs-Classification"


# ----------------------------------------------------------------------------- the three classes of a task
class BlobDataHandle(COINNDataHandle):
    def list_files(self):                      # items must be JSON-serialisable; here: row numbers of the CSV
        with open(os.path.join(self.state["baseDirectory"], self.cache["points_file"])) as f:
            return list(range(sum(1 for _ in f) - 1))


class BlobDataset(COINNDataset):
    def __init__(self, **kw):
        super().__init__(**kw)
        self.rows = None

    def load_index(self, ix):                  # called once per item of the split
        if self.rows is None:
            with open(self.path(cache_key="points_file")) as f:
                self.rows = [r for r in csv.reader(f)][1:]
        self.indices.append(int(ix))

    def __getitem__(self, i):
        x0, x1, y = self.rows[self.indices[i]]
        return {"inputs": torch.tensor([float(x0), float(x1)]), "labels": torch.tensor(int(y)), "ix": torch.tensor(i)}


class BlobTrainer(COINNTrainer):
    def _init_nn_model(self):
        self.nn["net"] = torch.nn.Sequential(torch.nn.Linear(2, 16), torch.nn.ReLU(), torch.nn.Linear(16, 2))

    def iteration(self, batch):
        x, y = batch["inputs"].to(self.device["gpu"]).float(), batch["labels"].to(self.device["gpu"]).long()
        out = self.nn["net"](x)
        loss = F.cross_entropy(out, y)
        score, avg = self.new_metrics(), self.new_averages()
        score.add(torch.softmax(out, 1)[:, 1].detach(), y)      # positive-class probability: AUC-style metric
        avg.add(loss.detach(), len(x))
        return {"loss": loss, "averages": avg, "metrics": score, "out": out}


# ----------------------------------------------------------------------------- entrypoints (reference local.py / remote.py)
def make_callbacks():
    """One `run(data)` per node, each with its own process-like cache (the simulator keeps them in one process)."""
    def local_run():
        cache = {}

        def run(data):
            node = COINNLocal(task_id=TASK, agg_engine="dSGD", cache=cache, input=data["input"], state=data["state"],
                              batch_size=16, epochs=15, patience=10, learning_rate=0.01, split_ratio=[0.7, 0.15, 0.15], monitor_metric="auc",
                              log_header="loss|auc", dataloader_args={"train": {"drop_last": True}})
            out = node(None, BlobTrainer, BlobDataset, BlobDataHandle)
            duration(cache, cache.setdefault("t0", __import__("time").time()), key="cumulative_total_duration")
            return out
        return run
    rcache = {}

    def remote_run(data):
        return COINNRemote(cache=rcache, input=data["input"], state=data["state"])(None, BlobTrainer)
    return local_run, remote_run


def write_sites(root, sizes, seed=0):
    rng = np.random.default_rng(seed)
    inputs = []
    for s, n in enumerate(sizes):
        d = os.path.join(root, "input", f"local{s}", "simulatorRun")
        os.makedirs(d, exist_ok=True)
        y = rng.integers(0, 2, n)
        x = rng.normal(0, 1.0, (n, 2)) + np.where(y[:, None] == 1, [1.5, 1.0], [-1.5, -1.0]) + 0.2 * s
        with open(os.path.join(d, "points.csv"), "w", newline="") as f:
            w = csv.writer(f)
            w.writerow(["x0", "x1", "label"])
            w.writerows([[f"{a:.4f}", f"{b:.4f}", int(c)] for (a, b), c in zip(x, y)])
        inputs.append({"points_file": "points.csv", "num_workers": 0, "num_reducers": 0, "gpus": []})
    with open(os.path.join(root, "inputspec.json"), "w") as f:
        json.dump([{k: {"value": v} for k, v in i.items()} for i in inputs], f)
    return inputs


def main(mode="file", root=None):
    root = root or tempfile.mkdtemp(prefix="blobs_")
    if mode == "file":
        from dinunet_implementations_b200.simulator import Simulator
        inputs = write_sites(root, (120, 80, 100))
        mk_local, remote_run = make_callbacks()
        in_dirs = [os.path.join(root, "input", f"local{i}", "simulatorRun") for i in range(3)]
        sim = Simulator([mk_local() for _ in range(3)], remote_run, inputs, os.path.join(root, "work"), in_dirs)
        out = sim.run()
        log = json.load(open(os.path.join(root, f"work/output/remote/simulatorRun/{TASK}/fold_0/logs.json")))
        print(json.dumps({"mode": mode, "phase": out["phase"], "rounds": sim.rounds, "test_metrics": log["test_metrics"]}))
        return log
    if mode == "site":
        from coinstac_dinunet.site_runner import SiteRunner
        write_sites(root, (150,))
        res = SiteRunner(taks_id=TASK, data_path=root, mode="train", split_ratio=[0.7, 0.15, 0.15], batch_size=16, epochs=15,
                         learning_rate=0.01, monitor_metric="auc", log_header="loss|auc", out_dir=os.path.join(root, "site_out")
                         ).run(BlobTrainer, BlobDataset, BlobDataHandle)
        print(json.dumps({"mode": mode, "test_metrics": res[0]["test_metrics"]}))
        return res[0]
    if mode == "device":
        from dinunet_implementations_b200.parallel.engine import init_distributed_from_env
        from dinunet_implementations_b200.parallel.federation import DeviceFederation
        import torch.distributed as dist
        rank, world, _ = init_distributed_from_env()
        root = os.environ.get("BLOBS_ROOT", os.path.join(tempfile.gettempdir(), "blobs_device"))
        if rank == 0:
            write_sites(root, [120, 80, 100, 90, 110, 70, 130, 60][:world])
        if world > 1:
            dist.barrier()
        fed = DeviceFederation(TASK, root, os.path.join(root, "work"),
                               overrides={"epochs": 15, "batch_size": 16, "learning_rate": 0.01, "monitor_metric": "auc",
                                          "log_header": "loss|auc"})
        res = fed.run(BlobTrainer, BlobDataset, BlobDataHandle)
        if rank == 0:
            print(json.dumps({"mode": mode, "sites": world, "test_metrics": res[0]["global_test_metrics"]}))
        return res[0]
    raise SystemExit(f"unknown mode {mode!r}")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "file")
