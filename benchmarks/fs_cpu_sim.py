#!/usr/bin/env python
"""BASELINE.json config 1: FreeSurfer MLP (66-256-128-64-32-2, batch 16) under dSGD with 2 simulated
sites on CPU, through the COINSTAC file mode (in-repo simulator).  Reports aggregation rounds/s and
train samples/s.  Arms: `ours` = computation/ (our task definitions), `ref-app` = the reference's own
local.py/remote.py/comps running unmodified on our coinstac_dinunet namespace (when the tree exists)."""
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("DINUNET_QUIET", "1")

from dinunet_implementations_b200.simulator import Simulator  # noqa: E402


def find_ref():
    for base in (os.path.join(ROOT, "baseline", "_ref", "reference_src"), "/root/reference"):
        if os.path.exists(os.path.join(base, "local.py")):
            return base
    return None


def run(comp_dir, data, sites, epochs, engine):
    work = tempfile.mkdtemp(prefix="fs_cpu_")
    sim = Simulator.from_data_path(comp_dir, data, work, num_sites=sites,
                                   overrides={"epochs": epochs, "num_reducers": 0, "num_workers": 0, "patience": 1000,
                                              "agg_engine": engine})
    t0 = time.time()
    sim.run(max_rounds=100000)
    dt = time.time() - t0
    log = json.load(open(os.path.join(work, "output/remote/simulatorRun/FS-Classification/fold_0/logs.json")))
    train_rounds = sum(1 for h in sim.history if h["phase"] == "computation" and h["modes"] and
                       all(m in ("train", "validation_waiting") for m in h["modes"].values()))
    return {"rounds": sim.rounds, "train_rounds": train_rounds, "seconds": round(dt, 2),
            "rounds_per_sec": round(sim.rounds / dt, 2), "train_samples_per_sec": round(train_rounds * 16 * sites / dt, 1),
            "test_metrics": log["test_metrics"], "best_val_epoch": log["best_val_epoch"]}


if __name__ == "__main__":
    ref = find_ref()
    data = os.path.join(ref, "datasets/test_fsl") if ref else None
    if data is None:
        print(json.dumps({"unavailable": "test_fsl dataset (reference tree) not found"}))
        sys.exit(0)
    epochs = int(os.environ.get("FS_EPOCHS", "21"))
    # untimed warm-up of both apps: the first optimiser step makes PyTorch import half of torch._dynamo / sympy
    # (~5 s), which used to be billed to whichever arm ran first
    run(os.path.join(ROOT, "computation"), data, 2, 1, "dSGD")
    run(ref, data, 2, 1, "dSGD")
    for engine in ("dSGD", "rankDAD", "powerSGD"):
        out = {"config": "FS-MLP 66-256-128-64-32-2, batch 16, 2 sites, CPU, file mode", "engine": engine, "epochs": epochs}
        out["ours"] = run(os.path.join(ROOT, "computation"), data, 2, epochs, engine)
        out["ref_app_on_our_engine"] = run(ref, data, 2, epochs, engine)
        print(json.dumps(out), flush=True)
