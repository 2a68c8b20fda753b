"""Synthetic datasets (``data/synthetic.py``) in the reference's on-disk layout, run through OUR
computation app (``computation/``) with the in-repo simulator — no dependency on the mounted
reference tree (counterpart of SURVEY R15/R16 + §4.3)."""
import json
import os

import numpy as np

from dinunet_implementations_b200.data import synthetic
from dinunet_implementations_b200.simulator import Simulator

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COMP = os.path.join(ROOT, "computation")


def test_fs_layout_matches_the_reference_toy_dataset(tmp_path):
    root = str(tmp_path / "fs")
    spec = synthetic.make_fs_dataset(root, seed=3)
    assert [len(os.listdir(os.path.join(root, "input", f"local{i}", "simulatorRun"))) - 1 for i in range(5)] == \
        list(synthetic.FS_SITE_SIZES)
    sub = os.path.join(root, "input/local0/simulatorRun/subject0_aseg_stats.txt")
    lines = open(sub).read().splitlines()
    assert lines[0] == "Measure:volume\tsubject0" and len(lines) == 1 + 66
    assert all(len(l.split("\t")) == 2 for l in lines[1:])
    raw = json.load(open(os.path.join(root, "inputspec.json")))
    assert len(raw) == 5 and raw[0]["labels_file"] == {"value": "site1_Covariate.csv"}
    assert spec[0]["input_size"] == 66 and spec[0]["hidden_sizes"] == [256, 128, 64, 32]
    from dinunet_implementations_b200.tasks.fs import parse_aseg_stats
    x = parse_aseg_stats(sub)
    assert x.shape == (66,) and abs(x.max() - 1.0) < 1e-12


def test_fs_three_sites_train_to_success_and_learn(tmp_path):
    root = str(tmp_path / "fs")
    synthetic.make_fs_dataset(root, site_sizes=(60, 40, 50), seed=1, effect=0.6)
    sim = Simulator.from_data_path(COMP, root, str(tmp_path / "work"), num_sites=3,
                                   overrides={"epochs": 6, "num_reducers": 0, "patience": 50, "num_workers": 0})
    out = sim.run(max_rounds=5000)
    assert out["phase"] == "success"
    log = json.load(open(os.path.join(str(tmp_path / "work"), "output/remote/simulatorRun/FS-Classification/fold_0/logs.json")))
    loss, auc = log["test_metrics"][0]
    assert np.isfinite(loss) and auc > 0.7, log["test_metrics"]      # planted signal is learnable
    # the log-summary tool (counterpart of the reference's nnlogs.ipynb) reads a work dir, a node dir or the zip
    from dinunet_implementations_b200.tools import nnlogs
    rdir = os.path.join(str(tmp_path / "work"), "output/remote/simulatorRun")
    z = [f for f in os.listdir(rdir) if f.endswith(".zip")][0]
    for src in (str(tmp_path / "work"), rdir, os.path.join(rdir, z)):
        rows = nnlogs.summarise(src)
        assert len(rows) == 1 and rows[0]["agg_engine"] == "dSGD" and rows[0]["fold"] == "fold_0"
        assert abs(rows[0]["test_score"] - auc) < 1e-9 and rows[0]["total_s"] > 0 and "accuracy" in rows[0]
    assert "dSGD [1 fold(s)]" in nnlogs.render(rows)


def test_ica_two_sites_cpu(tmp_path):
    root = str(tmp_path / "ica")
    spec = synthetic.make_ica_dataset(root, num_sites=2, subjects_per_site=12, num_components=6, temporal_size=40,
                                      window_size=5, window_stride=5, input_size=8, hidden_size=12, gpus_per_site=False)
    assert spec[1]["labels_file"] == "S2_2.csv"
    arr = np.load(os.path.join(root, "input/local0/simulatorRun", spec[0]["data_file"]))
    assert arr[arr.files[0]].shape == (12, 6, 40)
    sim = Simulator.from_data_path(COMP, root, str(tmp_path / "work"), num_sites=2,
                                   overrides={"task_id": "ICA-Classification", "epochs": 1, "num_reducers": 0,
                                              "batch_size": 4, "num_workers": 0, "split_ratio": [0.5, 0.25, 0.25]})
    out = sim.run(max_rounds=2000)
    assert out["phase"] == "success"
    assert os.path.exists(os.path.join(str(tmp_path / "work"),
                                       "output/remote/simulatorRun/ICA-Classification/fold_0/logs.json"))


def test_nnlogs_boxplots_compare_runs_over_folds(tmp_path):
    """Counterpart of NB.ipynb's 10-fold accuracy / F1 box plots and the scratch-vs-pretrain early-stop comparison."""
    from dinunet_implementations_b200.tools import nnlogs
    runs = []
    for name, pre in (("scratch", False), ("pretrained", True)):
        root = str(tmp_path / f"fs_{name}")
        synthetic.make_fs_dataset(root, site_sizes=(45, 45), seed=4, effect=0.6)
        ov = {"epochs": 3, "num_reducers": 0, "num_folds": 3, "num_workers": 0, "patience": 50}
        if pre:
            ov.update({"pretrain": True, "pretrain_args": {"epochs": 2, "batch_size": 16, "patience": 5}})
        sim = Simulator.from_data_path(COMP, root, str(tmp_path / f"work_{name}"), num_sites=2, overrides=ov)
        assert sim.run(max_rounds=8000)["phase"] == "success"
        runs.append(str(tmp_path / f"work_{name}"))
    rows = [r for p in runs for r in nnlogs.summarise(p)]
    assert len(rows) == 6
    written, epochs = nnlogs.boxplots(rows, str(tmp_path / "plots" / "cmp"), labels=["scratch", "pretrained"])
    assert any(w.endswith("_accuracy.svg") for w in written) and any(w.endswith("_f1.svg") for w in written)
    svg = open([w for w in written if w.endswith("_accuracy.svg")][0]).read()
    assert svg.startswith("<svg") and "scratch" in svg and "pretrained" in svg and svg.count("<rect") >= 3
    assert set(epochs) == {"scratch", "pretrained"}
