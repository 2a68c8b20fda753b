"""Device-mode federation (``parallel/federation.py``): the full run — folds, epochs, validation, best
checkpoint, early stop, test, logs in the reference layout — with the fused kernels and a CUDA graph
on the step path.  One GPU = one site here; the multi-process variant is in test_gpu_multi.py."""
import json
import os
import zipfile

import pytest
import torch

pytestmark = pytest.mark.gpu


def _out(work, *p):
    return os.path.join(str(work), "output", *p)


def test_ica_device_run_single_site(tmp_path):
    from dinunet_implementations_b200.data import synthetic
    from dinunet_implementations_b200.parallel.federation import DeviceFederation
    from dinunet_implementations_b200.ops import native
    root = str(tmp_path / "ica")
    synthetic.make_ica_dataset(root, num_sites=1, subjects_per_site=80, num_components=16, temporal_size=120,
                               window_size=10, window_stride=10, input_size=64, hidden_size=64, effect=1.0)
    n0 = native.launch_count()
    fed = DeviceFederation("ICA-Classification", root, str(tmp_path / "work"),
                           overrides={"epochs": 6, "patience": 10, "batch_size": 8, "num_workers": 0, "seed": 3})
    res = fed.run()
    assert native.launch_count() > n0, "the step path must run on our kernels"
    assert len(res) == 1 and res[0]["steps"] == 6 * 7                     # 56 train items / 8 per batch
    log = json.load(open(_out(tmp_path / "work", "local0/simulatorRun/ICA-Classification/fold_0/logs.json")))
    for k in ("agg_engine", "test_metrics", "cumulative_total_duration", "time_spent_on_computation", "best_val_epoch",
              "train_log", "validation_log"):
        assert k in log, k
    assert len(log["train_log"]) == 6 and len(log["validation_log"]) == 6 and len(log["test_metrics"][0]) == 2
    assert log["train_log"][-1][0] < log["train_log"][0][0], log["train_log"]      # loss went down
    rdir = _out(tmp_path / "work", "remote/simulatorRun")
    rows = open(os.path.join(rdir, "ICA-Classification/fold_0/test_metrics.csv")).read().splitlines()
    assert len(rows) == 2 and rows[0].startswith("fold,loss,auc")
    with zipfile.ZipFile(os.path.join(rdir, "ICA-Classification_global_results.zip")) as z:
        assert "fold_0/logs.json" in z.namelist()
    chk = torch.load(_out(tmp_path / "work", "local0/simulatorRun/ICA-Classification/fold_0/best.pt"), weights_only=False)
    assert chk["source"] == "coinstac" and "net" in chk["models"]
    assert "classifier.2.running_mean" in chk["models"]["net"]


@pytest.mark.parametrize("engine", ["dSGD", "rankDAD", "powerSGD"])
def test_fs_device_run_kfold(tmp_path, engine):
    from dinunet_implementations_b200.data import synthetic
    from dinunet_implementations_b200.parallel.federation import DeviceFederation
    root = str(tmp_path / "fs")
    synthetic.make_fs_dataset(root, site_sizes=(90,), seed=2, effect=0.6)
    fed = DeviceFederation("FS-Classification", root, str(tmp_path / "work"), agg_engine=engine,
                           overrides={"epochs": 4, "num_folds": 3, "num_workers": 0, "patience": 10})
    res = fed.run()
    assert len(res) == 3
    for k in range(3):
        log = json.load(open(_out(tmp_path / "work", f"remote/simulatorRun/FS-Classification/fold_{k}/logs.json")))
        assert log["agg_engine"] == engine and len(log["test_metrics"][0]) == 2
    aucs = [r["global_test_metrics"][0][1] for r in res]
    assert max(aucs) > 0.6, aucs
