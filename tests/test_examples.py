"""examples/custom_task.py: a task defined outside the package (dataset + trainer + data handle) runs through the
drop-in ``coinstac_dinunet`` namespace in file mode and alone on one site."""
import importlib.util
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load():
    spec = importlib.util.spec_from_file_location("custom_task_example", os.path.join(ROOT, "examples", "custom_task.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_custom_task_file_mode_learns(tmp_path):
    log = _load().main("file", root=str(tmp_path))
    loss, auc = log["test_metrics"][0]
    assert auc > 0.9 and loss < 0.5, log["test_metrics"]
    assert log["agg_engine"] == "dSGD" and log["best_val_epoch"] >= 1


def test_custom_task_single_site(tmp_path):
    res = _load().main("site", root=str(tmp_path))
    assert res["train_log"][-1][0] < res["train_log"][0][0]          # the training loss goes down
    assert len(res["test_metrics"][0]) == 2
