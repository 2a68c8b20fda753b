import importlib
import json
import os
import sys

import torch

from dinunet_implementations_b200 import SiteRunner
from dinunet_implementations_b200.utils import checkpoint as ckpt


def test_site_runner_reference_call_signature(reference_dir, tmp_path):
    """``SiteRunner(taks_id=…)`` exactly as in comps/fs/site_run.py:5-6 (typo included)."""
    if reference_dir not in sys.path:
        sys.path.insert(0, reference_dir)
    comps = importlib.import_module("comps")
    runner = SiteRunner(taks_id="FSL", data_path=os.path.join(reference_dir, "datasets/test_fsl"), mode="Train",
                        split_ratio=[0.8, 0.1, 0.1], out_dir=str(tmp_path), epochs=2, monitor_metric="auc",
                        log_header="loss|auc")
    res = runner.run(comps.FreeSurferTrainer, comps.FreeSurferDataset, comps.FSVDataHandle)
    assert len(res) == 1 and len(res[0]["test_metrics"][0]) == 2
    logs = json.load(open(os.path.join(str(tmp_path), "FSL/fold_0/logs.json")))
    assert "best_val_epoch" in logs and logs["task_id"] == "FSL"


def test_checkpoint_roundtrip_and_bare_state_dict(tmp_path):
    net = torch.nn.Sequential(torch.nn.Linear(4, 3), torch.nn.BatchNorm1d(3))
    opt = torch.optim.Adam(net.parameters(), lr=1e-3)
    net(torch.randn(8, 4)).sum().backward()
    opt.step()
    p = ckpt.save_checkpoint(str(tmp_path / "c.pt"), {"n": net}, {"adam": opt})
    chk = torch.load(p, weights_only=False)
    assert chk["source"] == "coinstac" and set(chk) >= {"models", "optimizers"}
    assert "0.weight" in chk["models"]["n"] and "1.running_mean" in chk["models"]["n"]   # buffers round-trip
    net2 = torch.nn.Sequential(torch.nn.Linear(4, 3), torch.nn.BatchNorm1d(3))
    opt2 = torch.optim.Adam(net2.parameters(), lr=1e-3)
    ckpt.load_checkpoint(p, {"n": net2}, {"adam": opt2})
    assert torch.equal(net2[0].weight, net[0].weight)
    assert opt2.state_dict()["state"][0]["step"] == opt.state_dict()["state"][0]["step"]
    torch.save(net.state_dict(), str(tmp_path / "bare.pt"))
    net3 = torch.nn.Sequential(torch.nn.Linear(4, 3), torch.nn.BatchNorm1d(3))
    ckpt.load_checkpoint(str(tmp_path / "bare.pt"), {"n": net3})
    assert torch.equal(net3[1].running_mean, net[1].running_mean)
