"""Device-resident data path (data/device.py): batches are cut on the GPU by a kernel that walks a device-side
permutation, so a training step copies nothing from the host (reference path: Dataset.__getitem__ + DataLoader +
.to(device), comps/icalstm/__init__.py:36-38,57-58)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_batcher_walks_a_shuffled_pass_and_replays_from_a_graph():
    from dinunet_implementations_b200.data.device import DeviceTable, DeviceBatcher
    N, B = 37, 8
    x = torch.arange(N * 6 * 5, dtype=torch.float32).view(N, 6, 5).cuda()
    y = (torch.arange(N) % 3).cuda()
    bat = DeviceBatcher(DeviceTable(x, y), B, seed=4)
    assert len(bat) == N // B
    seen = []
    for step in range(2 * len(bat)):                       # two passes, eager launches
        o = bat.next()
        torch.cuda.synchronize()
        ix = o["ix"].clone()
        assert torch.equal(o["inputs"], x[ix]) and torch.equal(o["labels"], y[ix])
        seen.append(ix.cpu())
    first, second = torch.cat(seen[:len(bat)]), torch.cat(seen[len(bat):])
    assert len(set(first.tolist())) == len(first) and len(set(second.tolist())) == len(second)   # no repeats inside a pass
    assert not torch.equal(first, second)                                                          # reshuffled between passes
    # captured once, replayed: the device-side cursor advances, the host only ticks
    g = torch.cuda.CUDAGraph()
    bat.tick()
    with torch.cuda.graph(g):
        o = bat.launch()
    got = []
    for _ in range(len(bat)):
        g.replay()
        torch.cuda.synchronize()
        got.append(o["ix"].clone().cpu())
        assert torch.equal(o["inputs"], x[o["ix"]])
        bat.tick()
    got = torch.cat(got)
    assert len(set(got.tolist())) == len(got)


def test_ica_dataset_device_table_matches_host_windows(tmp_path):
    from dinunet_implementations_b200.tasks.icalstm import ICADataset
    rng = np.random.default_rng(0)
    raw = rng.normal(size=(9, 4, 50)).astype(np.float32)
    np.save(tmp_path / "d.npy", raw)
    cache = {"window_size": 10, "window_stride": 10, "temporal_size": 50, "num_components": 4, "data_file": "d.npy"}
    ds = ICADataset(cache=cache, state={"baseDirectory": str(tmp_path)})
    ds.add(files=[[7, 1], [2, 0], [5, 1]])
    tbl = ds.device_table("cuda:0")
    assert tuple(tbl.inputs.shape) == (3, 5, 4, 10)
    for i in range(3):
        assert torch.equal(tbl.inputs[i].cpu(), ds[i]["inputs"].float())
    assert tbl.labels.tolist() == [1, 0, 1]


def test_federation_steps_without_per_step_host_copies(tmp_path):
    import json
    import os
    from dinunet_implementations_b200.data import synthetic
    from dinunet_implementations_b200.parallel.federation import DeviceFederation
    root = str(tmp_path / "ica")
    synthetic.make_ica_dataset(root, num_sites=1, subjects_per_site=80, num_components=16, temporal_size=120,
                               window_size=10, window_stride=10, input_size=64, hidden_size=64, effect=1.0)
    fed = DeviceFederation("ICA-Classification", root, str(tmp_path / "work"),
                           overrides={"epochs": 4, "patience": 10, "batch_size": 8, "num_workers": 0, "seed": 3})
    import torch.utils.data as tud
    made = []
    orig = tud.DataLoader.__init__

    def spy(self, *a, **k):
        made.append(1)
        return orig(self, *a, **k)
    tud.DataLoader.__init__ = spy
    try:
        res = fed.run()
    finally:
        tud.DataLoader.__init__ = orig
    log = json.load(open(os.path.join(str(tmp_path / "work"), "output/local0/simulatorRun/ICA-Classification/fold_0/logs.json")))
    assert log["data_path"]["mode"] == "device_resident" and log["data_path"]["h2d_bytes_per_step"] == 0
    assert res[0]["steps"] == 4 * 7
    # the only loader ever built is the one that counts the batches of a pass; none is iterated per step
    assert len(made) <= 2, len(made)
    assert log["train_log"][-1][0] < log["train_log"][0][0]
