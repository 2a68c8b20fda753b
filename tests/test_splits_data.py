import json
import os

from dinunet_implementations_b200.data import COINNDataset, COINNDataHandle
from dinunet_implementations_b200.data import splits


def test_ratio_split_partitions():
    files = [f"f{i}" for i in range(100)]
    sp = splits.ratio_split(files, [0.7, 0.15, 0.15], seed=3)
    assert sorted(sp["train"] + sp["validation"] + sp["test"]) == sorted(files)
    assert len(sp["train"]) == 70 and len(sp["validation"]) == 15
    assert sp == splits.ratio_split(files, [0.7, 0.15, 0.15], seed=3)
    two = splits.ratio_split(files, [0.8, 0.2], seed=3)
    assert len(two["validation"]) == 0 and len(two["test"]) == 20


def test_k_fold_each_item_tested_once():
    files = list(range(53))
    folds = splits.k_fold_splits(files, 5, seed=1)
    tested = [f for sp in folds for f in sp["test"]]
    assert sorted(tested) == files
    for sp in folds:
        assert not set(sp["train"]) & set(sp["test"]) and not set(sp["train"]) & set(sp["validation"])
        assert sorted(sp["train"] + sp["validation"] + sp["test"]) == files


def test_handle_writes_split_files(tmp_path):
    base = tmp_path / "in"
    base.mkdir()
    for i in range(20):
        (base / f"s{i}.txt").write_text("x")
    cache = {"task_id": "T", "num_folds": 4, "seed": 1}
    state = {"baseDirectory": str(base), "outputDirectory": str(tmp_path / "out")}
    h = COINNDataHandle(cache=cache, state=state)
    names = h.prepare_data()
    assert len(names) == 4 and cache["data_size"] == 20
    sp = json.load(open(os.path.join(h.split_dir(), names[2])))
    assert set(sp) == {"train", "validation", "test"}


def test_dataset_hooks_and_path(tmp_path):
    class D(COINNDataset):
        def load_index(self, file):
            self.indices.append([file, 1])

        def __getitem__(self, ix):
            return {"x": self.indices[ix][1]}

    cache = {"data_file": "blob.npy"}
    state = {"baseDirectory": str(tmp_path)}
    d = D(cache=cache, state=state)
    d.add(files=["a", "b", "c"], cache=cache, state=state)
    assert len(d) == 3 and d.indices[0] == ["a", 1]
    assert d.path() == str(tmp_path)
    assert d.path(cache_key="data_file") == os.path.join(str(tmp_path), "blob.npy")
