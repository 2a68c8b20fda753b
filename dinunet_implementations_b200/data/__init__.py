"""Dataset / data-handle base classes (the hooks ``comps/*`` subclass).

Contract evidenced by the reference's call sites:

* ``COINNDataset(**kw)``: ``self.state``, ``self.cache``, ``self.indices`` (appendable list),
  ``self.path()`` → the site's data dir, ``self.path(cache_key='data_file')`` → the file
  named by ``cache['data_file']`` (``comps/fs/__init__.py:35``,
  ``comps/icalstm/__init__.py:27``).  Two override levels: per-item ``load_index(file)``
  (``comps/fs/__init__.py:16``) and bulk ``_load_indices(files, **kw)``
  (``comps/icalstm/__init__.py:26``).
* ``COINNDataHandle``: ``self.state``, ``self.cache``, hook ``list_files()``
  (``comps/fs/__init__.py:67``, ``comps/icalstm/__init__.py:74``).
"""
from __future__ import annotations

import json
import os

import torch
from torch.utils.data import DataLoader, Dataset
from torch.utils.data._utils.collate import default_collate

from . import splits as _splits


def safe_collate(batch):
    return default_collate([b for b in batch if b is not None])


class COINNDataset(Dataset):
    def __init__(self, mode: str = "init", limit=None, cache=None, state=None, **kw):
        self.mode = mode
        self.limit = limit if limit is not None else float("inf")
        self.indices = []
        self.cache = cache if cache is not None else {}
        self.state = state if state is not None else {}
        self.input = kw.get("input", {})

    # -- hooks ------------------------------------------------------------------------
    def load_index(self, file):
        """Per-item hook; default keeps the item as the index entry."""
        self.indices.append([file])

    def _load_indices(self, files, **kw):
        """Bulk hook; default loops ``load_index`` and honours ``limit``."""
        for f in files:
            if len(self) >= self.limit:
                break
            self.load_index(f)

    def __getitem__(self, ix):
        raise NotImplementedError("Dataset subclasses return a dict of tensors here")

    def __len__(self):
        return len(self.indices)

    # -- framework side ---------------------------------------------------------------
    def path(self, cache_key=None):
        base = self.state.get("baseDirectory", ".")
        data_dir = self.cache.get("data_dir", "")
        root = os.path.join(base, data_dir) if data_dir else base
        if cache_key is None:
            return root
        return os.path.join(root, str(self.cache[cache_key]))

    def add(self, files, cache=None, state=None, **kw):
        if cache is not None:
            self.cache = cache
        if state is not None:
            self.state = state
        self._load_indices(files=files, **kw)
        return self

    def transforms(self, **kw):
        return None


class COINNDataHandle:
    """Lists a site's items, owns the fold splits, builds datasets and loaders."""

    def __init__(self, cache=None, input=None, state=None, dataloader_args=None, **kw):
        self.cache = cache if cache is not None else {}
        self.input = input if input is not None else {}
        self.state = state if state is not None else {}
        self.dataloader_args = dataloader_args or {}

    # -- hook -------------------------------------------------------------------------
    def list_files(self) -> list:
        """Default: every file in the site's data dir (override for CSV-indexed data)."""
        d = self.state.get("baseDirectory", ".")
        dd = self.cache.get("data_dir", "")
        d = os.path.join(d, dd) if dd else d
        return sorted(os.listdir(d))

    # -- splits -----------------------------------------------------------------------
    def split_dir(self):
        return os.path.join(self.state.get("outputDirectory", "."), str(self.cache.get("task_id", "task")),
                            "splits")

    def prepare_data(self):
        files = self.list_files()
        names = _splits.create_splits(
            files, self.split_dir(),
            split_ratio=self.cache.get("split_ratio"),
            num_folds=self.cache.get("num_folds"),
            seed=self.cache.get("seed", 0),
            split_files=self.cache.get("split_files") or None,
            base_dir=self.state.get("baseDirectory"),
        )
        self.cache["splits"] = names
        self.cache["data_size"] = len(files)
        return names

    def get_split(self, fold: int = None):
        fold = int(self.cache.get("fold", 0) if fold is None else fold)
        with open(os.path.join(self.split_dir(), self.cache["splits"][fold])) as f:
            return json.load(f)

    # -- datasets / loaders -----------------------------------------------------------
    def get_dataset(self, handle_key, files, dataset_cls, **kw):
        ds = dataset_cls(mode=handle_key, limit=self.cache.get("load_limit"), cache=self.cache,
                         state=self.state, input=self.input)
        ds.add(files=files, cache=self.cache, state=self.state, **kw)
        return ds

    def get_train_dataset(self, dataset_cls):
        return self.get_dataset("train", self.get_split()["train"], dataset_cls)

    def get_validation_dataset(self, dataset_cls):
        sp = self.get_split()
        return self.get_dataset("validation", sp["validation"] or sp["test"], dataset_cls)

    def get_test_dataset(self, dataset_cls):
        return self.get_dataset("test", self.get_split()["test"], dataset_cls)

    def get_loader(self, handle_key="", dataset=None, **kw):
        args = {
            "batch_size": int(self.cache.get("batch_size", 16)),
            "shuffle": handle_key == "train",
            "num_workers": int(self.cache.get("num_workers", 0)),
            "pin_memory": bool(self.cache.get("pin_memory", False)),
            "drop_last": False,
            "collate_fn": safe_collate,
        }
        args.update(self.dataloader_args.get(handle_key, {}))
        args.update(kw)
        if args.get("shuffle") and "generator" not in args:
            g = torch.Generator()
            g.manual_seed(int(self.cache.get("seed", 0)) + int(self.cache.get("epoch", 0)) * 7919 + 1)
            args["generator"] = g
        if len(dataset) < args["batch_size"]:
            args["drop_last"] = False
        if not args["drop_last"] and len(dataset) > 1 and len(dataset) % args["batch_size"] == 1:
            # a trailing batch of one sample breaks batch-statistics BatchNorm (the FreeSurfer
            # net normalises with batch stats even in eval, comps/fs/models.py:15,19)
            args["batch_size"] += 1
        return DataLoader(dataset, **args)


__all__ = ["COINNDataset", "COINNDataHandle", "safe_collate"]
