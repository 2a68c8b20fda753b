"""FreeSurfer-volume classification task (reference: ``comps/fs/__init__.py``).

Same behaviour as the reference's dataset/trainer/handle, different mechanics:

* the covariates CSV is parsed once per handle/dataset and kept as a dict;
* every subject's aseg-stats file (header line + 66 ``name<TAB>value`` rows,
  ``datasets/test_fsl/…/subject*_aseg_stats.txt``) is parsed once when it is indexed and held
  in one ``[n, 66]`` float32 tensor, instead of a pandas read per ``__getitem__``
  (``comps/fs/__init__.py:35``).  The per-subject normalisation — divide by that subject's own
  maximum (``:37``) — is applied at load;
* ``iteration`` keeps the loss on the device (no per-iteration ``.item()`` sync, ``:61``).
"""
from __future__ import annotations

import csv
import os

import numpy as np
import torch
import torch.nn.functional as F

from ..data import COINNDataset, COINNDataHandle
from ..models import MSANNet
from ..nn import COINNTrainer
from ..ops.elementwise import softmax_cross_entropy


def read_covariates(path, index_column=None):
    """CSV → ``(ordered keys, {key: row dict})``; key = ``index_column`` if present else row number."""
    with open(path, newline="") as f:
        rows = list(csv.DictReader(f))
    keys, table = [], {}
    for i, r in enumerate(rows):
        k = r[index_column] if index_column and index_column in r else i
        keys.append(k)
        table[k] = r
    return keys, table


def parse_aseg_stats(path, normalise=True):
    """One subject's volumes as float64 → divided by the subject's max (reference ``:37``).
    Uses the native parser of libdinunet_b200 when it is built, else pure Python."""
    from ..ops import native
    if native.available(cpu_only=True) and hasattr(native.lib(), "dnb_parse_aseg"):
        import ctypes
        buf = np.empty(4096, dtype=np.float64)
        n = native.lib().dnb_parse_aseg(os.fsencode(path), ctypes.c_void_p(buf.ctypes.data), ctypes.c_int(buf.size),
                                        ctypes.c_int(int(normalise)))
        if n >= 0:
            return buf[:n].copy()
    vals = []
    with open(path) as f:
        next(f, None)                     # "Measure:volume<TAB>subjectN"
        for line in f:
            parts = line.rstrip("\n").split("\t")
            if len(parts) >= 2 and parts[1] != "":
                vals.append(float(parts[1]))
    x = np.asarray(vals, dtype=np.float64)
    if normalise and x.size:
        x = x / x.max()
    return x


def _to_label(y):
    if isinstance(y, str):
        s = y.strip().lower()
        if s in ("true", "false"):
            return int(s == "true")
        return int(float(s))
    return int(y)


class FreeSurferDataset(COINNDataset):
    def __init__(self, **kw):
        super().__init__(**kw)
        self._table = None
        self._features = []

    def _labels(self):
        if self._table is None:
            p = os.path.join(self.state["baseDirectory"], self.cache["labels_file"])
            _, self._table = read_covariates(p, self.cache.get("data_column"))
        return self._table

    def load_index(self, file):
        y = _to_label(self._labels()[file][self.cache["labels_column"]])
        self.indices.append([file, y])                       # JSON-serialisable, like the reference
        self._features.append(parse_aseg_stats(os.path.join(self.path(), file)))

    def __getitem__(self, ix):
        _, y = self.indices[ix]
        return {"inputs": torch.from_numpy(self._features[ix]), "labels": torch.tensor(y), "ix": torch.tensor(ix)}

    def device_table(self, device):
        """The split's 66-feature table on the GPU (one upload; reference ``comps/fs/__init__.py:33-39`` re-reads a
        file per item)."""
        import numpy as np
        from ..data.device import DeviceTable
        x = torch.from_numpy(np.stack(self._features).astype(np.float32)).to(device)
        y = torch.as_tensor([int(y) for _, y in self.indices], dtype=torch.int64)
        return DeviceTable(x.contiguous(), y.to(device))


class FreeSurferTrainer(COINNTrainer):
    def _init_nn_model(self):
        self.nn["fs_net"] = MSANNet(in_size=self.cache["input_size"], hidden_sizes=self.cache["hidden_sizes"],
                                    out_size=self.cache["num_class"], fused=bool(self.cache.get("fused", True)))

    def iteration(self, batch):
        dev = self.device["gpu"]
        inputs = batch["inputs"].to(dev, non_blocking=True).float()
        labels = batch["labels"].to(dev, non_blocking=True).long()
        indices = batch["ix"].to(dev, non_blocking=True).long() if "ix" in batch else None
        loss, out = softmax_cross_entropy(self.nn["fs_net"](inputs), labels, log_softmax_out=True)
        predicted = out.argmax(1)
        score = self.new_metrics()
        score.add(predicted, labels)                         # hard predictions (reference :57-59)
        val = self.new_averages()
        val.add(loss.detach(), len(inputs))
        return {"out": out, "loss": loss, "averages": val, "metrics": score, "prediction": predicted,
                "indices": indices}


class FSVDataHandle(COINNDataHandle):
    def list_files(self):
        keys, _ = read_covariates(os.path.join(self.state["baseDirectory"], self.cache["labels_file"]),
                                  self.cache.get("data_column"))
        return list(keys)
