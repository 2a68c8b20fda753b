"""ICA time-course classification task (reference: ``comps/icalstm/__init__.py``).

* ``ICADataset``: loads ``[subjects, components, time]`` once, cuts it into
  ``temporal_size // window_size`` windows of ``window_size`` samples every ``window_stride``
  (reference ``:28-32``) — here as one strided view + copy (or the CUDA gather kernel when the
  data already lives on the device) instead of a Python double loop — and serves
  ``[S, comps, window]`` items.
* ``ICATrainer``: ``ICALstm`` → softmax / cross-entropy; metrics are fed the positive-class
  probability (``:65``).  Loss stays on the device.
* ``ICADataHandle``: rows of the labels CSV as ``[[data_index, label], …]`` (``:74-77``).
"""
from __future__ import annotations

import csv
import os

import numpy as np
import torch
import torch.nn.functional as F

from ..data import COINNDataset, COINNDataHandle
from ..models import ICALstm
from ..nn import COINNTrainer
from ..ops.elementwise import softmax_cross_entropy


def window_view(data: np.ndarray, window_size: int, window_stride: int, temporal_size: int) -> np.ndarray:
    """``[N, C, T] → [N, S, C, W]`` with ``S = temporal_size // window_size``; window ``j`` is
    ``data[:, :, j*stride : j*stride + W]`` — same indexing as the reference loop."""
    N, C, T = data.shape
    S = int(temporal_size // window_size)
    need = (S - 1) * window_stride + window_size
    if need > T:
        raise ValueError(f"time axis has {T} samples, windows need {need}")
    s0, s1, s2 = data.strides
    v = np.lib.stride_tricks.as_strided(data, shape=(N, S, C, window_size),
                                        strides=(s0, s2 * window_stride, s1, s2), writeable=False)
    return np.ascontiguousarray(v)


def load_timecourses(path):
    arr = np.load(path, allow_pickle=False)
    if isinstance(arr, np.lib.npyio.NpzFile):     # the inputspec names an .npz
        arr = arr[arr.files[0]]
    return np.asarray(arr)


class ICADataset(COINNDataset):
    def __init__(self, **kw):
        super().__init__(**kw)
        self._raw, self._data = None, None
        self.window_size = int(self.cache["window_size"])
        self.window_stride = int(self.cache["window_stride"])
        self.temporal_size = int(self.cache["temporal_size"])
        self.num_components = int(self.cache["num_components"])

    def _load_indices(self, files, **kw):
        self._raw = load_timecourses(self.path(cache_key="data_file"))
        self._data = None
        self.indices += [[int(ix), int(y)] for ix, y in files]

    @property
    def data(self):
        """Host copy of the windows ``[N, S, C, W]`` — built on first use (the device path never needs it)."""
        if self._data is None and getattr(self, "_raw", None) is not None:
            self._data = window_view(self._raw, self.window_size, self.window_stride, self.temporal_size)
        return self._data

    @data.setter
    def data(self, v):
        self._data = v

    def device_table(self, device):
        """The split on the GPU: ONE upload of the subjects this split uses (raw ``[n, C, T]``), windows cut there by
        the K10 gather kernel (``dnb_window_gather``) — the reference's Python double loop (``:29-32``) on the device."""
        from ..data.device import DeviceTable
        from ..ops.elementwise import window_gather
        import numpy as _np
        rows = _np.asarray([ix for ix, _ in self.indices], dtype=_np.int64)
        raw = torch.from_numpy(_np.ascontiguousarray(self._raw[rows], dtype=_np.float32)).to(device)
        x = window_gather(raw, self.window_size, self.window_stride, self.temporal_size)
        y = torch.as_tensor([int(y) for _, y in self.indices], dtype=torch.int64)
        return DeviceTable(x.contiguous(), y.to(device))

    def __getitem__(self, ix):
        data_index, y = self.indices[ix]
        return {"inputs": torch.from_numpy(self.data[data_index]), "labels": y}


class ICATrainer(COINNTrainer):
    def _init_nn_model(self):
        self.nn["net"] = ICALstm(
            window_size=self.cache["window_size"], input_size=self.cache["input_size"],
            hidden_size=self.cache["hidden_size"], num_comps=self.cache["num_components"],
            num_cls=self.cache["num_class"], num_layers=self.cache.setdefault("num_layers", 1),
            bidirectional=self.cache.setdefault("bidirectional", True), fused=bool(self.cache.get("fused", True)))

    def iteration(self, batch):
        dev = self.device["gpu"]
        inputs = batch["inputs"].to(dev, non_blocking=True).float()
        labels = batch["labels"].to(dev, non_blocking=True).long()
        out, _ = self.nn["net"](inputs)
        loss, prob = softmax_cross_entropy(out, labels)       # K6: softmax + CE + gradient in one launch on CUDA
        score = self.new_metrics()
        score.add(prob[:, 1], labels)
        val = self.new_averages()
        val.add(loss.detach(), len(inputs))
        return {"out": prob, "loss": loss, "averages": val, "metrics": score}


class ICADataHandle(COINNDataHandle):
    def list_files(self):
        with open(os.path.join(self.state["baseDirectory"], self.cache["labels_file"]), newline="") as f:
            rows = list(csv.reader(f))
        out = []
        for r in rows[1:]:                      # first row is the header, as pandas.read_csv assumes
            out.append([int(float(r[0])), int(float(r[1]))])
        return out
