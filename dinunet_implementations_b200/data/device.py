"""Device-resident datasets: a site's whole split lives in GPU memory and batches are cut on the device.

The reference serves every batch through ``Dataset.__getitem__`` → ``DataLoader`` → ``default_collate`` →
``.to(device)`` (``comps/icalstm/__init__.py:36-38,57-58``, ``comps/fs/__init__.py:33-39,51-52``): per step a Python
loop over the items, a pageable host copy and a host-to-device transfer.  Here the split is uploaded ONCE
(``DeviceTable``; the ICA windows are cut on the GPU by the K10 gather kernel) and ``DeviceBatcher`` selects the rows
of a step with one kernel that reads a device-side permutation and advances a device-side cursor — so the launch
can sit inside the step's CUDA graph and a training step moves no bytes over PCIe.  The host only reshuffles at pass
boundaries (one small asynchronous copy of the new permutation).
"""
from __future__ import annotations

import ctypes

import torch

from ..ops import native


class DeviceTable:
    """``inputs [N, ...] fp32``, ``labels [N] int64`` (+ ``ix [N]``) of one dataset on the device."""

    def __init__(self, inputs: torch.Tensor, labels: torch.Tensor):
        assert inputs.is_cuda and inputs.dtype == torch.float32 and inputs.is_contiguous()
        self.inputs = inputs
        self.labels = labels.to(device=inputs.device, dtype=torch.int64).contiguous()
        self.ix = torch.arange(inputs.shape[0], device=inputs.device, dtype=torch.int64)
        self.h2d_bytes = inputs.numel() * 4 + self.labels.numel() * 8

    def __len__(self):
        return int(self.inputs.shape[0])

    @property
    def item_shape(self):
        return tuple(self.inputs.shape[1:])

    def slice(self, lo, hi):
        return {"inputs": self.inputs[lo:hi], "labels": self.labels[lo:hi], "ix": self.ix[lo:hi]}

    @classmethod
    def from_dataset(cls, dataset, device):
        """Uses the dataset's own ``device_table(device)`` when it has one (bulk upload, device-side preprocessing);
        otherwise stacks ``dataset[i]`` on the host once and uploads the result."""
        device = torch.device(device)
        if len(dataset) == 0:
            return None
        if hasattr(dataset, "device_table"):
            return dataset.device_table(device)
        items = [dataset[i] for i in range(len(dataset))]
        x = torch.stack([torch.as_tensor(it["inputs"]).float() for it in items]).contiguous()
        y = torch.as_tensor([int(it["labels"]) for it in items], dtype=torch.int64)
        return cls(x.to(device), y.to(device))


class DeviceBatcher:
    """Endless shuffled batches of ``batch_size`` rows out of a ``DeviceTable`` (full batches only: a pass is
    ``max(N // B, 1)`` steps, like the train loader's ``drop_last=True``, ``local.py:29``; a split smaller than one
    batch wraps around).  ``tick()`` is the per-step host bookkeeping, ``launch()`` the gather kernel."""

    def __init__(self, table: DeviceTable, batch_size: int, seed: int = 0):
        self.table, self.B, self.seed = table, int(batch_size), int(seed)
        dev = table.inputs.device
        n = len(table)
        self.n_perm = n
        self.steps_per_pass = max(n // self.B, 1)
        self.perm = torch.zeros(n, device=dev, dtype=torch.int64)
        self._pinned = [torch.zeros(n, dtype=torch.int64).pin_memory() for _ in range(2)]
        self.state = torch.zeros(2, device=dev, dtype=torch.int64)       # {cursor, block counter}
        self.out = {"inputs": torch.zeros((self.B,) + table.item_shape, device=dev),
                    "labels": torch.zeros(self.B, device=dev, dtype=torch.int64),
                    "ix": torch.zeros(self.B, device=dev, dtype=torch.int64)}
        self.row_floats = int(table.inputs[0].numel())
        self.passes = 0
        self._step_in_pass = self.steps_per_pass       # forces a shuffle before the first step
        self.h2d_bytes_per_pass = n * 8

    def __len__(self):
        return self.steps_per_pass

    def _reshuffle(self):
        g = torch.Generator()
        g.manual_seed(self.seed + 7919 * self.passes + 1)
        host = self._pinned[self.passes & 1]
        host.copy_(torch.randperm(self.n_perm, generator=g))
        self.perm.copy_(host, non_blocking=True)
        native.memset_async(self.state)
        self.passes += 1
        self._step_in_pass = 0

    def tick(self):
        """Call once per step BEFORE the step is launched / replayed."""
        if self._step_in_pass >= self.steps_per_pass:
            self._reshuffle()
        self._step_in_pass += 1

    def launch(self):
        """Enqueue the gather of the next batch into the static output tensors (capturable)."""
        lib = native.require_gpu()
        lib.dnb_batch_gather.argtypes = [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p, ctypes.c_void_p,
                                         ctypes.c_longlong, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                         ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        t, o = self.table, self.out
        native.check(lib.dnb_batch_gather(t.inputs.data_ptr(), self.row_floats, t.labels.data_ptr(), self.perm.data_ptr(),
                                          self.n_perm, self.state.data_ptr(), self.B, o["inputs"].data_ptr(),
                                          o["labels"].data_ptr(), o["ix"].data_ptr(), native.stream()), "dnb_batch_gather")
        return o

    def next(self):
        self.tick()
        return self.launch()
