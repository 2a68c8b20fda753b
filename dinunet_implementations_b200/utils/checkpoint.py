"""Checkpoint format.

``{'source': 'coinstac', 'models': {name: state_dict}, 'optimizers': {name: state_dict}}``
saved with ``torch.save``; the loader also accepts a bare ``state_dict``.  Tensors are
keyed by the reference's module names and stored as canonical fp32 on CPU,
so a checkpoint written by the fused sm_100a path loads into the reference's own
``models.py`` and vice versa.
"""
from __future__ import annotations

import os

import torch

SOURCE_TAG = "coinstac"


def _cpu_state(sd):
    out = {}
    for k, v in sd.items():
        if isinstance(v, torch.Tensor):
            v = v.detach().to("cpu")
            if v.is_floating_point():
                v = v.float()
            out[k] = v.clone()
        elif isinstance(v, dict):
            out[k] = _cpu_state(v)
        else:
            out[k] = v
    return out


def _unwrap(m):
    return m.module if hasattr(m, "module") and isinstance(m.module, torch.nn.Module) else m


def save_checkpoint(path: str, nn: dict, optimizer: dict | None = None, **extra):
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    chk = {"source": SOURCE_TAG, "models": {}, "optimizers": {}}
    for name, model in nn.items():
        chk["models"][name] = _cpu_state(_unwrap(model).state_dict())
    for name, opt in (optimizer or {}).items():
        if opt is not None and hasattr(opt, "state_dict"):
            chk["optimizers"][name] = _cpu_state(opt.state_dict())
    chk.update(extra)
    tmp = path + ".tmp"
    torch.save(chk, tmp)
    os.replace(tmp, path)
    return path


def load_checkpoint(path: str, nn: dict, optimizer: dict | None = None, map_location="cpu",
                    load_optimizer_state: bool = True, strict: bool = True, trusted: bool = False):
    """Checkpoints cross the site trust boundary (``pretrained_<site>.pt`` is written by another site and relayed
    by the aggregator; ``pretrained_path`` is user supplied), so they are read with the restricted unpickler
    (tensors, containers, strings, numbers: everything ``save_checkpoint`` writes).  ``trusted=True`` falls back
    to full unpickling for a file this process's owner wrote with extra Python objects in it."""
    try:
        chk = torch.load(path, map_location=map_location, weights_only=True)
    except Exception as e:      # noqa: BLE001  (pickle.UnpicklingError and friends)
        if not trusted:
            raise RuntimeError(f"{path}: not a plain tensor checkpoint ({type(e).__name__}: {e}); refusing to unpickle "
                               "arbitrary objects from a file that may come from another site") from e
        chk = torch.load(path, map_location=map_location, weights_only=False)
    if isinstance(chk, dict) and chk.get("source") == SOURCE_TAG:
        for name, model in nn.items():
            if name in chk["models"]:
                _unwrap(model).load_state_dict(chk["models"][name], strict=strict)
        if optimizer and load_optimizer_state:
            for name, opt in optimizer.items():
                if name in chk.get("optimizers", {}) and hasattr(opt, "load_state_dict"):
                    opt.load_state_dict(chk["optimizers"][name])
    else:
        # a bare state_dict: goes into the first (usually only) model
        first = next(iter(nn))
        _unwrap(nn[first]).load_state_dict(chk, strict=strict)
    return chk
