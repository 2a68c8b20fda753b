"""Small utilities used on both sides of the round protocol."""
from __future__ import annotations

import json
import os
import random
import time

import numpy as np
import torch

from . import logger  # noqa: F401  (re-export)


def duration(cache: dict, begin: float, key: str):
    """Append ``time.time() - begin`` to the list ``cache[key]``.

    Contract from ``local.py:51-52`` / ``remote.py:39-40``; the notebook proves the value is
    a per-round list (``nnlogs.ipynb:87-92`` sums it / takes ``[-1]``).
    """
    t = time.time() - begin
    cache.setdefault(key, []).append(t)
    return t


def set_seed(seed: int):
    seed = int(seed) % (2 ** 31 - 1)
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)


def unwrap_inputspec(entry: dict) -> dict:
    """``{"k": {"value": v}}`` → ``{"k": v}`` (``datasets/*/inputspec.json`` layout)."""
    out = {}
    for k, v in entry.items():
        out[k] = v["value"] if isinstance(v, dict) and set(v.keys()) == {"value"} else v
    return out


def clean_json(obj):
    """Make ``obj`` JSON-serialisable (numpy scalars/arrays, tensors, enums, sets)."""
    from enum import Enum
    if isinstance(obj, Enum):
        return obj.value
    if isinstance(obj, dict):
        return {str(clean_json(k)): clean_json(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple, set)):
        return [clean_json(v) for v in obj]
    if isinstance(obj, np.generic):
        return obj.item()
    if isinstance(obj, np.ndarray):
        return obj.tolist()
    if isinstance(obj, torch.Tensor):
        return obj.detach().cpu().tolist()
    if isinstance(obj, (str, int, float, bool)) or obj is None:
        return obj
    return str(obj)


def save_json(path: str, obj, indent=None):
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    with open(path, "w") as f:
        json.dump(clean_json(obj), f, indent=indent)


def load_json(path: str):
    with open(path) as f:
        return json.load(f)


def stringify_enum(v):
    from enum import Enum
    return v.value if isinstance(v, Enum) else v
