"""Train/validation/test splitting: ratio split, k-fold, or explicit split files.

Fields: ``split_ratio`` (``compspec.json:205-216``), ``num_folds`` (``:217-224``, mutually
exclusive with the ratio), ``split_files`` (``:239,265``).  Splits are plain JSON
(``{"train": [...], "validation": [...], "test": [...]}``) written to the site's output
dir, one file per fold, so a run can be audited or replayed.
"""
from __future__ import annotations

import json
import os
import random


def _shuffled(files, seed):
    files = list(files)
    random.Random(int(seed)).shuffle(files)
    return files


def ratio_split(files, ratio=(0.8, 0.1, 0.1), seed=0, shuffle=True):
    """2 numbers → train/test (validation empty→ uses test), 3 numbers → train/val/test."""
    ratio = list(ratio)
    if len(ratio) not in (2, 3):
        raise ValueError(f"split_ratio needs 2 or 3 numbers, got {ratio}")
    if abs(sum(ratio) - 1.0) > 1e-6:
        s = float(sum(ratio))
        ratio = [r / s for r in ratio]
    files = _shuffled(files, seed) if shuffle else list(files)
    n = len(files)
    n_train = int(round(ratio[0] * n))
    if len(ratio) == 2:
        return {"train": files[:n_train], "validation": [], "test": files[n_train:]}
    n_val = int(round(ratio[1] * n))
    n_val = min(n_val, n - n_train)
    return {"train": files[:n_train], "validation": files[n_train:n_train + n_val],
            "test": files[n_train + n_val:]}


def k_fold_splits(files, k, seed=0, shuffle=True):
    """Fold i: test = chunk i, validation = chunk (i+1) mod k, train = the rest."""
    k = int(k)
    if k < 3:
        raise ValueError("num_folds must be >= 3 (test, validation and train chunks)")
    files = _shuffled(files, seed) if shuffle else list(files)
    chunks = [files[i::k] for i in range(k)]
    out = []
    for i in range(k):
        test, val = chunks[i], chunks[(i + 1) % k]
        train = [f for j, c in enumerate(chunks) if j not in (i, (i + 1) % k) for f in c]
        out.append({"train": train, "validation": val, "test": test})
    return out


def create_splits(files, out_dir, split_ratio=None, num_folds=None, seed=0, split_files=None,
                  base_dir=None):
    """Materialise ``SPLIT_<k>.json`` files under ``out_dir``; returns their names in order."""
    os.makedirs(out_dir, exist_ok=True)
    if split_files:
        names = []
        for i, sf in enumerate(split_files):
            src = sf if os.path.isabs(sf) else os.path.join(base_dir or ".", sf)
            with open(src) as f:
                sp = json.load(f)
            nm = f"SPLIT_{i}.json"
            with open(os.path.join(out_dir, nm), "w") as f:
                json.dump({k: sp.get(k, []) for k in ("train", "validation", "test")}, f)
            names.append(nm)
        return names
    if num_folds:
        splits = k_fold_splits(files, num_folds, seed)
    else:
        splits = [ratio_split(files, split_ratio or (0.8, 0.1, 0.1), seed)]
    names = []
    for i, sp in enumerate(splits):
        nm = f"SPLIT_{i}.json"
        with open(os.path.join(out_dir, nm), "w") as f:
            json.dump(sp, f)
        names.append(nm)
    return names


def num_runs(split_ratio=None, num_folds=None, split_files=None):
    if split_files:
        return len(split_files)
    return int(num_folds) if num_folds else 1
