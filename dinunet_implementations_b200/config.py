"""Keys, phases, modes and default arguments shared by both ends of the round protocol.

The reference only shows the *outside* of this protocol (``local.py:31-49``,
``remote.py:23-37``, ``compspec.json``); the wire vocabulary between our ``COINNLocal``
and ``COINNRemote`` is ours to define because both ends are ours.
"""
from __future__ import annotations

from enum import Enum


class Phase(str, Enum):
    INIT_RUNS = "init_runs"            # sites report data size + shared args
    NEXT_RUN = "next_run"              # remote assigned fold k / seed; sites build fold
    PRE_COMPUTATION = "pre_computation"  # optional pre-trained weights are distributed
    COMPUTATION = "computation"        # train / validation / test rounds
    NEXT_RUN_WAITING = "next_run_waiting"
    SUCCESS = "success"


class Mode(str, Enum):
    PRE_TRAIN = "pre_train"
    TRAIN = "train"
    VALIDATION = "validation"
    TEST = "test"
    VALIDATION_WAITING = "validation_waiting"   # site finished its epoch, still contributing
    TRAIN_WAITING = "train_waiting"             # site sent val scores, waits for the verdict
    TEST_WAITING = "test_waiting"


class AggEngine(str, Enum):
    """Same strings as ``comps/__init__.py:13-16``."""
    DECENTRALIZED_SGD = "dSGD"
    RANK_DAD = "rankDAD"
    POWER_SGD = "powerSGD"


class Key:
    TRAIN_SERIALIZABLE = "train_serializable"
    VALIDATION_SERIALIZABLE = "validation_serializable"
    TEST_SERIALIZABLE = "test_serializable"
    GLOBAL_TEST_SERIALIZABLE = "global_test_serializable"
    TRAIN_LOG = "train_log"
    VALIDATION_LOG = "validation_log"
    TEST_METRICS = "test_metrics"
    GLOBAL_TEST_METRICS = "global_test_metrics"
    ARGS_CACHED = "args_cached"


# Defaults mirror ``compspec.json`` (field -> lines): batch_size 80-87, local_iterations
# 88-95, learning_rate 96-107, epochs 108-119, validation_epochs 149-160, precision_bits
# 161-176, pin_memory/num_workers 177-192, patience 193-204, split_ratio 205-216,
# num_folds 217-224, num_reducers 72-79, dad_* 236-238.
DEFAULT_ARGS = {
    "mode": "train",
    "agg_engine": "dSGD",
    "num_reducers": 2,
    "batch_size": 16,
    "local_iterations": 1,
    "learning_rate": 0.001,
    "epochs": 101,
    "validation_epochs": 1,
    "precision_bits": "32",
    "pin_memory": False,
    "num_workers": 0,
    "patience": 35,
    "split_ratio": None,
    "num_folds": None,
    "split_files": None,
    "pretrain": False,
    "pretrain_args": None,
    "pretrained_path": None,
    "monitor_metric": "f1",
    "metric_direction": "maximize",
    "log_header": "loss|precision,recall,f1,accuracy",
    "seed": 244627,
    "gpus": [],
    "dad_reduction_rank": 10,
    "dad_num_pow_iters": 5,
    "dad_tol": 1e-3,
    "powersgd_rank": 4,
    "powersgd_start_iter": 0,
    "optimizer": "adam",
    "cpu_threads": "auto",
    "profile_enabled": False,
    "device_resident": True,       # device mode: keep every split on the GPU and cut batches there (data/device.py)
    "verbose": False,
    "persist_logs": True,
}

# Arguments the consortium owner sets once (``"source": "owner"`` in compspec.json); the
# first site's values are echoed back by the remote so every site trains with the same.
SHARED_ARGS = [
    "task_id", "mode", "agg_engine", "num_reducers", "batch_size", "local_iterations",
    "learning_rate", "epochs", "validation_epochs", "precision_bits", "patience",
    "split_ratio", "num_folds", "pretrain", "pretrain_args", "monitor_metric",
    "metric_direction", "log_header", "seed", "dad_reduction_rank", "dad_num_pow_iters",
    "dad_tol", "powersgd_rank", "optimizer", "profile_enabled",
]


def metric_is_better(new, old, direction="maximize", delta=0.0):
    if old is None:
        return True
    if str(direction).startswith("max"):
        return new > old + delta
    return new < old - delta
