"""Task definitions ("comps"): dataset + trainer + data-handle + model per task.

Same task ids and engine names as the reference (``comps/__init__.py:7-16``).  ``TASKS`` maps a
task id to its ``(Trainer, Dataset, DataHandle)`` triple — the dispatch that the reference
spells out with if/elif in ``local.py:40-47`` / ``remote.py:28-35``; add a new task by
registering it here.
"""
from enum import Enum

from ..config import AggEngine  # noqa: F401  (re-export, same strings as the reference)
from .fs import FreeSurferDataset, FreeSurferTrainer, FSVDataHandle
from .icalstm import ICADataset, ICATrainer, ICADataHandle


class NNComputation(str, Enum):
    TASK_FREE_SURFER = "FS-Classification"
    TASK_ICA = "ICA-Classification"


TASKS = {
    NNComputation.TASK_FREE_SURFER.value: (FreeSurferTrainer, FreeSurferDataset, FSVDataHandle),
    NNComputation.TASK_ICA.value: (ICATrainer, ICADataset, ICADataHandle),
}


def register_task(task_id, trainer_cls, dataset_cls, datahandle_cls):
    TASKS[str(getattr(task_id, "value", task_id))] = (trainer_cls, dataset_cls, datahandle_cls)


def get_task(task_id):
    key = str(getattr(task_id, "value", task_id))
    if key not in TASKS:
        raise ValueError(f"Invalid task: {task_id!r}; known: {sorted(TASKS)}")
    return TASKS[key]


__all__ = ["NNComputation", "AggEngine", "TASKS", "register_task", "get_task", "FreeSurferDataset",
           "FreeSurferTrainer", "FSVDataHandle", "ICADataset", "ICATrainer", "ICADataHandle"]
