"""Site entrypoint of the COINSTAC computation (counterpart of the reference ``local.py``).

COINSTAC calls ``run(data)`` once per round in a long-lived process; ``CACHE`` and the reducer
pool survive between calls.  Task dispatch is table-driven (``tasks.get_task``) instead of an
if/elif chain; defaults match the reference's (``local.py:29-37``).
"""
import multiprocessing as mp
import time

from dinunet_implementations_b200 import COINNLocal
from dinunet_implementations_b200.tasks import NNComputation, AggEngine, get_task
from dinunet_implementations_b200.utils import duration

DEFAULT_TASK = NNComputation.TASK_FREE_SURFER
DEFAULT_ENGINE = AggEngine.DECENTRALIZED_SGD

CACHE = {}
MP_POOL = None

DEFAULTS = dict(
    batch_size=16, epochs=21, patience=31, split_ratio=[0.7, 0.15, 0.15], pretrain_args=None,
    dataloader_args={"train": {"drop_last": True}}, num_class=2, monitor_metric="auc", log_header="loss|auc",
)


def run(data):
    global MP_POOL
    t_round = time.time()
    t_first = CACHE.setdefault("start_time", t_round)
    if MP_POOL is None and CACHE.get("num_reducers"):
        MP_POOL = mp.Pool(processes=int(CACHE["num_reducers"]))

    node = COINNLocal(task_id=DEFAULT_TASK, agg_engine=DEFAULT_ENGINE, cache=CACHE, input=data["input"],
                      state=data["state"], **DEFAULTS)
    out = node(MP_POOL, *get_task(node.cache["task_id"]))

    duration(CACHE, t_round, key="time_spent_on_computation")
    duration(CACHE, t_first, key="cumulative_total_duration")
    return out
