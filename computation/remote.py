"""Aggregator entrypoint (counterpart of the reference ``remote.py``)."""
import multiprocessing as mp
import time

from dinunet_implementations_b200 import COINNRemote
from dinunet_implementations_b200.tasks import get_task
from dinunet_implementations_b200.utils import duration

CACHE = {}
MP_POOL = None


def run(data):
    global MP_POOL
    t_round = time.time()
    t_first = CACHE.setdefault("start_time", t_round)
    if MP_POOL is None and CACHE.get("num_reducers"):
        MP_POOL = mp.Pool(processes=int(CACHE["num_reducers"]))

    node = COINNRemote(cache=CACHE, input=data["input"], state=data["state"])
    trainer_cls = get_task(node.cache["task_id"])[0]
    out = node(MP_POOL, trainer_cls)

    duration(CACHE, t_round, key="time_spent_on_computation")
    duration(CACHE, t_first, key="cumulative_total_duration")
    return out
