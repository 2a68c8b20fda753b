"""utils/profiler.py on the CPU fallback (the CUDA-event path is exercised by tests/multigpu/mp_dsgd_check.py
and tests/test_gpu_engine.py)."""
import time

from dinunet_implementations_b200.utils.profiler import PhaseTimer, timed


def test_phase_timer_cpu():
    t = PhaseTimer(device="cpu")
    for _ in range(3):
        with t.phase("a"):
            time.sleep(0.002)
        with t.phase("b"):
            pass
    s = t.summary()
    assert list(s) == ["a", "b"] and s["a"]["calls"] == 3
    assert s["a"]["ms_mean"] >= 1.5 and s["a"]["ms_max"] >= s["a"]["ms_mean"] > s["b"]["ms_mean"]
    t.reset()
    assert t.summary() == {}
    off = PhaseTimer(device="cpu", enabled=False)
    with off.phase("x"):
        pass
    assert off.summary() == {}


def test_timed_decorator_writes_into_cache():
    class Node:
        def __init__(self, on):
            self.cache = {"profile_enabled": on}

        @timed("work")
        def work(self, x):
            return x + 1

    n = Node(True)
    assert n.work(1) == 2 and n.work(2) == 3
    assert len(n.cache["profile"]["work"]) == 2
    m = Node(False)
    assert m.work(1) == 2 and "profile" not in m.cache
