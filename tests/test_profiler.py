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


def test_profile_enabled_reaches_logs_in_file_mode(tmp_path):
    """``profile_enabled`` (compspec field): per-call wall times of the engine's learner / reducer methods are
    written into every node's logs.json."""
    import json
    import os
    from dinunet_implementations_b200.data import synthetic
    from dinunet_implementations_b200.simulator import Simulator
    root = str(tmp_path / "fs")
    synthetic.make_fs_dataset(root, site_sizes=(40, 30), seed=1)
    comp = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "computation")
    sim = Simulator.from_data_path(comp, root, str(tmp_path / "work"), num_sites=2,
                                   overrides={"epochs": 1, "num_reducers": 0, "num_workers": 0, "profile_enabled": True})
    assert sim.run()["phase"] == "success"
    llog = json.load(open(str(tmp_path / "work/output/local0/simulatorRun/FS-Classification/fold_0/logs.json")))
    rlog = json.load(open(str(tmp_path / "work/output/remote/simulatorRun/FS-Classification/fold_0/logs.json")))
    assert {"learner_backward", "learner_to_reduce", "learner_step"} <= set(llog["profile"])
    assert "reducer_reduce" in rlog["profile"] and all(t >= 0 for t in rlog["profile"]["reducer_reduce"])
