"""Multi-process CPU test (gloo, world_size 3): the federation driver's host-side logic without a GPU."""
import os
import subprocess
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    import socket
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return str(s.getsockname()[1])


def test_federation_three_processes_gloo(tmp_path):
    env = dict(os.environ, FED_TMP=str(tmp_path), CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="2")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=3", "--master-addr", "127.0.0.1",
           "--master-port", _free_port(), os.path.join(ROOT, "tests/multiproc/mp_federation_cpu.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "cpu federation on 3 sites: OK" in r.stdout


def test_single_process_train_then_test_only_mode(tmp_path):
    """World size 1 needs no launcher; ``mode: test`` with ``pretrained_path`` skips training and reproduces the
    trained model's test metrics (reference compspec: mode train | test)."""
    import json
    from dinunet_implementations_b200.data import synthetic
    from dinunet_implementations_b200.parallel.federation import DeviceFederation
    if torch.cuda.is_available():
        import pytest
        pytest.skip("CPU path (HostEngine); the CUDA path is covered by tests/test_gpu_federation.py")
    root = str(tmp_path / "fs")
    synthetic.make_fs_dataset(root, site_sizes=(80,), seed=5, effect=0.6)
    ov = {"epochs": 4, "num_workers": 0, "batch_size": 8, "patience": 10}
    res = DeviceFederation("FS-Classification", root, str(tmp_path / "w1"), overrides=ov).run()
    best = str(tmp_path / "w1/output/local0/simulatorRun/FS-Classification/fold_0/best.pt")
    assert os.path.exists(best) and res[0]["steps"] > 0
    res2 = DeviceFederation("FS-Classification", root, str(tmp_path / "w2"),
                            overrides=dict(ov, mode="test", pretrained_path=best)).run()
    assert res2[0]["steps"] == 0 and res2[0]["train_log"] == []
    a, b = res[0]["global_test_metrics"][0], res2[0]["global_test_metrics"][0]
    assert abs(a[0] - b[0]) < 1e-5 and abs(a[1] - b[1]) < 1e-9, (a, b)
    log = json.load(open(str(tmp_path / "w2/output/remote/simulatorRun/FS-Classification/fold_0/logs.json")))
    assert log["mode"] == "test"
