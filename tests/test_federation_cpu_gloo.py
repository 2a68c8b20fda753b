"""Multi-process CPU test (gloo, world_size 3): the federation driver's host-side logic without a GPU."""
import os
import subprocess
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_federation_three_processes_gloo(tmp_path):
    env = dict(os.environ, FED_TMP=str(tmp_path), CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="2")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=3", "--master-addr", "127.0.0.1",
           "--master-port", "29655", os.path.join(ROOT, "tests/multiproc/mp_federation_cpu.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "cpu federation on 3 sites: OK" in r.stdout
