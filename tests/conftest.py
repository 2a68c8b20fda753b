import importlib.util
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
os.environ.setdefault("DINUNET_QUIET", "1")

REFERENCE = os.environ.get("DINUNET_REFERENCE", "/root/reference")
if not os.path.isdir(REFERENCE):
    # the GPU boxes have no /root/reference: __graft_entry__.build() stages a verbatim copy (git-ignored) for them
    _staged = os.path.join(ROOT, "baseline", "_ref", "reference_src")
    if os.path.isdir(_staged):
        REFERENCE = _staged


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: long-running CPU test")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


def load_by_path(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="session")
def reference_dir():
    if not os.path.isdir(REFERENCE):
        pytest.skip("reference tree not mounted")
    return REFERENCE


@pytest.fixture(scope="session")
def ref_fs_models(reference_dir):
    """The reference's own FreeSurfer model file, loaded by path (numerical oracle)."""
    return load_by_path(os.path.join(reference_dir, "comps/fs/models.py"), "_ref_fs_models")


@pytest.fixture(scope="session")
def ref_ica_models(reference_dir):
    return load_by_path(os.path.join(reference_dir, "comps/icalstm/models.py"), "_ref_ica_models")
