import importlib.util
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REFERENCE_DIR = os.environ.get("TDL_REFERENCE_DIR", "/root/reference")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")


def pytest_collection_modifyitems(config, items):
    import torch
    have = torch.cuda.is_available()
    ngpu = torch.cuda.device_count() if have else 0
    skip_gpu = pytest.mark.skip(reason="no CUDA device")
    skip_multi = pytest.mark.skip(reason="needs >= 2 CUDA devices")
    for item in items:
        if "gpu" in item.keywords and not have:
            item.add_marker(skip_gpu)
        if "multigpu" in item.keywords and ngpu < 2:
            item.add_marker(skip_multi)


def load_reference_module(name: str):
    """Import an unmodified reference module straight from the read-only mount (oracle)."""
    path = os.path.join(REFERENCE_DIR, f"{name}.py")
    if not os.path.exists(path):
        pytest.skip(f"reference module {path} not available")
    spec = importlib.util.spec_from_file_location(f"_reference_{name}", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="session")
def ref_trust():
    return load_reference_module("trust_manager")


@pytest.fixture(scope="session")
def ref_detector():
    return load_reference_module("attack_detector")
