import os
import pathlib
import subprocess
import sys

import pytest

ROOT = pathlib.Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "metal-renderer_b200"))
sys.path.insert(0, str(ROOT / "oracle"))
sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")
    # build the oracle and the product library if they are missing (no-ops when up to date / prebuilt)
    if not (ROOT / "oracle" / "liboracle.so").exists():
        subprocess.check_call(["make", "-s", "-C", str(ROOT / "oracle"), "liboracle.so", "liboracle_fma.so"])
    if not (ROOT / "metal-renderer_b200" / "lib" / "libmrt_b200.so").exists():
        subprocess.check_call(["make", "-s", "-C", str(ROOT / "metal-renderer_b200"), "lib/libmrt_b200.so"])


def has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if has_cuda():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden_dir():
    return ROOT / "tests" / "golden"
