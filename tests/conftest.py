"""pytest configuration: the `gpu` marker, repo root on sys.path, in-tree builds.

`-m "not gpu"` runs on a CPU-only container (oracle vs golden vectors, host
logic, C-ABI symbol check, gloo sharding tests); `-m gpu` runs the parity tests
proper on a B200 through the C ABI.
"""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_cuda():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Build the product library (nvcc cross-compiles without a GPU) and the C
    oracle once per session; both are no-ops when the outputs are fresh."""
    from instag_b200 import build as B
    B.build()
    B.build_binding()
    if not os.path.exists(os.path.join(ROOT, "oracle", "liboracle_cpu.so")):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "cpu"])
    yield


@pytest.fixture(scope="session")
def ref_available():
    from oracle import ref_cuda
    return ref_cuda.available()
