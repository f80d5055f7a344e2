import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real NVIDIA B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session", autouse=True)
def _built_libraries():
    """Build what is missing.  On the GPU box the .so files travel with the snapshot."""
    lib = os.path.join(ROOT, "gpumounter_b200", "libgpumounter_probe.so")
    if not os.path.exists(lib):
        subprocess.run(["make", "-s", "-j8", "-C", os.path.join(ROOT, "gpumounter_b200", "csrc")], check=True)
    if not os.path.exists(os.path.join(ROOT, "oracle", "libgmo_oracle.so")):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle")], check=True)
    yield


@pytest.fixture(scope="session")
def cuda():
    """torch on cuda:0 — used only to own device memory and streams in the GPU tests."""
    import torch

    assert torch.cuda.is_available(), "GPU test selected on a box without CUDA"
    torch.cuda.set_device(0)
    return torch
