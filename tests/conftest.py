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


FAKE_NVML = os.path.join(ROOT, "tests", "fake_nvml", "libfake_nvml.so")
FAKE_UUIDS = [f"GPU-{i:08x}-{i:04x}-{i:04x}-{i:04x}-{i:012x}" for i in (0xA0, 0xB1, 0xC2, 0xD3, 0xE4, 0xF5, 0x16, 0x27)]


@pytest.fixture
def fake_nvml(tmp_path, monkeypatch):
    """Point the library at tests/fake_nvml and return a function that (re)writes its config.

    Usage: fake_nvml(gpus=[(uuid, minor, bus)], procs=[(uuid, pid, bytes, 'compute')], p2p=[(i, j, caps, status)],
                     fail=[('nvmlDeviceGetUUID', 999)], nvlink=[(uuid, 18)])
    """
    subprocess.run(["make", "-s", "-C", os.path.dirname(FAKE_NVML)], check=True)  # no-op when up to date
    cfg = tmp_path / "fake_nvml.cfg"
    monkeypatch.setenv("GPUMOUNTER_NVML_LIB", FAKE_NVML)
    monkeypatch.setenv("GMP_FAKE_NVML_CONFIG", str(cfg))

    def write(gpus=None, procs=(), p2p=(), fail=(), nvlink=(), extra=()):
        if gpus is None:
            # NVML index order deliberately differs from minor order, as on real HGX boxes
            gpus = [(FAKE_UUIDS[i], minor, f"00000000:{0x10 + i:02X}:00.0") for i, minor in enumerate([3, 2, 1, 0, 7, 6, 5, 4])]
        lines = [f"gpu {u} {m} {b}" for u, m, b in gpus]
        lines += [f"proc {u} {pid} {mem} {kind}" for u, pid, mem, kind in procs]
        lines += [f"p2p {i} {j} {c} {s}" for i, j, c, s in p2p]
        lines += [f"nvlink {u} {n}" for u, n in nvlink]
        lines += [f"fail {fn} {code}" for fn, code in fail]
        lines += list(extra)
        cfg.write_text("\n".join(lines) + "\n")
        return gpus

    write()
    return write
