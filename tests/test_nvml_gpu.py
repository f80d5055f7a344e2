"""The NVML rows of the hot path (a12, a13, a7) against the REAL driver (GPU only).

The fake NVML of the CPU suite cannot catch what goes wrong at this seam: the process queries use the `_v3` 24-byte
record while the reference's vendored header describes the 16-byte v1 one (ref: lib/nvml.h:225-231,
pkg/util/gpu/collector/nvml/nvml.go:33-73).  Each answer here is therefore compared with an independent reader of the
same driver: NVIDIA's own Python binding (pynvml), `nvidia-smi`, and the driver's /proc files.
"""

import glob
import os
import re
import shutil
import subprocess

import pytest

from gpumounter_b200 import device, probe

pytestmark = pytest.mark.gpu


def pynvml_processes(uuid):
    import pynvml

    pynvml.nvmlInit()
    try:
        h = pynvml.nvmlDeviceGetHandleByUUID(uuid)
        out = []
        for fn in (pynvml.nvmlDeviceGetGraphicsRunningProcesses, pynvml.nvmlDeviceGetComputeRunningProcesses):
            out += [(int(p.pid), None if p.usedGpuMemory is None else int(p.usedGpuMemory)) for p in fn(h)]
        return out
    finally:
        pynvml.nvmlShutdown()


def test_running_processes_lists_this_process(cuda):
    """ref: pkg/device/nvidia.go:58-87 - init, handle by UUID, graphics then compute processes, shutdown."""
    torch = cuda
    keep = torch.ones(1 << 20, device="cuda:0")  # this process holds a context (and memory) on GPU 0
    torch.cuda.synchronize()
    cu = probe.enumerate_cuda()
    nv = {d.uuid: d for d in probe.enumerate_nvml()}
    rec = device.New(nv[cu[0].uuid].minor, cu[0].uuid)
    mine = rec.GetRunningProcess()
    theirs = pynvml_processes(cu[0].uuid)
    assert [p.Pid for p in mine] == [pid for pid, _ in theirs], (mine, theirs)  # same order: graphics, then compute
    for p, (_, mem) in zip(mine, theirs):
        if mem is not None:
            assert abs(p.UsedGPUMemory - mem) <= 64 << 20, (p, mem)  # two readings a moment apart
    assert len(mine) >= 1, "this process holds a context on GPU 0; NVML must list at least it"
    if os.getpid() not in [pid for pid, _ in theirs]:
        pytest.skip(f"NVML on this box reports PIDs of another PID namespace ({theirs}; we are {os.getpid()}): "
                    "the library agrees with pynvml, the own-PID check needs hostPID as the worker has it")
    me = [p for p in mine if p.Pid == os.getpid()]
    assert len(me) == 1 and me[0].UsedGPUMemory > 0, (os.getpid(), mine)
    if len(cu) >= 2:  # nothing of ours runs on GPU 1
        other = device.New(nv[cu[1].uuid].minor, cu[1].uuid)
        assert os.getpid() not in [p.Pid for p in other.GetRunningProcess()]
    assert float(keep[0]) == 1.0


@pytest.mark.skipif(shutil.which("nvidia-smi") is None, reason="no nvidia-smi on this box")
def test_enumeration_equals_nvidia_smi():
    """ref: collector.go:56-77 - index order, UUID exactly as NVML prints it; plus the bus id used for the CUDA join."""
    out = subprocess.run(["nvidia-smi", "--query-gpu=index,uuid,pci.bus_id", "--format=csv,noheader"],
                         capture_output=True, text=True, check=True).stdout
    smi = [tuple(x.strip() for x in line.split(",")) for line in out.splitlines() if line.strip()]
    ours = [(str(d.nvml_index), d.uuid, d.pci_bus_id) for d in probe.enumerate_nvml()]
    assert ours == smi


def test_minor_numbers_equal_the_drivers_proc_files():
    """/proc/driver/nvidia/gpus/<bus id>/information carries 'Device Minor': the number in /dev/nvidia<N>."""
    files = glob.glob("/proc/driver/nvidia/gpus/*/information")
    if not files:
        pytest.skip("no /proc/driver/nvidia/gpus on this box")
    by_bus = {}
    for f in files:
        text = open(f).read()
        m = re.search(r"Device Minor:\s*(\d+)", text)
        u = re.search(r"GPU UUID:\s*(GPU-[0-9a-f-]+)", text)
        assert m, text
        by_bus[os.path.basename(os.path.dirname(f)).lower()] = (int(m.group(1)), u.group(1) if u else None)
    for d in probe.enumerate_nvml():
        bus = d.pci_bus_id.lower()
        key = bus if bus in by_bus else bus[4:]  # NVML prints an 8-digit domain, /proc a 4-digit one
        assert key in by_bus, (bus, sorted(by_bus))
        minor, uuid = by_bus[key]
        assert d.minor == minor
        if uuid and "?" not in uuid:
            assert d.uuid == uuid
