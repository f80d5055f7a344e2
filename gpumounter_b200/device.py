"""ref: pkg/device/nvidia.go — the NvidiaGPU record, over the C ABI."""

from __future__ import annotations

import ctypes
from typing import List, Sequence

from ._lib import check, lib
from ._model import gmp_gpu, gmp_process_info

DEFAULT_NVIDA_MAJOR_NUMBER = 195  # ref: nvidia.go:37 (spelling kept)
DEFAULT_CGROUP_PERMISSION = "rw"
DEFAULT_DEVICE_FILE_PERMISSION = "666"
NVIDIA_DEVICE_FILE_PREFIX = "/dev/nvidia"
GPU_FREE_STATE = "GPU_FREE_STATE"
GPU_ALLOCATED_STATE = "GPU_ALLOCATED_STATE"


class ProcessInfo:
    """ref: pkg/util/gpu/collector/nvml/nvml.go:12-15"""

    def __init__(self, pid: int, used_gpu_memory: int):
        self.Pid = pid
        self.UsedGPUMemory = used_gpu_memory

    def __repr__(self):
        return f"ProcessInfo(Pid={self.Pid}, UsedGPUMemory={self.UsedGPUMemory})"


class NvidiaGPU:
    """ref: nvidia.go:10-17.  Wraps one gmp_gpu; attribute names follow the Go struct."""

    def __init__(self, c: gmp_gpu):
        self._c = c

    MinorNumber = property(lambda self: int(self._c.minor_number))
    DeviceFilePath = property(lambda self: self._c.device_file_path.decode())
    UUID = property(lambda self: self._c.uuid.decode())

    @property
    def State(self) -> str:
        return self._c.state.decode()

    @State.setter
    def State(self, v: str):
        self._c.state = v.encode()

    @property
    def PodName(self) -> str:
        return self._c.pod_name.decode()

    @PodName.setter
    def PodName(self, v: str):
        self._c.pod_name = v.encode()

    @property
    def Namespace(self) -> str:
        return self._c.ns.decode()

    @Namespace.setter
    def Namespace(self, v: str):
        self._c.ns = v.encode()

    def String(self) -> str:  # ref: nvidia.go:43-50
        buf = ctypes.create_string_buffer(2048)
        check(lib().gmp_device_string(ctypes.byref(self._c), buf, 2048), "gmp_device_string")
        return buf.value.decode()

    __str__ = String

    def ResetState(self) -> None:  # ref: nvidia.go:52-56
        check(lib().gmp_device_reset_state(ctypes.byref(self._c)), "gmp_device_reset_state")

    def GetRunningProcess(self) -> List[ProcessInfo]:  # ref: nvidia.go:58-87
        cap = 2048
        arr = (gmp_process_info * cap)()
        n = ctypes.c_int(0)
        check(lib().gmp_gpu_running_processes(self._c.uuid, arr, cap, ctypes.byref(n)), "gmp_gpu_running_processes")
        return [ProcessInfo(int(arr[i].pid), int(arr[i].used_gpu_memory)) for i in range(n.value)]

    def __repr__(self):
        return f"NvidiaGPU({self.String()})"


def New(minorNumber: int, uuid: str) -> NvidiaGPU:  # ref: nvidia.go:25-34
    c = gmp_gpu()
    check(lib().gmp_device_new(minorNumber, uuid.encode(), ctypes.byref(c)), "gmp_device_new")
    return NvidiaGPU(c)


def as_c_array(gpus: Sequence[NvidiaGPU]):
    arr = (gmp_gpu * max(1, len(gpus)))()
    for i, g in enumerate(gpus):
        ctypes.memmove(ctypes.byref(arr[i]), ctypes.byref(g._c), ctypes.sizeof(gmp_gpu))
    return arr
