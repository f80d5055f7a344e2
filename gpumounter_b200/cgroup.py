"""ref: pkg/util/cgroup/cgroup.go — device-cgroup path arithmetic and permission writes, over the C ABI."""

from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence

from ._lib import check, lib
from ._model import MountEnv, Pod, c_strings, env_ptr
from .device import NvidiaGPU

PodQOSGuaranteed, PodQOSBurstable, PodQOSBestEffort = "Guaranteed", "Burstable", "BestEffort"
_QOS = {0: PodQOSGuaranteed, 1: PodQOSBurstable, 2: PodQOSBestEffort}


def GetPodQOS(pod: Pod) -> str:  # ref: cgroup.go:177-237
    q = ctypes.c_int()
    c = pod.to_c()
    check(lib().gmp_get_pod_qos(ctypes.byref(c), ctypes.byref(q)), "gmp_get_pod_qos")
    return _QOS[q.value]


class CgroupName(list):
    """ref: cgroup.go:19-22 — a list of name components."""

    def ToSystemd(self) -> str:  # ref: cgroup.go:52-68
        buf = ctypes.create_string_buffer(4096)
        check(lib().gmp_cgroup_name_to_systemd(c_strings(self), len(self), buf, 4096), "gmp_cgroup_name_to_systemd")
        return buf.value.decode()

    def ToCgroupfs(self) -> str:  # ref: cgroup.go:74-76
        buf = ctypes.create_string_buffer(4096)
        check(lib().gmp_cgroup_name_to_cgroupfs(c_strings(self), len(self), buf, 4096), "gmp_cgroup_name_to_cgroupfs")
        return buf.value.decode()


def NewCgroupName(base: Sequence[str], *components: str) -> CgroupName:  # ref: cgroup.go:32-45 (raises where Go panics)
    for c in components:
        check(lib().gmp_cgroup_name_check(c.encode()), "NewCgroupName")
    return CgroupName(list(base) + list(components))


def GetCgroupDriver() -> str:  # ref: cgroup.go:78-84, reads $CGROUP_DRIVER
    buf = ctypes.create_string_buffer(64)
    check(lib().gmp_get_cgroup_driver(buf, 64), "GetCgroupDriver")
    return buf.value.decode()


def GetCgroupName(cgroupDriver: str, pod: Pod, containerID: str) -> str:  # ref: cgroup.go:86-113
    buf = ctypes.create_string_buffer(4096)
    c = pod.to_c()
    check(lib().gmp_get_cgroup_name(cgroupDriver.encode(), ctypes.byref(c), containerID.encode(), buf, 4096), "GetCgroupName")
    return buf.value.decode()


def GetCgroupNameOfPod(pod: Pod, env: Optional[MountEnv] = None) -> str:
    """ref: util.go:21-34: strip "docker://", read the driver, GetCgroupName - what MountGPU resolves first."""
    buf = ctypes.create_string_buffer(4096)
    c = pod.to_c()
    check(lib().gmp_get_pod_cgroup_path(env_ptr(env), ctypes.byref(c), buf, 4096), "gmp_get_pod_cgroup_path")
    return buf.value.decode()


def DetectCgroupVersion(env: Optional[MountEnv] = None) -> int:
    """1: the v1 `devices` controller is mounted (also on hybrid hosts); 2: unified hierarchy only.  Extension, header 7."""
    v = ctypes.c_int()
    check(lib().gmp_detect_cgroup_version(env_ptr(env), ctypes.byref(v)), "gmp_detect_cgroup_version")
    return v.value


def GetDeviceGroupPath(cgroupPath: str, env: Optional[MountEnv] = None) -> str:  # ref: cgroup.go:115-118
    buf = ctypes.create_string_buffer(8192)
    check(lib().gmp_get_device_group_path(env_ptr(env), cgroupPath.encode(), buf, 8192), "GetDeviceGroupPath")
    return buf.value.decode()


def GetCgroupPIDs(cgroupPath: str, env: Optional[MountEnv] = None) -> List[str]:  # ref: cgroup.go:120-141
    cap = 8192
    arr = ((ctypes.c_char * 16) * cap)()
    n = ctypes.c_int()
    check(lib().gmp_get_cgroup_pids(env_ptr(env), cgroupPath.encode(), arr, cap, ctypes.byref(n)), "GetCgroupPIDs")
    return [arr[i].value.decode() for i in range(n.value)]


def AddGPUDevicePermission(cgroupPath: str, gpu: NvidiaGPU, env: Optional[MountEnv] = None) -> None:  # ref: cgroup.go:143-155
    check(lib().gmp_add_gpu_device_permission(env_ptr(env), cgroupPath.encode(), ctypes.byref(gpu._c)), "AddGPUDevicePermission")


def RemoveGPUDevicePermission(cgroupPath: str, gpu: NvidiaGPU, env: Optional[MountEnv] = None) -> None:  # ref: cgroup.go:157-169
    check(lib().gmp_remove_gpu_device_permission(env_ptr(env), cgroupPath.encode(), ctypes.byref(gpu._c)),
          "RemoveGPUDevicePermission")
