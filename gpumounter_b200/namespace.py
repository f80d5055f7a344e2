"""ref: pkg/util/namespace/namespace.go — nsenter configuration; device-node work done natively (setns + mknod)."""

from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

from ._lib import check, lib
from ._model import MountEnv, c_strings, enc, env_ptr, gmp_ns_config
from .device import NvidiaGPU


@dataclass
class Config:
    """ref: namespace.go:16-39, same field names."""
    Cgroup: bool = False
    CgroupFile: str = ""
    FollowContext: bool = False
    GID: int = 0
    IPC: bool = False
    IPCFile: str = ""
    Mount: bool = False
    MountFile: str = ""
    Net: bool = False
    NetFile: str = ""
    NoFork: bool = False
    PID: bool = False
    PIDFile: str = ""
    PreserveCredentials: bool = False
    RootDirectory: str = ""
    Target: int = 0
    UID: int = 0
    User: bool = False
    UserFile: str = ""
    UTS: bool = False
    UTSFile: str = ""
    WorkingDirectory: str = ""

    def to_c(self) -> gmp_ns_config:
        return gmp_ns_config(int(self.Cgroup), enc(self.CgroupFile), int(self.FollowContext), self.GID, int(self.IPC),
                             enc(self.IPCFile), int(self.Mount), enc(self.MountFile), int(self.Net), enc(self.NetFile),
                             int(self.NoFork), int(self.PID), enc(self.PIDFile), int(self.PreserveCredentials),
                             enc(self.RootDirectory), self.Target, self.UID, int(self.User), enc(self.UserFile),
                             int(self.UTS), enc(self.UTSFile), enc(self.WorkingDirectory))

    def buildCommand(self) -> List[str]:  # ref: namespace.go:70-165
        buf = ctypes.create_string_buffer(16384)
        argc = ctypes.c_int()
        c = self.to_c()
        check(lib().gmp_ns_build_command(ctypes.byref(c), buf, 16384, ctypes.byref(argc)), "buildCommand")
        return [s.decode() for s in buf.raw.split(b"\0")[:argc.value]]

    def Execute(self, program: str, *args: str) -> Tuple[str, str]:  # ref: namespace.go:42-68
        out = ctypes.create_string_buffer(65536)
        err = ctypes.create_string_buffer(65536)
        c = self.to_c()
        check(lib().gmp_ns_execute(ctypes.byref(c), program.encode(), c_strings(args), len(args), out, 65536, err, 65536),
              "Execute")
        return out.value.decode(errors="replace"), err.value.decode(errors="replace")


def AddGPUDeviceFile(config: Config, gpu: NvidiaGPU, env: Optional[MountEnv] = None) -> None:  # ref: namespace.go:167-177
    c = config.to_c()
    check(lib().gmp_ns_add_gpu_device_file(env_ptr(env), ctypes.byref(c), ctypes.byref(gpu._c)), "AddGPUDeviceFile")


def RemoveGPUDeviceFile(config: Config, gpu: NvidiaGPU, env: Optional[MountEnv] = None) -> None:  # ref: namespace.go:179-189
    c = config.to_c()
    check(lib().gmp_ns_remove_gpu_device_file(env_ptr(env), ctypes.byref(c), ctypes.byref(gpu._c)), "RemoveGPUDeviceFile")


def KillRunningGPUProcesses(config: Config, podGPUProcesses: Sequence[str]) -> None:  # ref: namespace.go:191-201
    c = config.to_c()
    check(lib().gmp_ns_kill_running_gpu_processes(ctypes.byref(c), c_strings(podGPUProcesses), len(podGPUProcesses)),
          "KillRunningGPUProcesses")
