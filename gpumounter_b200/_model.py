"""ctypes structures of section 6 of include/gpumounter_probe.h and the small Python pod model."""

from __future__ import annotations

import ctypes
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

GMP_UUID_LEN = 96


class gmp_gpu(ctypes.Structure):
    _fields_ = [("minor_number", ctypes.c_int32), ("device_file_path", ctypes.c_char * 64),
                ("uuid", ctypes.c_char * GMP_UUID_LEN), ("state", ctypes.c_char * 24),
                ("pod_name", ctypes.c_char * 256), ("ns", ctypes.c_char * 256)]


class gmp_container_resources(ctypes.Structure):
    _fields_ = [("cpu_request", ctypes.c_char_p), ("memory_request", ctypes.c_char_p),
                ("cpu_limit", ctypes.c_char_p), ("memory_limit", ctypes.c_char_p)]


class gmp_pod(ctypes.Structure):
    _fields_ = [("name", ctypes.c_char_p), ("ns", ctypes.c_char_p), ("uid", ctypes.c_char_p),
                ("node_name", ctypes.c_char_p), ("container_id", ctypes.c_char_p),
                ("n_containers", ctypes.c_int32), ("containers", ctypes.POINTER(gmp_container_resources))]


class gmp_mount_env(ctypes.Structure):
    _fields_ = [("devices_root", ctypes.c_char_p), ("dev_prefix", ctypes.c_char_p),
                ("cgroup_driver", ctypes.c_char_p), ("proc_root", ctypes.c_char_p),
                ("cgroup2_root", ctypes.c_char_p), ("extensions", ctypes.c_uint32), ("probe_timeout_ms", ctypes.c_uint32)]


EXT_CRI_RUNTIMES, EXT_CGROUP_V2, EXT_PROBE_KEEP_PRIVILEGES = 1, 2, 4  # GMP_EXT_*: off by default (section 7 of the header)


class gmp_ns_config(ctypes.Structure):
    _fields_ = [("cgroup", ctypes.c_int32), ("cgroup_file", ctypes.c_char_p), ("follow_context", ctypes.c_int32),
                ("gid", ctypes.c_int32), ("ipc", ctypes.c_int32), ("ipc_file", ctypes.c_char_p),
                ("mount", ctypes.c_int32), ("mount_file", ctypes.c_char_p), ("net", ctypes.c_int32),
                ("net_file", ctypes.c_char_p), ("no_fork", ctypes.c_int32), ("pid", ctypes.c_int32),
                ("pid_file", ctypes.c_char_p), ("preserve_credentials", ctypes.c_int32),
                ("root_directory", ctypes.c_char_p), ("target", ctypes.c_int32), ("uid", ctypes.c_int32),
                ("user", ctypes.c_int32), ("user_file", ctypes.c_char_p), ("uts", ctypes.c_int32),
                ("uts_file", ctypes.c_char_p), ("working_directory", ctypes.c_char_p)]


class gmp_process_info(ctypes.Structure):
    _fields_ = [("pid", ctypes.c_uint32), ("reserved", ctypes.c_uint32), ("used_gpu_memory", ctypes.c_uint64)]


class gmp_container_devices(ctypes.Structure):
    _fields_ = [("resource_name", ctypes.c_char_p), ("n_ids", ctypes.c_int32),
                ("device_ids", ctypes.POINTER(ctypes.c_char_p))]


class gmp_container_entry(ctypes.Structure):
    _fields_ = [("name", ctypes.c_char_p), ("n_devices", ctypes.c_int32),
                ("devices", ctypes.POINTER(gmp_container_devices))]


class gmp_pod_resources(ctypes.Structure):
    _fields_ = [("name", ctypes.c_char_p), ("ns", ctypes.c_char_p), ("n_containers", ctypes.c_int32),
                ("containers", ctypes.POINTER(gmp_container_entry))]


def enc(s: Optional[str]):
    return None if s is None else s.encode()


def c_strings(items: Sequence[str]):
    arr = (ctypes.c_char_p * max(1, len(items)))()
    for i, s in enumerate(items):
        arr[i] = s.encode()
    return arr


@dataclass
class ContainerResources:
    """Resources.Requests / Resources.Limits of one container; only cpu and memory take part in QoS."""
    cpu_request: Optional[str] = None
    memory_request: Optional[str] = None
    cpu_limit: Optional[str] = None
    memory_limit: Optional[str] = None


@dataclass
class Pod:
    """The fields of corev1.Pod the mount path reads."""
    name: str
    namespace: str = "default"
    uid: str = ""
    node_name: str = ""
    container_id: Optional[str] = None  # Status.ContainerStatuses[0].ContainerID
    containers: List[ContainerResources] = field(default_factory=lambda: [ContainerResources()])

    def to_c(self) -> gmp_pod:
        arr = (gmp_container_resources * max(1, len(self.containers)))()
        for i, r in enumerate(self.containers):
            arr[i] = gmp_container_resources(enc(r.cpu_request), enc(r.memory_request), enc(r.cpu_limit), enc(r.memory_limit))
        p = gmp_pod(enc(self.name), enc(self.namespace), enc(self.uid), enc(self.node_name), enc(self.container_id),
                    len(self.containers), arr)
        p._keepalive = arr
        return p


@dataclass
class MountEnv:
    """Where the node's files live; None = the reference's constants (see gmp_mount_env)."""
    devices_root: Optional[str] = None
    dev_prefix: Optional[str] = None
    cgroup_driver: Optional[str] = None
    proc_root: Optional[str] = None
    cgroup2_root: Optional[str] = None
    extensions: int = 0
    probe_timeout_ms: int = 0  # ProbeInPod: kill the probe after this long (0 = no limit)

    def to_c(self):
        return gmp_mount_env(enc(self.devices_root), enc(self.dev_prefix), enc(self.cgroup_driver), enc(self.proc_root),
                             enc(self.cgroup2_root), self.extensions, self.probe_timeout_ms)


def env_ptr(env: Optional[MountEnv]):
    if env is None:
        return None
    c = env.to_c()
    return ctypes.pointer(c)
