"""ctypes view of oracle/gmo_reference.c — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Python-friendly wrappers over the plain-C restatement of the reference's Go functions.  Function names
follow the reference (snake-cased) so that tests read like its call sites.
"""

from __future__ import annotations

import ctypes
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

from . import gmo_ctypes

E_OK, E_PANIC_NAME, E_DRIVER, E_SLICE, E_BUF, E_QUANTITY, E_TARGET, E_IO, E_EXEC, E_PANIC_INDEX, E_BUSY, E_ATOI = (
    0, -1, -2, -3, -4, -5, -6, -7, -8, -9, -10, -11)
MOUNT_ENTIRE, MOUNT_SINGLE, MOUNT_NONE, MOUNT_UNKNOWN = 0, 1, 2, 3
QOS_NAMES = {0: "Guaranteed", 1: "Burstable", 2: "BestEffort"}


class OracleError(Exception):
    def __init__(self, code: int, msg: str = ""):
        self.code = code
        super().__init__(f"oracle error {code}: {msg}")


class gmo_gpu(ctypes.Structure):
    _fields_ = [("minor_number", ctypes.c_int), ("device_file_path", ctypes.c_char * 64), ("uuid", ctypes.c_char * 96),
                ("state", ctypes.c_char * 24), ("pod_name", ctypes.c_char * 256), ("ns", ctypes.c_char * 256)]


class gmo_container_resources(ctypes.Structure):
    _fields_ = [("cpu_request", ctypes.c_char_p), ("memory_request", ctypes.c_char_p),
                ("cpu_limit", ctypes.c_char_p), ("memory_limit", ctypes.c_char_p)]


class gmo_pod(ctypes.Structure):
    _fields_ = [("name", ctypes.c_char_p), ("ns", ctypes.c_char_p), ("uid", ctypes.c_char_p),
                ("node_name", ctypes.c_char_p), ("container_id", ctypes.c_char_p), ("n_containers", ctypes.c_int),
                ("containers", ctypes.POINTER(gmo_container_resources))]


class gmo_nsenter_config(ctypes.Structure):
    _fields_ = [("cgroup", ctypes.c_int), ("cgroup_file", ctypes.c_char_p), ("follow_context", ctypes.c_int),
                ("gid", ctypes.c_int), ("ipc", ctypes.c_int), ("ipc_file", ctypes.c_char_p),
                ("mount", ctypes.c_int), ("mount_file", ctypes.c_char_p), ("net", ctypes.c_int),
                ("net_file", ctypes.c_char_p), ("no_fork", ctypes.c_int), ("pid", ctypes.c_int),
                ("pid_file", ctypes.c_char_p), ("preserve_credentials", ctypes.c_int),
                ("root_directory", ctypes.c_char_p), ("target", ctypes.c_int), ("uid", ctypes.c_int),
                ("user", ctypes.c_int), ("user_file", ctypes.c_char_p), ("uts", ctypes.c_int),
                ("uts_file", ctypes.c_char_p), ("working_directory", ctypes.c_char_p)]


class gmo_container_devices(ctypes.Structure):
    _fields_ = [("resource_name", ctypes.c_char_p), ("n_ids", ctypes.c_int),
                ("device_ids", ctypes.POINTER(ctypes.c_char_p))]


class gmo_container_res_entry(ctypes.Structure):
    _fields_ = [("name", ctypes.c_char_p), ("n_devices", ctypes.c_int),
                ("devices", ctypes.POINTER(gmo_container_devices))]


class gmo_pod_resources(ctypes.Structure):
    _fields_ = [("name", ctypes.c_char_p), ("ns", ctypes.c_char_p), ("n_containers", ctypes.c_int),
                ("containers", ctypes.POINTER(gmo_container_res_entry))]


class gmo_mount_env(ctypes.Structure):
    _fields_ = [("devices_root", ctypes.c_char_p), ("dev_prefix", ctypes.c_char_p), ("cgroup_driver", ctypes.c_char_p)]


def _b(s: Optional[str]):
    return None if s is None else s.encode()


@dataclass
class Resources:
    cpu_request: Optional[str] = None
    memory_request: Optional[str] = None
    cpu_limit: Optional[str] = None
    memory_limit: Optional[str] = None


@dataclass
class Pod:
    name: str = "gpu-pod"
    namespace: str = "default"
    uid: str = "1234-abcd-5678-efgh"
    node_name: str = "node-1"
    container_id: str = "docker://" + "ab" * 32
    containers: List[Resources] = field(default_factory=lambda: [Resources()])

    def c(self):
        arr = (gmo_container_resources * max(1, len(self.containers)))()
        for i, r in enumerate(self.containers):
            arr[i] = gmo_container_resources(_b(r.cpu_request), _b(r.memory_request), _b(r.cpu_limit), _b(r.memory_limit))
        p = gmo_pod(_b(self.name), _b(self.namespace), _b(self.uid), _b(self.node_name), _b(self.container_id),
                    len(self.containers), arr)
        p._keep = arr
        return p


@dataclass
class GPU:
    minor_number: int
    device_file_path: str
    uuid: str
    state: str
    pod_name: str
    namespace: str


def _lib():
    l = gmo_ctypes.lib()
    if getattr(l, "_ref_ready", False):
        return l
    l.gmo_mount_type_string.restype = ctypes.c_char_p
    l.gmo_qos_string.restype = ctypes.c_char_p
    l.gmo_nsenter_argv.argtypes = [ctypes.POINTER(gmo_nsenter_config), ctypes.c_char_p, ctypes.POINTER(ctypes.c_char_p),
                                   ctypes.c_int, ctypes.POINTER(ctypes.c_char_p), ctypes.c_int, ctypes.c_char_p,
                                   ctypes.c_size_t]
    l._ref_ready = True
    return l


def _from_c(g: gmo_gpu) -> GPU:
    return GPU(g.minor_number, g.device_file_path.decode(), g.uuid.decode(), g.state.decode(), g.pod_name.decode(),
               g.ns.decode())


def _to_c_list(gpus: Sequence[GPU]):
    arr = (gmo_gpu * max(1, len(gpus)))()
    for i, g in enumerate(gpus):
        arr[i].minor_number = g.minor_number
        arr[i].device_file_path = g.device_file_path.encode()
        arr[i].uuid = g.uuid.encode()
        arr[i].state = g.state.encode()
        arr[i].pod_name = g.pod_name.encode()
        arr[i].ns = g.namespace.encode()
    return arr


def device_new(minor: int, uuid: str) -> GPU:
    g = gmo_gpu()
    _lib().gmo_device_new(minor, uuid.encode(), ctypes.byref(g))
    return _from_c(g)


def device_string(g: GPU) -> str:
    buf = ctypes.create_string_buffer(2048)
    n = _lib().gmo_device_string(ctypes.byref(_to_c_list([g])[0]), buf, 2048)
    if n < 0:
        raise OracleError(E_BUF)
    return buf.value.decode()


def pod_qos(pod: Pod) -> str:
    rc = _lib().gmo_pod_qos(ctypes.byref(pod.c()))
    if rc < 0:
        raise OracleError(rc, "quantity")
    return QOS_NAMES[rc]


def quantity_cmp(a: str, b: str) -> int:
    out = ctypes.c_int()
    rc = _lib().gmo_quantity_cmp(a.encode(), b.encode(), ctypes.byref(out))
    if rc != 0:
        raise OracleError(rc, f"quantity {a!r} / {b!r}")
    return out.value


def _strs(parts: Sequence[str]):
    arr = (ctypes.c_char_p * max(1, len(parts)))()
    for i, p in enumerate(parts):
        arr[i] = p.encode()
    return arr


def new_cgroup_name_check(component: str) -> int:
    return _lib().gmo_new_cgroup_name_check(component.encode())


def expand_slice(s: str) -> str:
    buf = ctypes.create_string_buffer(4096)
    rc = _lib().gmo_expand_slice(s.encode(), buf, 4096)
    if rc != 0:
        raise OracleError(rc, f"invalid slice name: {s}")
    return buf.value.decode()


def to_systemd(parts: Sequence[str]) -> str:
    buf = ctypes.create_string_buffer(4096)
    rc = _lib().gmo_to_systemd(_strs(parts), len(parts), buf, 4096)
    if rc != 0:
        raise OracleError(rc)
    return buf.value.decode()


def to_cgroupfs(parts: Sequence[str]) -> str:
    buf = ctypes.create_string_buffer(4096)
    rc = _lib().gmo_to_cgroupfs(_strs(parts), len(parts), buf, 4096)
    if rc != 0:
        raise OracleError(rc)
    return buf.value.decode()


def get_cgroup_driver(env_value: Optional[str]) -> str:
    out = ctypes.create_string_buffer(64)
    err = ctypes.create_string_buffer(256)
    rc = _lib().gmo_get_cgroup_driver(_b(env_value), out, 64, err, 256)
    if rc != 0:
        raise OracleError(rc, err.value.decode())
    return out.value.decode()


def get_cgroup_name(driver: str, pod: Pod, container_id: str) -> str:
    buf = ctypes.create_string_buffer(4096)
    rc = _lib().gmo_get_cgroup_name(driver.encode(), ctypes.byref(pod.c()), container_id.encode(), buf, 4096)
    if rc != 0:
        raise OracleError(rc)
    return buf.value.decode()


def device_group_path(cgroup_path: str, devices_root: Optional[str] = None) -> str:
    buf = ctypes.create_string_buffer(8192)
    rc = _lib().gmo_device_group_path(_b(devices_root), cgroup_path.encode(), buf, 8192)
    if rc != 0:
        raise OracleError(rc)
    return buf.value.decode()


def get_cgroup_pids(devices_root: str, cgroup_path: str) -> List[str]:
    cap = 4096
    arr = ((ctypes.c_char * 16) * cap)()
    n = _lib().gmo_get_cgroup_pids(devices_root.encode(), cgroup_path.encode(), arr, cap)
    if n < 0:
        raise OracleError(n, "open cgroup.procs")
    return [arr[i].value.decode() for i in range(min(n, cap))]


def permission_command(cgroup_path: str, minor: int, allow: bool, devices_root: Optional[str] = None) -> str:
    buf = ctypes.create_string_buffer(8192)
    rc = _lib().gmo_permission_command(_b(devices_root), cgroup_path.encode(), minor, 1 if allow else 0, buf, 8192)
    if rc != 0:
        raise OracleError(rc)
    return buf.value.decode()


def nsenter_argv(program: str, args: Sequence[str], **cfg) -> List[str]:
    c = gmo_nsenter_config()
    for k, v in cfg.items():
        setattr(c, k, v.encode() if isinstance(v, str) else int(v))
    argv = (ctypes.c_char_p * 64)()
    pool = ctypes.create_string_buffer(16384)
    n = _lib().gmo_nsenter_argv(ctypes.byref(c), program.encode(), _strs(args), len(args), argv, 64, pool, 16384)
    if n < 0:
        raise OracleError(n, "Target must be specified " if n == E_TARGET else "")
    return [argv[i].decode() for i in range(n)]


def mknod_command(minor: int, dev_prefix: Optional[str] = None) -> str:
    buf = ctypes.create_string_buffer(1024)
    _lib().gmo_mknod_command(_b(dev_prefix), minor, buf, 1024)
    return buf.value.decode()


def rm_command(minor: int, dev_prefix: Optional[str] = None) -> str:
    buf = ctypes.create_string_buffer(1024)
    _lib().gmo_rm_command(_b(dev_prefix), minor, buf, 1024)
    return buf.value.decode()


def kill_command(pids: Sequence[str]) -> str:
    buf = ctypes.create_string_buffer(65536)
    _lib().gmo_kill_command(_strs(pids), len(pids), buf, 65536)
    return buf.value.decode()


def can_mount(mount_type: int, is_entire: bool) -> bool:
    return bool(_lib().gmo_can_mount(mount_type, 1 if is_entire else 0))


def pod_gpu_processes(cgroup_pids: Sequence[str], gpu_pids: Sequence[int]) -> Optional[List[str]]:
    arr = (ctypes.c_uint * max(1, len(gpu_pids)))(*gpu_pids)
    out = (ctypes.c_int * max(1, len(gpu_pids)))()
    n = _lib().gmo_pod_gpu_processes(_strs(cgroup_pids), len(cgroup_pids), arr, len(gpu_pids), out)
    return None if n == 0 else [str(gpu_pids[out[i]]) for i in range(n)]


def update_gpu_status(gpus: Sequence[GPU], pod_resources) -> List[GPU]:
    """pod_resources: list of (name, ns, [(container, [(resource_name, [ids])])]). Raises on an unknown UUID."""
    keep = []
    pods = (gmo_pod_resources * max(1, len(pod_resources)))()
    for i, (name, ns, containers) in enumerate(pod_resources):
        cs = (gmo_container_res_entry * max(1, len(containers)))()
        for j, (cname, devices) in enumerate(containers):
            ds = (gmo_container_devices * max(1, len(devices)))()
            for k, (rname, ids) in enumerate(devices):
                ida = _strs(ids)
                keep.append(ida)
                ds[k] = gmo_container_devices(rname.encode(), len(ids), ida)
            keep.append(ds)
            cs[j] = gmo_container_res_entry(cname.encode(), len(devices), ds)
        keep.append(cs)
        pods[i] = gmo_pod_resources(name.encode(), ns.encode(), len(containers), cs)
    arr = _to_c_list(gpus)
    rc = _lib().gmo_update_gpu_status(arr, len(gpus), pods, len(pod_resources))
    out = [_from_c(arr[i]) for i in range(len(gpus))]
    if rc != 0:
        raise OracleError(rc, "No GPU with UUID")
    return out


def pod_gpu_resources(gpus: Sequence[GPU], pod_name: str, ns: str) -> List[int]:
    out = (ctypes.c_int * max(1, len(gpus)))()
    n = _lib().gmo_pod_gpu_resources(_to_c_list(gpus), len(gpus), pod_name.encode(), ns.encode(), out)
    return [out[i] for i in range(n)]


def get_mount_type(gpus: Sequence[GPU], pod_name: str, ns: str) -> int:
    return _lib().gmo_get_mount_type(_to_c_list(gpus), len(gpus), pod_name.encode(), ns.encode())


def mount_type_string(t: int) -> str:
    return _lib().gmo_mount_type_string(t).decode()


def get_remove_gpu(gpus: Sequence[GPU], owner: str, owner_ns: str, uuids: Sequence[str]) -> List[int]:
    out = (ctypes.c_int * max(1, len(gpus)))()
    n = _lib().gmo_get_remove_gpu(_to_c_list(gpus), len(gpus), owner.encode(), owner_ns.encode(), _strs(uuids),
                                  len(uuids), out)
    return [out[i] for i in range(n)]


def slave_pod_name(owner: str, rand3: bytes) -> str:
    buf = ctypes.create_string_buffer(512)
    rc = _lib().gmo_slave_pod_name(owner.encode(), (ctypes.c_ubyte * 3)(*rand3), buf, 512)
    if rc != 0:
        raise OracleError(rc)
    return buf.value.decode()


def slave_pod_count(gpu_num: int, is_entire: bool):
    per = ctypes.c_int()
    n = _lib().gmo_slave_pod_count(gpu_num, 1 if is_entire else 0, ctypes.byref(per))
    return n, per.value


def add_result_http(result: int, pod: str, node: str):
    buf = ctypes.create_string_buffer(2048)
    code = _lib().gmo_add_result_http(result, pod.encode(), node.encode(), buf, 2048)
    return code, buf.value.decode()


def remove_result_http(result: int, pod: str, node: str, uuids: Sequence[str]):
    buf = ctypes.create_string_buffer(4096)
    code = _lib().gmo_remove_result_http(result, pod.encode(), node.encode(), _strs(uuids), len(uuids), buf, 4096)
    return code, buf.value.decode()


def _env(devices_root, dev_prefix, cgroup_driver):
    return gmo_mount_env(_b(devices_root), _b(dev_prefix), _b(cgroup_driver))


def mount_gpu(pod: Pod, gpu: GPU, *, devices_root: str, dev_prefix: str, cgroup_driver: Optional[str]) -> int:
    err = ctypes.create_string_buffer(4096)
    env = _env(devices_root, dev_prefix, cgroup_driver)
    rc = _lib().gmo_mount_gpu(ctypes.byref(env), ctypes.byref(pod.c()), ctypes.byref(_to_c_list([gpu])[0]), err, 4096)
    if rc != 0:
        raise OracleError(rc, err.value.decode(errors="replace"))
    return rc


def unmount_gpu(pod: Pod, gpu: GPU, force: bool, gpu_pids: Sequence[int], *, devices_root: str, dev_prefix: str,
                cgroup_driver: Optional[str]) -> int:
    err = ctypes.create_string_buffer(4096)
    env = _env(devices_root, dev_prefix, cgroup_driver)
    arr = (ctypes.c_uint * max(1, len(gpu_pids)))(*gpu_pids)
    rc = _lib().gmo_unmount_gpu(ctypes.byref(env), ctypes.byref(pod.c()), ctypes.byref(_to_c_list([gpu])[0]),
                                1 if force else 0, arr, len(gpu_pids), err, 4096)
    if rc != 0:
        raise OracleError(rc, err.value.decode(errors="replace"))
    return rc
