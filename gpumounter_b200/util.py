"""ref: pkg/util/util.go — node-local mount / unmount orchestration, over the C ABI."""

from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence

from ._lib import check, lib
from ._model import MountEnv, Pod, env_ptr
from .device import NvidiaGPU, as_c_array

# ref: pkg/util/gpu/types.go:22-27
EntireMount, SingleMount, NoMount, UnknownMount = "entire-mount", "single-mount", "no-mount", "unknown-mount"
MOUNT_TYPE_CODE = {EntireMount: 0, SingleMount: 1, NoMount: 2, UnknownMount: 3}
MOUNT_TYPE_NAME = {v: k for k, v in MOUNT_TYPE_CODE.items()}


def MountGPU(pod: Pod, gpu: NvidiaGPU, env: Optional[MountEnv] = None) -> None:  # ref: util.go:17-71
    c = pod.to_c()
    check(lib().gmp_mount_gpu(env_ptr(env), ctypes.byref(c), ctypes.byref(gpu._c)), "MountGPU")


def UnmountGPU(pod: Pod, gpu: NvidiaGPU, forceRemove: bool, env: Optional[MountEnv] = None) -> None:  # ref: util.go:73-147
    c = pod.to_c()
    check(lib().gmp_unmount_gpu(env_ptr(env), ctypes.byref(c), ctypes.byref(gpu._c), int(forceRemove)), "UnmountGPU")


def MountGPUs(pod: Pod, gpus: Sequence[NvidiaGPU], env: Optional[MountEnv] = None) -> int:
    """Batched form of the loop at ref: pkg/server/gpu-mount/server.go:80-95 (no reference counterpart)."""
    c = pod.to_c()
    done = ctypes.c_int(0)
    check(lib().gmp_mount_gpus(env_ptr(env), ctypes.byref(c), as_c_array(gpus), len(gpus), ctypes.byref(done)), "MountGPUs")
    return done.value


def UnmountGPUs(pod: Pod, gpus: Sequence[NvidiaGPU], forceRemove: bool, env: Optional[MountEnv] = None) -> int:
    c = pod.to_c()
    done = ctypes.c_int(0)
    check(lib().gmp_unmount_gpus(env_ptr(env), ctypes.byref(c), as_c_array(gpus), len(gpus), int(forceRemove),
                                 ctypes.byref(done)), "UnmountGPUs")
    return done.value


def GetPodGPUProcesses(pod: Pod, gpu: NvidiaGPU, env: Optional[MountEnv] = None) -> Optional[List[str]]:  # ref: util.go:152-196
    cap = 2048
    arr = ((ctypes.c_char * 16) * cap)()
    n = ctypes.c_int()
    c = pod.to_c()
    check(lib().gmp_get_pod_gpu_processes(env_ptr(env), ctypes.byref(c), ctypes.byref(gpu._c), arr, cap, ctypes.byref(n)),
          "GetPodGPUProcesses")
    return None if n.value == 0 else [arr[i].value.decode() for i in range(n.value)]


def ContainString(stringList: Sequence[str], aimString: str) -> bool:  # ref: util.go:198-205
    return any(s == aimString for s in stringList)


def CanMount(mountType: str, isEntireMount: bool) -> bool:  # ref: util.go:207-226 (request.IsEntireMount passed directly)
    return bool(lib().gmp_can_mount(MOUNT_TYPE_CODE[mountType], int(isEntireMount)))


class NvmlAttached:
    """`with NvmlAttached():` keeps NVML attached for the block (gmp_nvml_hold), so that the per-GPU process queries made
    inside it do not each pay a cold nvmlInit_v2.  Best effort: where NVML cannot be attached the block runs anyway and
    the queries report the error themselves, exactly as without the hold."""

    def __enter__(self):
        self.held = lib().gmp_nvml_hold(1) == 0
        return self

    def __exit__(self, *exc):
        if self.held:
            lib().gmp_nvml_hold(0)
        return False


def CheckNodesInPod(pod: Pod, gpus: Sequence[NvidiaGPU], env: Optional[MountEnv] = None, raise_on_error: bool = True):
    """Do /dev/nvidiactl, /dev/nvidia-uvm and the GPUs' nodes open from inside the pod (its device cgroup, its mount
    namespace)?  Nothing of the pod's image is executed.  Returns (status, index of the failing GPU or -1).
    No reference counterpart; see gmp_check_nodes_in_pod."""
    failed = ctypes.c_int(-1)
    c = pod.to_c()
    rc = lib().gmp_check_nodes_in_pod(env_ptr(env), ctypes.byref(c), as_c_array(gpus), len(gpus), ctypes.byref(failed))
    if raise_on_error:
        check(rc, "CheckNodesInPod")
    return rc, failed.value


def VerifyGoneInPod(pod: Pod, gpus: Sequence[NvidiaGPU], env: Optional[MountEnv] = None, raise_on_error: bool = True):
    """After UnmountGPU(s): none of the GPUs' nodes may exist in the pod any more; checked by a helper in the pod's mount
    namespace, nothing executed there.  Returns (status, index of the first GPU still present or -1)."""
    failed = ctypes.c_int(-1)
    c = pod.to_c()
    rc = lib().gmp_verify_gone_in_pod(env_ptr(env), ctypes.byref(c), as_c_array(gpus), len(gpus), ctypes.byref(failed))
    if raise_on_error:
        check(rc, "VerifyGoneInPod")
    return rc, failed.value


def ProbeInPod(pod: Pod, gpus: Sequence[NvidiaGPU], extra_args: Sequence[str] = (), env: Optional[MountEnv] = None,
               probe_exe: Optional[str] = None, raise_on_error: bool = True):
    """Run gpumounter-probe inside the pod (device cgroup + mount namespace) and return (status, report dict).
    No reference counterpart; this is the hook after ref: pkg/server/gpu-mount/server.go:97."""
    import json as _json

    from . import probe as _probe
    from ._model import c_strings

    exe = probe_exe or _probe.probe_executable()
    buf = ctypes.create_string_buffer(1 << 16)
    code = ctypes.c_int(-1)
    c = pod.to_c()
    rc = lib().gmp_probe_in_pod(env_ptr(env), ctypes.byref(c), as_c_array(gpus), len(gpus), exe.encode(),
                                c_strings(list(extra_args)), len(extra_args), buf, 1 << 16, ctypes.byref(code))
    text = buf.value.decode(errors="replace")
    report = _json.loads(text) if text.startswith("{") else {}
    if raise_on_error:
        check(rc, "ProbeInPod")
    return rc, report


def ProbeInPodFanout(pod: Pod, gpus: Sequence[NvidiaGPU], extra_args: Sequence[str] = (), env: Optional[MountEnv] = None,
                     probe_exe: Optional[str] = None, raise_on_error: bool = True):
    """ProbeInPod with one probe process per GPU, each confined to its own GPU by a child device cgroup, then one
    whole-pod process for the device-set check and the P2P step (gmp_probe_in_pod_fanout; no reference counterpart).
    Returns (status, report) where the report has ProbeInPod's shape - "hbm", "devices", "health" gathered from the per-GPU
    reports in the order of `gpus`; device-set / NVML status, "p2p" and "ms_p2p" from the whole-pod process; "ms_cuda_init"
    the slowest per-GPU one - plus the raw "reports" and "whole_pod"."""
    import json as _json

    from . import probe as _probe
    from ._model import c_strings

    exe = probe_exe or _probe.probe_executable()
    buf = ctypes.create_string_buffer(1 << 18)
    code = ctypes.c_int(-1)
    c = pod.to_c()
    rc = lib().gmp_probe_in_pod_fanout(env_ptr(env), ctypes.byref(c), as_c_array(gpus), len(gpus), exe.encode(),
                                       c_strings(list(extra_args)), len(extra_args), buf, 1 << 18, ctypes.byref(code))
    text = buf.value.decode(errors="replace")
    rep = _json.loads(text) if text.startswith("{") else {}
    if rep:
        subs = [r for r in rep.get("reports", []) if r]
        whole = rep.get("whole_pod") or {}
        for key in ("hbm", "devices", "health"):
            rep[key] = [x for r in subs for x in r.get(key, [])]
        for key in ("ms_cuda_init", "ms_nvml", "ms_nodes", "ms_hbm"):
            rep[key] = max([r.get(key, 0.0) for r in subs], default=0.0)
        rep["ms_to_usable"] = (rep["mono_usable_ns"] - rep["mono_start_ns"]) / 1e6 if rep.get("mono_usable_ns") else 0.0
        rep["nodes_status"] = min([r.get("nodes_status", 0) for r in subs], default=0)
        for key in ("device_set_status", "nvml_status"):
            rep[key] = whole.get(key, min([r.get(key, 0) for r in subs], default=0))
        if "p2p" in whole:
            rep["p2p"] = whole["p2p"]
        rep["ms_p2p"] = whole.get("ms_p2p", 0.0)
    if raise_on_error:
        check(rc, "ProbeInPodFanout")
    return rc, rep


def VerifyUnmountedInPod(pod: Pod, gpus: Sequence[NvidiaGPU], env: Optional[MountEnv] = None, dev_dir: Optional[str] = None,
                         probe_exe: Optional[str] = None, raise_on_error: bool = True):
    """After UnmountGPU(s): from inside the pod, none of the GPUs' device nodes may exist any more.  No reference
    counterpart (it returns Success once `rm` exited 0, ref: util.go:131-134).  Returns (status, report dict)."""
    import json as _json

    from . import probe as _probe
    from ._model import c_strings

    args = []
    for g in gpus:
        args += ["--gone", f"{g.UUID}:{g.MinorNumber}"]
    if dev_dir:
        args += ["--dev-dir", dev_dir]
    exe = probe_exe or _probe.probe_executable()
    buf = ctypes.create_string_buffer(1 << 16)
    code = ctypes.c_int(-1)
    c = pod.to_c()
    rc = lib().gmp_probe_in_pod(env_ptr(env), ctypes.byref(c), None, 0, exe.encode(), c_strings(args), len(args), buf, 1 << 16,
                                ctypes.byref(code))
    text = buf.value.decode(errors="replace")
    report = _json.loads(text) if text.startswith("{") else {}
    if raise_on_error:
        check(rc, "VerifyUnmountedInPod")
    return rc, report
