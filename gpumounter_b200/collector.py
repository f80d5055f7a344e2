"""ref: pkg/util/gpu/collector/collector.go — the GPU list and its ownership bookkeeping, over the C ABI.

The kubelet PodResources gRPC call itself (collector.go:165-194) is control-plane plumbing outside the hot
path's scope; UpdateGPUStatus takes the already fetched listing.
"""

from __future__ import annotations

import ctypes
from typing import List, Sequence, Tuple

from . import device, probe
from ._lib import check, lib
from ._model import (c_strings, gmp_container_devices, gmp_container_entry, gmp_gpu, gmp_pod_resources)

SocketPath = "/var/lib/kubelet/pod-resources/kubelet.sock"  # ref: types.go:6-7
NvidiaResourceName = "nvidia.com/gpu"  # ref: types.go:10
GPUPoolNamespace = "gpu-pool"  # ref: types.go:18

# listing = [(pod_name, namespace, [(container_name, [(resource_name, [device_ids])])])]
Listing = Sequence[Tuple[str, str, Sequence[Tuple[str, Sequence[Tuple[str, Sequence[str]]]]]]]


class GPUCollector:
    """ref: collector.go:19-21"""

    def __init__(self, gpus: Sequence[device.NvidiaGPU] = ()):
        self.GPUList: List[device.NvidiaGPU] = list(gpus)

    def GetGPUInfo(self) -> None:  # ref: collector.go:40-79 — NVML index order: minor, UUID
        for d in probe.enumerate_nvml():
            self.GPUList.append(device.New(d.minor, d.uuid))

    def GetGPUByUUID(self, uuid: str) -> device.NvidiaGPU:  # ref: collector.go:81-88
        arr = device.as_c_array(self.GPUList)
        idx = ctypes.c_int()
        check(lib().gmp_get_gpu_by_uuid(arr, len(self.GPUList), uuid.encode(), ctypes.byref(idx)), "GetGPUByUUID")
        return self.GPUList[idx.value]

    def UpdateGPUStatus(self, listing: Listing) -> None:  # ref: collector.go:113-135
        keep = []
        pods = (gmp_pod_resources * max(1, len(listing)))()
        for i, (name, ns, containers) in enumerate(listing):
            cs = (gmp_container_entry * max(1, len(containers)))()
            for j, (cname, devices) in enumerate(containers):
                ds = (gmp_container_devices * max(1, len(devices)))()
                for k, (rname, ids) in enumerate(devices):
                    ida = c_strings(ids)
                    keep.append(ida)
                    ds[k] = gmp_container_devices(rname.encode(), len(ids), ida)
                keep.append(ds)
                cs[j] = gmp_container_entry(cname.encode(), len(devices), ds)
            keep.append(cs)
            pods[i] = gmp_pod_resources(name.encode(), ns.encode(), len(containers), cs)
        arr = device.as_c_array(self.GPUList)
        rc = lib().gmp_update_gpu_status(arr, len(self.GPUList), pods, len(listing))
        for i, g in enumerate(self.GPUList):  # the in-place update is visible even when an unknown UUID aborts it
            ctypes.memmove(ctypes.byref(g._c), ctypes.byref(arr[i]), ctypes.sizeof(gmp_gpu))
        check(rc, "UpdateGPUStatus")

    def GetPodGPUResources(self, podName: str, namespace: str) -> List[device.NvidiaGPU]:  # ref: collector.go:149-163
        arr = device.as_c_array(self.GPUList)
        idx = (ctypes.c_int32 * max(1, len(self.GPUList)))()
        n = ctypes.c_int()
        check(lib().gmp_get_pod_gpu_resources(arr, len(self.GPUList), podName.encode(), namespace.encode(), idx,
                                              ctypes.byref(n)), "GetPodGPUResources")
        return [self.GPUList[idx[i]] for i in range(n.value)]
