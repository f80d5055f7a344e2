"""ref: pkg/util/gpu/allocator/allocator.go — the selection logic (mount type, GPUs to remove, slave-pod naming).

Creating / polling / deleting slave pods through the k8s API (allocator.go:40-99,128-156,236-317) is the control
plane and stays with the Go worker (measured, not rebuilt).
"""

from __future__ import annotations

import ctypes
import os
from typing import List, Sequence

from . import device
from ._lib import check, lib
from ._model import c_strings
from .collector import GPUCollector
from .util import MOUNT_TYPE_NAME


class GPUAllocator(GPUCollector):
    """ref: allocator.go:23-25 (embeds the collector)"""

    def GetMountType(self, podName: str, namespace: str) -> str:  # ref: allocator.go:158-187
        arr = device.as_c_array(self.GPUList)
        t = ctypes.c_int()
        check(lib().gmp_get_mount_type(arr, len(self.GPUList), podName.encode(), namespace.encode(), ctypes.byref(t)),
              "GetMountType")
        return MOUNT_TYPE_NAME[t.value]

    def GetRemoveGPU(self, ownerPodName: str, ownerNamespace: str, uuids: Sequence[str]) -> List[device.NvidiaGPU]:
        # ref: allocator.go:101-126
        arr = device.as_c_array(self.GPUList)
        idx = (ctypes.c_int32 * max(1, len(self.GPUList)))()
        n = ctypes.c_int()
        check(lib().gmp_get_remove_gpu(arr, len(self.GPUList), ownerPodName.encode(), ownerNamespace.encode(),
                                       c_strings(uuids), len(uuids), idx, ctypes.byref(n)), "GetRemoveGPU")
        return [self.GPUList[idx[i]] for i in range(n.value)]


def newGPUSlavePodName(ownerPodName: str, randBytes: bytes = b"") -> str:  # ref: allocator.go:191-199
    rb = randBytes or os.urandom(3)
    buf = ctypes.create_string_buffer(512)
    check(lib().gmp_slave_pod_name(ownerPodName.encode(), (ctypes.c_ubyte * 3)(*rb[:3]), buf, 512), "newGPUSlavePodName")
    return buf.value.decode()
