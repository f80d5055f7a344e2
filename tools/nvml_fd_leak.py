#!/usr/bin/env python
"""Which part of an NVML session leaves descriptors behind: dlopen/dlclose, or nvmlInit_v2/nvmlShutdown?
Driver-library behaviour, measured on the GPU box; decides how NvmlSession (csrc/gmp_nvml.cpp) holds the library."""
import ctypes
import json
import os
import _ctypes


def fds():
    out = {}
    for f in os.listdir("/proc/self/fd"):
        try:
            out[f] = os.readlink(f"/proc/self/fd/{f}")
        except OSError:
            pass
    return out


def diff(a, b):
    return sorted(v for k, v in b.items() if k not in a)


NAME = os.environ.get("GPUMOUNTER_NVML_LIB", "libnvidia-ml.so.1")
res = {}

# (a) load/unload only, no init
a = fds()
for _ in range(5):
    h = _ctypes.dlopen(NAME, os.RTLD_LAZY | os.RTLD_LOCAL)
    _ctypes.dlclose(h)
res["dlopen_dlclose_x5"] = diff(a, fds())

# (b) loaded once, init/shutdown five times
lib = ctypes.CDLL(NAME)
a = fds()
for _ in range(5):
    assert lib.nvmlInit_v2() == 0
    n = ctypes.c_uint()
    assert lib.nvmlDeviceGetCount_v2(ctypes.byref(n)) == 0
    assert lib.nvmlShutdown() == 0
res["init_shutdown_x5_loaded_once"] = diff(a, fds())

# (c) one outer init held, five nested init/shutdown pairs
assert lib.nvmlInit_v2() == 0
a = fds()
for _ in range(5):
    assert lib.nvmlInit_v2() == 0
    assert lib.nvmlShutdown() == 0
res["nested_init_shutdown_x5"] = diff(a, fds())
assert lib.nvmlShutdown() == 0

# (d) full cycle as the reference does it (nvml_dl.go): dlopen, init, shutdown, dlclose
a = fds()
for _ in range(5):
    h = _ctypes.dlopen(NAME, os.RTLD_LAZY | os.RTLD_LOCAL)
    l2 = ctypes.CDLL(NAME, handle=h)
    assert l2.nvmlInit_v2() == 0
    assert l2.nvmlShutdown() == 0
    _ctypes.dlclose(h)
res["full_cycle_x5_while_loaded_elsewhere"] = diff(a, fds())
print(json.dumps(res, indent=1))
