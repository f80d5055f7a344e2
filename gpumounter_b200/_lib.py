"""ctypes loader for libgpumounter_probe.so.  Fails loudly when the library is missing."""

from __future__ import annotations

import ctypes
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_NAME = "libgpumounter_probe.so"

# status codes (include/gpumounter_probe.h section 1)
GMP_OK = 0
GMP_ERR_INVALID = -1
GMP_ERR_UNAVAILABLE = -2
GMP_ERR_NO_DEVICE = -3
GMP_ERR_CUDA = -4
GMP_ERR_NVML = -5
GMP_ERR_NVML_NOT_FOUND = -6
GMP_ERR_VERIFY = -7
GMP_ERR_NOMEM = -8
GMP_ERR_SYS = -9
GMP_ERR_BUFFER_TOO_SMALL = -10
GMP_ERR_DEVICE_SET = -11
GMP_ERR_P2P_MATRIX = -12
GMP_ERR_CGROUP_DRIVER = -13
GMP_ERR_NAME = -14
GMP_ERR_EMPTY_CGROUP = -15
GMP_ERR_TARGET = -16
GMP_ERR_BUSY = -17
GMP_ERR_NODE_CTL = -18
GMP_ERR_NODE_UVM = -19
GMP_ERR_NODE_GPU = -20
GMP_ERR_NODE_DENIED = -21
GMP_ERR_DRIVER_LIB = -22
GMP_ERR_STILL_MOUNTED = -23
GMP_ERR_UNSUPPORTED = -24
GMP_ERR_SLOW = -25
GMP_ERR_UNHEALTHY = -26
GMP_ERR_TIMEOUT = -27


class ProbeError(RuntimeError):
    """A GMP_ERR_* status from the C ABI, with the library's detail string."""

    def __init__(self, code: int, detail: str = "", where: str = ""):
        self.code = code
        self.detail = detail
        name = ""
        try:
            name = lib().gmp_strerror(code).decode()
        except Exception:  # pragma: no cover - only when the library itself is broken
            pass
        msg = f"{where + ': ' if where else ''}{name} ({code})"
        if detail:
            msg += f": {detail}"
        super().__init__(msg)


_lock = threading.Lock()
_lib = None


def lib_path() -> str:
    return os.environ.get("GPUMOUNTER_PROBE_LIB", os.path.join(_HERE, _LIB_NAME))


def lib() -> ctypes.CDLL:
    """The loaded shared library.  There is no fallback: a missing build is an error."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            path = lib_path()
            if not os.path.exists(path):
                raise ImportError(
                    f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                    f"or `make -C gpumounter_b200/csrc`.  gpumounter_b200 has no CPU fallback."
                )
            handle = ctypes.CDLL(path)
            handle.gmp_strerror.restype = ctypes.c_char_p
            handle.gmp_strerror.argtypes = [ctypes.c_int]
            handle.gmp_last_error.restype = ctypes.c_char_p
            handle.gmp_last_error.argtypes = []
            handle.gmp_abi_version.restype = ctypes.c_int
            _lib = handle
    return _lib


def check(rc: int, where: str = "") -> int:
    if rc < 0:
        raise ProbeError(rc, lib().gmp_last_error().decode(errors="replace"), where)
    return rc
