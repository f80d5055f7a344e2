"""ctypes view of oracle/libgmo_oracle.so — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
import this module; nothing under gpumounter_b200/ does.
"""

from __future__ import annotations

import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "libgmo_oracle.so")
UINT64_MAX = (1 << 64) - 1


class gmo_verify_result(ctypes.Structure):
    _fields_ = [
        ("mismatch_lanes", ctypes.c_uint64),
        ("first_bad_lane", ctypes.c_uint64),
        ("xor64", ctypes.c_uint64),
        ("sum64", ctypes.c_uint64),
    ]


_lib = None


def build() -> str:
    subprocess.run(["make", "-s", "-C", _HERE], check=True)
    return _PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_PATH):
            build()
        l = ctypes.CDLL(_PATH)
        l.gmo_splitmix64_next.restype = ctypes.c_uint64
        l.gmo_splitmix64_next.argtypes = [ctypes.POINTER(ctypes.c_uint64)]
        l.gmo_lane_words.argtypes = [ctypes.c_int, ctypes.c_uint64, ctypes.c_uint64,
                                     ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint64)]
        l.gmo_pattern_fill.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint64]
        l.gmo_pattern_verify.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint64,
                                         ctypes.POINTER(gmo_verify_result)]
        l.gmo_pattern_checksum.argtypes = [ctypes.c_size_t, ctypes.c_int, ctypes.c_uint64,
                                           ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint64)]
        l.gmo_cpu_probe.argtypes = [ctypes.c_size_t, ctypes.c_int, ctypes.c_uint64,
                                    ctypes.POINTER(ctypes.c_double * 3), ctypes.POINTER(gmo_verify_result)]
        _lib = l
    return _lib


def splitmix64_stream(seed: int, n: int):
    s = ctypes.c_uint64(seed)
    return [lib().gmo_splitmix64_next(ctypes.byref(s)) for _ in range(n)]


def lane_words(pattern: int, seed: int, lane: int):
    w0, w1 = ctypes.c_uint64(), ctypes.c_uint64()
    rc = lib().gmo_lane_words(pattern, seed & UINT64_MAX, lane, ctypes.byref(w0), ctypes.byref(w1))
    if rc != 0:
        raise ValueError(f"unknown pattern {pattern}")
    return w0.value, w1.value


def pattern_fill(nbytes: int, pattern: int, seed: int) -> bytearray:
    buf = bytearray(nbytes)
    if nbytes:
        c = (ctypes.c_char * nbytes).from_buffer(buf)
        if lib().gmo_pattern_fill(c, nbytes, pattern, seed & UINT64_MAX) != 0:
            raise ValueError(f"unknown pattern {pattern}")
    elif not 0 <= pattern < 5:
        raise ValueError(f"unknown pattern {pattern}")
    return buf


def pattern_verify(data, pattern: int, seed: int):
    """data: any writable buffer (bytearray / numpy array).  Returns (mismatch, first_bad, xor64, sum64)."""
    mv = memoryview(data).cast("B")
    n = mv.nbytes
    r = gmo_verify_result()
    if n:
        c = (ctypes.c_char * n).from_buffer(mv)
        rc = lib().gmo_pattern_verify(c, n, pattern, seed & UINT64_MAX, ctypes.byref(r))
    else:
        rc = lib().gmo_pattern_verify(None, 0, pattern, seed & UINT64_MAX, ctypes.byref(r))
    if rc != 0:
        raise ValueError(f"unknown pattern {pattern}")
    return r.mismatch_lanes, r.first_bad_lane, r.xor64, r.sum64


def pattern_checksum(nbytes: int, pattern: int, seed: int):
    x, s = ctypes.c_uint64(), ctypes.c_uint64()
    if lib().gmo_pattern_checksum(nbytes, pattern, seed & UINT64_MAX, ctypes.byref(x), ctypes.byref(s)) != 0:
        raise ValueError(f"unknown pattern {pattern}")
    return x.value, s.value


def cpu_probe(nbytes: int, threads: int, seed: int):
    """Timed fill/verify/copy on host threads.  Returns (seconds[3], (mismatch, first_bad, xor64, sum64))."""
    secs = (ctypes.c_double * 3)()
    r = gmo_verify_result()
    if lib().gmo_cpu_probe(nbytes, threads, seed & UINT64_MAX, ctypes.byref(secs), ctypes.byref(r)) != 0:
        raise MemoryError("gmo_cpu_probe")
    return list(secs), (r.mismatch_lanes, r.first_bad_lane, r.xor64, r.sum64)
