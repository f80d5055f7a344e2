"""The post-mount probe: ctypes mirror of sections 2-5 of include/gpumounter_probe.h.

No reference counterpart (the reference never touches the device it mounts); the hook
point is right after the mount loop, ref: pkg/server/gpu-mount/server.go:97-98.
"""

from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import List, Optional, Sequence

from . import _lib
from ._lib import ProbeError, check, lib

GMP_UUID_LEN = 96
GMP_PCI_LEN = 32
GMP_MAX_DEVICES = 16

PAT_SPLITMIX, PAT_ZERO, PAT_ONES, PAT_CHECKER, PAT_ADDRESS = range(5)
PAT_COUNT = 5
PATTERN_NAMES = ("splitmix", "zero", "ones", "checker", "address")

HBM_F_SKIP_COPY = 1 << 0
HBM_F_KEEP_DEVICE = 1 << 1
HBM_F_LIB_BASELINES = 1 << 2
HBM_F_RETAIN = 1 << 3  # keep buffers, streams and the pinned staging block for the next call on the device (implies KEEP_DEVICE)
UINT64_MAX = (1 << 64) - 1


class gmp_device(ctypes.Structure):
    _fields_ = [
        ("uuid", ctypes.c_char * GMP_UUID_LEN),
        ("minor", ctypes.c_uint32),
        ("pci_bus_id", ctypes.c_char * GMP_PCI_LEN),
        ("nvml_index", ctypes.c_int32),
        ("cuda_ordinal", ctypes.c_int32),
    ]


class gmp_verify_result(ctypes.Structure):
    _fields_ = [
        ("mismatch_lanes", ctypes.c_uint64),
        ("first_bad_lane", ctypes.c_uint64),
        ("xor64", ctypes.c_uint64),
        ("sum64", ctypes.c_uint64),
    ]


class gmp_hbm_opts(ctypes.Structure):
    _fields_ = [
        ("bytes", ctypes.c_uint64),
        ("iters", ctypes.c_uint32),
        ("pattern_mask", ctypes.c_uint32),
        ("seed", ctypes.c_uint64),
        ("dma_bytes", ctypes.c_uint64),
        ("flags", ctypes.c_uint32),
        ("reserved", ctypes.c_uint32),
    ]


class gmp_hbm_result(ctypes.Structure):
    _fields_ = [
        ("uuid", ctypes.c_char * GMP_UUID_LEN),
        ("cuda_ordinal", ctypes.c_int32),
        ("status", ctypes.c_int32),
        ("bytes", ctypes.c_uint64),
        ("iters", ctypes.c_uint32),
        ("patterns_run", ctypes.c_uint32),
        ("write_gbs_best", ctypes.c_double),
        ("write_gbs_median", ctypes.c_double),
        ("read_gbs_best", ctypes.c_double),
        ("read_gbs_median", ctypes.c_double),
        ("copy_gbs_best", ctypes.c_double),
        ("copy_gbs_median", ctypes.c_double),
        ("h2d_gbs", ctypes.c_double),
        ("d2h_gbs", ctypes.c_double),
        ("mismatch_lanes", ctypes.c_uint64),
        ("first_bad_lane", ctypes.c_uint64),
        ("first_bad_pattern", ctypes.c_int32),
        ("sm_count", ctypes.c_uint32),
        ("xor64", ctypes.c_uint64),
        ("sum64", ctypes.c_uint64),
        ("h2d_bytes", ctypes.c_uint64),
        ("d2h_bytes", ctypes.c_uint64),
        ("ns_select_device", ctypes.c_uint64),
        ("ns_context", ctypes.c_uint64),
        ("ns_alloc", ctypes.c_uint64),
        ("ns_first_kernel", ctypes.c_uint64),
        ("ns_kernels", ctypes.c_uint64),
        ("ns_dma", ctypes.c_uint64),
        ("ns_teardown", ctypes.c_uint64),
        ("ns_total", ctypes.c_uint64),
        ("mono_first_verified_ns", ctypes.c_uint64),
        ("lib_memset_gbs", ctypes.c_double),
        ("lib_memcpy_gbs", ctypes.c_double),
        ("mono_context_begin_ns", ctypes.c_uint64),
        ("mono_context_end_ns", ctypes.c_uint64),
    ]

    def as_dict(self) -> dict:
        d = {}
        for name, _ in self._fields_:
            v = getattr(self, name)
            d[name] = v.decode() if isinstance(v, bytes) else v
        return d


@dataclass
class Device:
    uuid: str
    minor: Optional[int]
    pci_bus_id: str
    nvml_index: int
    cuda_ordinal: int


@dataclass
class VerifyResult:
    mismatch_lanes: int
    first_bad_lane: int
    xor64: int
    sum64: int

    @property
    def ok(self) -> bool:
        return self.mismatch_lanes == 0


def _devices(arr, n) -> List[Device]:
    out = []
    for i in range(n):
        d = arr[i]
        out.append(
            Device(
                uuid=d.uuid.decode(),
                minor=None if d.minor == 0xFFFFFFFF else int(d.minor),
                pci_bus_id=d.pci_bus_id.decode(),
                nvml_index=int(d.nvml_index),
                cuda_ordinal=int(d.cuda_ordinal),
            )
        )
    return out


def _to_c_devices(devs: Sequence[Device]):
    arr = (gmp_device * max(1, len(devs)))()
    for i, d in enumerate(devs):
        arr[i].uuid = d.uuid.encode()
        arr[i].minor = 0xFFFFFFFF if d.minor is None else d.minor
        arr[i].pci_bus_id = d.pci_bus_id.encode()
        arr[i].nvml_index = d.nvml_index
        arr[i].cuda_ordinal = d.cuda_ordinal
    return arr


def abi_version() -> int:
    return lib().gmp_abi_version()


def enumerate_nvml() -> List[Device]:
    """NVML enumeration in index order (ref: pkg/util/gpu/collector/collector.go:40-79)."""
    arr = (gmp_device * GMP_MAX_DEVICES)()
    n = ctypes.c_int(0)
    check(lib().gmp_enumerate_nvml(arr, GMP_MAX_DEVICES, ctypes.byref(n)), "gmp_enumerate_nvml")
    return _devices(arr, n.value)


def enumerate_cuda() -> List[Device]:
    arr = (gmp_device * GMP_MAX_DEVICES)()
    n = ctypes.c_int(0)
    check(lib().gmp_enumerate_cuda(arr, GMP_MAX_DEVICES, ctypes.byref(n)), "gmp_enumerate_cuda")
    return _devices(arr, n.value)


def device_sets_equal(a: Sequence[Device], b: Sequence[Device]) -> bool:
    rc = lib().gmp_device_sets_equal(_to_c_devices(a), len(a), _to_c_devices(b), len(b))
    if rc == _lib.GMP_ERR_DEVICE_SET:
        return False
    check(rc, "gmp_device_sets_equal")
    return True


def _ptr(p) -> ctypes.c_void_p:
    return ctypes.c_void_p(int(p))


def hbm_write_pattern(dptr: int, nbytes: int, pattern: int = PAT_SPLITMIX, seed: int = 0, stream: int = 0) -> None:
    f = lib().gmp_hbm_write_pattern
    f.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint64, ctypes.c_void_p]
    check(f(_ptr(dptr), nbytes, pattern, seed & UINT64_MAX, _ptr(stream)), "gmp_hbm_write_pattern")


def hbm_read_verify(dptr: int, nbytes: int, pattern: int = PAT_SPLITMIX, seed: int = 0, stream: int = 0) -> VerifyResult:
    f = lib().gmp_hbm_read_verify
    f.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint64,
                  ctypes.POINTER(gmp_verify_result), ctypes.c_void_p]
    r = gmp_verify_result()
    check(f(_ptr(dptr), nbytes, pattern, seed & UINT64_MAX, ctypes.byref(r), _ptr(stream)), "gmp_hbm_read_verify")
    return VerifyResult(r.mismatch_lanes, r.first_bad_lane, r.xor64, r.sum64)


NOALLOC_MAX_BYTES = 8 << 30


def hbm_read_verify_noalloc(dptr: int, nbytes: int, pattern: int = PAT_SPLITMIX, seed: int = 0, stream: int = 0) -> VerifyResult:
    """hbm_read_verify() for callers that must not allocate device memory (a sidecar on a GPU the tenant has
    filled, a monitor that must not trigger cudaFree's device-wide synchronisation, a caller whose own rules forbid
    allocator calls at that point).  Scratch is a fixed array inside the library: at most NOALLOC_MAX_BYTES per call,
    one call at a time per device; make the first call on a device early (it loads the kernels) and before any
    stream capture."""
    f = lib().gmp_hbm_read_verify_noalloc
    f.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint64,
                  ctypes.POINTER(gmp_verify_result), ctypes.c_void_p]
    r = gmp_verify_result()
    check(f(_ptr(dptr), nbytes, pattern, seed & UINT64_MAX, ctypes.byref(r), _ptr(stream)), "gmp_hbm_read_verify_noalloc")
    return VerifyResult(r.mismatch_lanes, r.first_bad_lane, r.xor64, r.sum64)


def hbm_read_verify_async(dptr: int, nbytes: int, pattern: int, seed: int, d_acc: int, stream: int = 0) -> None:
    f = lib().gmp_hbm_read_verify_async
    f.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_void_p]
    check(f(_ptr(dptr), nbytes, pattern, seed & UINT64_MAX, _ptr(d_acc), _ptr(stream)), "gmp_hbm_read_verify_async")


def hbm_verify_fetch(d_acc: int, stream: int = 0) -> VerifyResult:
    f = lib().gmp_hbm_verify_fetch
    f.argtypes = [ctypes.c_void_p, ctypes.POINTER(gmp_verify_result), ctypes.c_void_p]
    r = gmp_verify_result()
    check(f(_ptr(d_acc), ctypes.byref(r), _ptr(stream)), "gmp_hbm_verify_fetch")
    return VerifyResult(r.mismatch_lanes, r.first_bad_lane, r.xor64, r.sum64)


def hbm_copy(dst: int, src: int, nbytes: int, stream: int = 0) -> None:
    f = lib().gmp_hbm_copy
    f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
    check(f(_ptr(dst), _ptr(src), nbytes, _ptr(stream)), "gmp_hbm_copy")


def hbm_corrupt(dptr: int, nbytes: int, byte_offset: int, mask: int, stream: int = 0) -> None:
    f = lib().gmp_hbm_corrupt
    f.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_uint8, ctypes.c_void_p]
    check(f(_ptr(dptr), nbytes, byte_offset, mask, _ptr(stream)), "gmp_hbm_corrupt")


def kernel_launches() -> int:
    f = lib().gmp_kernel_launches
    f.restype = ctypes.c_ulonglong
    return int(f())


def tune(knob: str, value: int = -1) -> int:
    f = lib().gmp_tune
    f.argtypes = [ctypes.c_char_p, ctypes.c_int]
    return check(f(knob.encode(), value), "gmp_tune")


def probe_hbm(uuid: Optional[str] = None, *, nbytes: int = 0, iters: int = 0, pattern_mask: int = 0, seed: int = 0,
              dma_bytes: int = UINT64_MAX, flags: int = 0, raise_on_error: bool = True) -> gmp_hbm_result:
    """One-shot probe of one mounted GPU (see gmp_probe_hbm in the header)."""
    f = lib().gmp_probe_hbm
    f.argtypes = [ctypes.c_char_p, ctypes.POINTER(gmp_hbm_opts), ctypes.POINTER(gmp_hbm_result)]
    o = gmp_hbm_opts(nbytes, iters, pattern_mask, seed & UINT64_MAX, dma_bytes, flags, 0)
    r = gmp_hbm_result()
    rc = f(uuid.encode() if uuid else None, ctypes.byref(o), ctypes.byref(r))
    if raise_on_error:
        check(rc, "gmp_probe_hbm")
    return r


def probe_hbm_release(uuid: Optional[str] = None) -> None:
    """Free what HBM_F_RETAIN calls left on one device (or on every device)."""
    f = lib().gmp_probe_hbm_release
    f.argtypes = [ctypes.c_char_p]
    check(f(uuid.encode() if uuid else None), "gmp_probe_hbm_release")


GMP_P2P_CELLS = GMP_MAX_DEVICES * GMP_MAX_DEVICES
P2P_F_KEEP_DEVICE, P2P_F_SKIP_BIDIR, P2P_F_SKIP_ALLTOALL, P2P_F_MATRIX_ONLY = 1, 2, 4, 8
PROBE_F_SKIP_NODES, PROBE_F_SKIP_HBM, PROBE_F_SKIP_P2P, PROBE_F_SKIP_NVML, PROBE_F_KEEP_DEVICE = 1, 2, 4, 8, 16
PROBE_F_NODES_ONLY, PROBE_F_HOST_VIEW = 32, 64


class gmp_p2p_opts(ctypes.Structure):
    _fields_ = [("bytes", ctypes.c_uint64), ("iters", ctypes.c_uint32), ("flags", ctypes.c_uint32),
                ("seed", ctypes.c_uint64)]


class gmp_p2p_result(ctypes.Structure):
    _fields_ = [
        ("n", ctypes.c_int32), ("status", ctypes.c_int32),
        ("uuid", (ctypes.c_char * GMP_UUID_LEN) * GMP_MAX_DEVICES),
        ("cuda_ordinal", ctypes.c_int32 * GMP_MAX_DEVICES),
        ("nvlink_active", ctypes.c_int32 * GMP_MAX_DEVICES),
        ("nvml_available", ctypes.c_int32), ("matrix_match", ctypes.c_int32),
        ("can_access", ctypes.c_int32 * GMP_P2P_CELLS),
        ("nvml_read", ctypes.c_int32 * GMP_P2P_CELLS), ("nvml_write", ctypes.c_int32 * GMP_P2P_CELLS),
        ("nvml_nvlink", ctypes.c_int32 * GMP_P2P_CELLS), ("nvml_atomics", ctypes.c_int32 * GMP_P2P_CELLS),
        ("write_gbs", ctypes.c_double * GMP_P2P_CELLS), ("read_gbs", ctypes.c_double * GMP_P2P_CELLS),
        ("memcpy_gbs", ctypes.c_double * GMP_P2P_CELLS), ("atomics_ok", ctypes.c_int32 * GMP_P2P_CELLS),
        ("bidir_gbs", ctypes.c_double * GMP_MAX_DEVICES), ("alltoall_gbs", ctypes.c_double),
        ("bytes", ctypes.c_uint64), ("mismatch_lanes", ctypes.c_uint64),
        ("first_bad_src", ctypes.c_int32), ("first_bad_dst", ctypes.c_int32),
        ("ns_setup", ctypes.c_uint64), ("ns_pairs", ctypes.c_uint64), ("ns_bidir", ctypes.c_uint64),
        ("ns_alltoall", ctypes.c_uint64), ("ns_teardown", ctypes.c_uint64), ("ns_total", ctypes.c_uint64),
    ]

    def matrix(self, name: str):
        n = self.n
        flat = getattr(self, name)
        return [[flat[i * n + j] for j in range(n)] for i in range(n)]


class gmp_gpu_health(ctypes.Structure):
    _fields_ = [("sm_mhz", ctypes.c_uint32), ("mem_mhz", ctypes.c_uint32), ("sm_max_mhz", ctypes.c_uint32),
                ("mem_max_mhz", ctypes.c_uint32), ("temperature_c", ctypes.c_uint32), ("power_mw", ctypes.c_uint32),
                ("throttle_reasons", ctypes.c_uint64), ("ecc_uncorrected_volatile", ctypes.c_uint64),
                ("ecc_corrected_volatile", ctypes.c_uint64), ("remapped_rows_correctable", ctypes.c_uint32),
                ("remapped_rows_uncorrectable", ctypes.c_uint32), ("remap_pending", ctypes.c_uint32),
                ("remap_failure", ctypes.c_uint32), ("nvlink_crc_flit_errors", ctypes.c_uint64),
                ("nvlink_crc_data_errors", ctypes.c_uint64), ("nvlink_replay_errors", ctypes.c_uint64),
                ("nvlink_recovery_errors", ctypes.c_uint64)]


def nvml_health(uuid: str) -> gmp_gpu_health:
    """Clocks, throttle reasons, temperature, power and ECC counters of one GPU (unknown fields are all-ones)."""
    h = gmp_gpu_health()
    check(lib().gmp_nvml_health(uuid.encode(), ctypes.byref(h)), "gmp_nvml_health")
    return h


class gmp_probe_opts(ctypes.Structure):
    _fields_ = [("hbm", gmp_hbm_opts), ("p2p", gmp_p2p_opts), ("flags", ctypes.c_uint32), ("reserved", ctypes.c_uint32),
                ("dev_dir", ctypes.c_char_p), ("also_visible", ctypes.POINTER(gmp_device)), ("n_also_visible", ctypes.c_int32),
                ("reserved2", ctypes.c_int32)]


class gmp_report(ctypes.Structure):
    _fields_ = [
        ("status", ctypes.c_int32), ("n", ctypes.c_int32), ("nodes_status", ctypes.c_int32),
        ("device_set_status", ctypes.c_int32), ("nvml_status", ctypes.c_int32), ("reserved", ctypes.c_int32),
        ("devices", gmp_device * GMP_MAX_DEVICES), ("hbm", gmp_hbm_result * GMP_MAX_DEVICES), ("p2p", gmp_p2p_result),
        ("health", gmp_gpu_health * GMP_MAX_DEVICES),
        ("ns_nodes", ctypes.c_uint64), ("ns_enumerate", ctypes.c_uint64), ("ns_hbm", ctypes.c_uint64),
        ("ns_p2p", ctypes.c_uint64), ("ns_total", ctypes.c_uint64), ("ns_cuda_init", ctypes.c_uint64),
        ("ns_nvml", ctypes.c_uint64), ("mono_start_ns", ctypes.c_uint64), ("mono_usable_ns", ctypes.c_uint64),
        ("detail", ctypes.c_char * 256),
    ]


def probe_p2p(uuids: Sequence[str], *, nbytes: int = 0, iters: int = 0, flags: int = 0, seed: int = 0,
              raise_on_error: bool = True) -> gmp_p2p_result:
    """Capability matrices (CUDA vs NVML) and NVLink bandwidth across the GPUs named by UUID."""
    f = lib().gmp_probe_p2p
    f.argtypes = [ctypes.POINTER(ctypes.c_char_p), ctypes.c_int, ctypes.POINTER(gmp_p2p_opts), ctypes.POINTER(gmp_p2p_result)]
    arr = (ctypes.c_char_p * max(1, len(uuids)))(*[u.encode() for u in uuids])
    o = gmp_p2p_opts(nbytes, iters, flags, seed & UINT64_MAX)
    r = gmp_p2p_result()
    rc = f(arr, len(uuids), ctypes.byref(o), ctypes.byref(r))
    if raise_on_error:
        check(rc, "gmp_probe_p2p")
    return r


def peer_access(device: int, peer: int, enable: bool = True) -> None:
    check(lib().gmp_peer_access(device, peer, int(enable)), "gmp_peer_access")


def probe_all(expected: Sequence[Device] = (), *, hbm_bytes: int = 0, hbm_iters: int = 0, pattern_mask: int = 0,
              dma_bytes: int = UINT64_MAX, p2p_bytes: int = 0, p2p_iters: int = 0, flags: int = 0, hbm_flags: int = 0,
              p2p_flags: int = 0, dev_dir: Optional[str] = None, also_visible: Sequence[Device] = (),
              raise_on_error: bool = True) -> gmp_report:
    """The whole-pod probe the worker's hook runs after AddGPU's mount loop (ref: server.go:97-98)."""
    f = lib().gmp_probe_all
    f.argtypes = [ctypes.POINTER(gmp_device), ctypes.c_int, ctypes.POINTER(gmp_probe_opts), ctypes.POINTER(gmp_report)]
    o = gmp_probe_opts()
    o.hbm = gmp_hbm_opts(hbm_bytes, hbm_iters, pattern_mask, 0, dma_bytes, hbm_flags, 0)
    o.p2p = gmp_p2p_opts(p2p_bytes, p2p_iters, p2p_flags, 0)
    o.flags = flags
    o.dev_dir = dev_dir.encode() if dev_dir else None
    also = _to_c_devices(also_visible) if also_visible else None  # GPUs the pod already holds: visible, not probed
    o.also_visible = also
    o.n_also_visible = len(also_visible)
    r = gmp_report()
    rc = f(_to_c_devices(expected) if expected else None, len(expected), ctypes.byref(o), ctypes.byref(r))
    if raise_on_error and rc < 0:
        raise ProbeError(rc, r.detail.decode(errors="replace"), "gmp_probe_all")
    return r


def report_json(report: gmp_report) -> str:
    buf = ctypes.create_string_buffer(1 << 16)
    check(lib().gmp_report_json(ctypes.byref(report), buf, 1 << 16), "gmp_report_json")
    return buf.value.decode()


def probe_executable() -> str:
    """Path of the standalone gpumounter-probe binary built next to the library."""
    import os
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "gpumounter-probe")


__all__ = [n for n in dir() if not n.startswith("_")] + ["ProbeError"]
