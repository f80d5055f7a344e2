"""The dlopen NVML layer against a fake libnvidia-ml (CPU only): enumeration order, strings, processes,
P2P status, fault injection.  Mirrors what the reference's nvml_test.go / collector_test.go poke at by hand."""

import ctypes
import os

import pytest

from gpumounter_b200 import ProbeError, _lib, collector, device, lib, probe
from tests.conftest import FAKE_UUIDS


def test_enumeration_is_in_nvml_index_order_with_minor_and_uuid(fake_nvml):
    gpus = fake_nvml()
    devs = probe.enumerate_nvml()
    assert [(d.uuid, d.minor, d.pci_bus_id) for d in devs] == [tuple(g) for g in gpus]
    assert [d.nvml_index for d in devs] == list(range(8))
    assert all(d.cuda_ordinal == -1 for d in devs)


def test_collector_get_gpu_info_builds_the_reference_records(fake_nvml):
    gpus = fake_nvml()
    c = collector.GPUCollector()
    c.GetGPUInfo()  # ref: collector.go:40-79
    assert [(g.MinorNumber, g.UUID, g.DeviceFilePath, g.State) for g in c.GPUList] == [
        (m, u, f"/dev/nvidia{m}", "GPU_FREE_STATE") for u, m, _ in gpus]
    assert c.GetGPUByUUID(FAKE_UUIDS[2]).MinorNumber == 1
    with pytest.raises(ProbeError) as e:
        c.GetGPUByUUID("GPU-nope")
    assert e.value.code == _lib.GMP_ERR_NO_DEVICE and "No GPU with UUIDGPU-nope" in e.value.detail


def test_no_devices_and_capacity(fake_nvml):
    fake_nvml(gpus=[])
    assert probe.enumerate_nvml() == []
    fake_nvml()
    arr = (probe.gmp_device * 4)()
    n = ctypes.c_int()
    rc = lib().gmp_enumerate_nvml(arr, 4, ctypes.byref(n))
    assert rc == _lib.GMP_ERR_BUFFER_TOO_SMALL and n.value == 8


def test_library_not_found_message_matches_the_reference(monkeypatch):
    monkeypatch.setenv("GPUMOUNTER_NVML_LIB", "/nonexistent/libnvidia-ml.so.1")
    with pytest.raises(ProbeError) as e:
        probe.enumerate_nvml()
    assert e.value.code == _lib.GMP_ERR_NVML_NOT_FOUND
    assert e.value.detail.startswith("could not load NVML library")  # ref: bindings.go:53-55


@pytest.mark.parametrize("fn", ["nvmlInit_v2", "nvmlDeviceGetCount_v2", "nvmlDeviceGetHandleByIndex_v2",
                                "nvmlDeviceGetMinorNumber", "nvmlDeviceGetUUID"])
def test_nvml_errors_surface_as_nvml_colon_message(fake_nvml, fn):
    fake_nvml(fail=[(fn, 15)])  # NVML_ERROR_GPU_IS_LOST
    with pytest.raises(ProbeError) as e:
        probe.enumerate_nvml()
    assert e.value.code == _lib.GMP_ERR_NVML
    assert e.value.detail.startswith("nvml: GPU is lost")  # ref: bindings.go:47-48


def test_versions(fake_nvml):
    fake_nvml(extra=["driver 580.159", "nvml 12.580.159"])
    d, n = ctypes.create_string_buffer(80), ctypes.create_string_buffer(80)
    assert lib().gmp_nvml_versions(d, 80, n, 80) == 0
    assert (d.value, n.value) == (b"580.159", b"12.580.159")


def test_running_processes_graphics_first_then_compute(fake_nvml):
    u = FAKE_UUIDS
    fake_nvml(procs=[(u[0], 111, 1 << 20, "compute"), (u[0], 222, 2 << 20, "graphics"), (u[1], 333, 5, "compute"),
                     (u[0], 444, 7, "compute")])
    g = device.New(3, u[0])
    procs = g.GetRunningProcess()  # ref: nvidia.go:58-87 appends graphics, then compute
    assert [(p.Pid, p.UsedGPUMemory) for p in procs] == [(222, 2 << 20), (111, 1 << 20), (444, 7)]
    assert device.New(9, u[5]).GetRunningProcess() == []
    with pytest.raises(ProbeError) as e:
        device.New(9, "GPU-unknown").GetRunningProcess()
    assert e.value.code == _lib.GMP_ERR_NVML and "Not Found" in e.value.detail


def test_p2p_status_matrix(fake_nvml):
    fake_nvml(p2p=[(0, 2, 0, 5), (2, 0, 1, 3)])  # READ 0->2 NOT_SUPPORTED(5); WRITE 2->0 GPU_NOT_SUPPORTED(2)+1
    uuids = FAKE_UUIDS[:3]
    arr = (ctypes.c_char_p * 3)(*[u.encode() for u in uuids])
    st = (ctypes.c_int32 * 9)()
    assert lib().gmp_nvml_p2p_status(arr, 3, 0, st) == 0
    assert list(st) == [0, 0, 5, 0, 0, 0, 0, 0, 0]
    assert lib().gmp_nvml_p2p_status(arr, 3, 1, st) == 0
    assert list(st) == [0, 0, 0, 0, 0, 0, 3, 0, 0]
    assert lib().gmp_nvml_p2p_status(arr, 3, 9, st) == _lib.GMP_ERR_INVALID


def test_nvlink_active_links(fake_nvml):
    fake_nvml(nvlink=[(FAKE_UUIDS[0], 18), (FAKE_UUIDS[1], 4)])
    a, m = ctypes.c_int(), ctypes.c_int()
    assert lib().gmp_nvml_nvlink_active(FAKE_UUIDS[0].encode(), ctypes.byref(a), ctypes.byref(m)) == 0
    assert (a.value, m.value) == (18, 18)  # toolkit nvml.h: NVML_NVLINK_MAX_LINKS 18 (the reference's header says 12)
    assert lib().gmp_nvml_nvlink_active(FAKE_UUIDS[1].encode(), ctypes.byref(a), None) == 0 and a.value == 4


def test_device_set_comparison(fake_nvml):
    fake_nvml()
    nv = probe.enumerate_nvml()
    cu = [probe.Device(d.uuid, None, "", -1, i) for i, d in enumerate(reversed(nv))]  # other order, minors unknown
    assert probe.device_sets_equal(nv, cu)
    assert not probe.device_sets_equal(nv, cu[:-1])
    cu[0] = probe.Device("GPU-other", None, "", -1, 0)
    assert not probe.device_sets_equal(nv, cu)
    wrong_minor = [probe.Device(d.uuid, (d.minor + 1) % 8, "", d.nvml_index, -1) for d in nv]
    assert not probe.device_sets_equal(nv, wrong_minor)


def test_health_reports_clocks_reasons_and_ecc(fake_nvml):
    fake_nvml(extra=[f"health {FAKE_UUIDS[1]} 1357 3996 83 999600 4 2"])   # sw_power_cap, two uncorrected ECC errors
    h = probe.nvml_health(FAKE_UUIDS[0])
    assert (h.sm_mhz, h.mem_mhz, h.sm_max_mhz, h.mem_max_mhz, h.throttle_reasons, h.ecc_uncorrected_volatile) == (
        1965, 3996, 1965, 3996, 0, 0)
    h = probe.nvml_health(FAKE_UUIDS[1])
    assert (h.sm_mhz, h.temperature_c, h.power_mw, h.throttle_reasons, h.ecc_uncorrected_volatile, h.ecc_corrected_volatile) == (
        1357, 83, 999600, 4, 2, 0)
    # a query NVML refuses leaves its field "unknown" instead of failing the whole call
    fake_nvml(fail=[("nvmlDeviceGetTemperature", 3), ("nvmlDeviceGetTotalEccErrors", 3)])
    h = probe.nvml_health(FAKE_UUIDS[0])
    assert h.temperature_c == 0xFFFFFFFF and h.ecc_uncorrected_volatile == (1 << 64) - 1 and h.sm_mhz == 1965
    with pytest.raises(ProbeError) as e:
        probe.nvml_health("GPU-unknown")
    assert e.value.code == _lib.GMP_ERR_NVML


def _maps_count(needle):
    with open("/proc/self/maps") as f:
        return len({line.split()[-1] for line in f if needle in line})


def test_sessions_reuse_one_library_image_and_leak_no_descriptors(fake_nvml):
    """The driver's libnvidia-ml leaves descriptors behind per fresh image (profiles/r01_nvml_fd_leak.json), so the
    image is loaded once per process; nvmlInit_v2/nvmlShutdown still bracket every call, so a changed device list
    is seen by the next call."""
    import os

    fake_nvml()
    probe.enumerate_nvml()
    fds0 = len(os.listdir("/proc/self/fd"))
    for i in range(50):
        gpus = fake_nvml(gpus=[(FAKE_UUIDS[j], j, f"00000000:{0x10 + j:02X}:00.0") for j in range(1 + i % 8)])
        assert [d.uuid for d in probe.enumerate_nvml()] == [g[0] for g in gpus]
        d, n = ctypes.create_string_buffer(80), ctypes.create_string_buffer(80)
        assert lib().gmp_nvml_versions(d, 80, n, 80) == 0
    assert len(os.listdir("/proc/self/fd")) == fds0
    assert _maps_count("libfake_nvml") == 1


def test_hold_makes_nested_sessions_share_one_attach(fake_nvml):
    fake_nvml()
    cold = ctypes.CDLL(os.environ["GPUMOUNTER_NVML_LIB"]).fake_nvml_cold_inits
    cold.restype = ctypes.c_uint
    before = cold()
    for _ in range(3):
        probe.enumerate_nvml()
    assert cold() - before == 3                      # as the reference: one attach per query
    before = cold()
    assert lib().gmp_nvml_hold(1) == 0
    assert lib().gmp_nvml_hold(1) == 0               # holds nest
    for _ in range(3):
        probe.enumerate_nvml()
    assert lib().gmp_nvml_hold(0) == 0
    probe.enumerate_nvml()
    assert lib().gmp_nvml_hold(0) == 0
    assert cold() - before == 1
    assert lib().gmp_nvml_hold(0) == _lib.GMP_ERR_INVALID     # unmatched release
    probe.enumerate_nvml()
    assert cold() - before == 2                      # released: the next query attaches again (fresh device list)


def test_hold_without_nvml_is_an_error_and_changes_nothing(monkeypatch):
    monkeypatch.setenv("GPUMOUNTER_NVML_LIB", "/nonexistent/libnvidia-ml.so.1")
    assert lib().gmp_nvml_hold(1) == _lib.GMP_ERR_NVML_NOT_FOUND
    assert lib().gmp_nvml_hold(0) == _lib.GMP_ERR_INVALID
    from gpumounter_b200 import util
    with util.NvmlAttached() as h:
        assert h.held is False


def test_health_reports_hbm_row_remapping(fake_nvml):
    """nvmlDeviceGetRemappedRows: what the HBM has already repaired, and whether a repair is pending or has failed."""
    fake_nvml(extra=[f"remap {FAKE_UUIDS[2]} 3 1 1 0", f"remap {FAKE_UUIDS[3]} 0 8 0 1"])
    h = probe.nvml_health(FAKE_UUIDS[0])
    assert (h.remapped_rows_correctable, h.remapped_rows_uncorrectable, h.remap_pending, h.remap_failure) == (0, 0, 0, 0)
    h = probe.nvml_health(FAKE_UUIDS[2])
    assert (h.remapped_rows_correctable, h.remapped_rows_uncorrectable, h.remap_pending, h.remap_failure) == (3, 1, 1, 0)
    # the health policy on a report carrying these samples
    check = lib().gmp_report_check_health
    check.argtypes = [ctypes.POINTER(probe.gmp_report)]
    r = probe.gmp_report()
    r.n = 2
    for i, u in enumerate((FAKE_UUIDS[0], FAKE_UUIDS[2])):
        r.devices[i].uuid = u.encode()
        r.health[i] = probe.nvml_health(u)
    assert check(ctypes.byref(r)) == _lib.GMP_ERR_UNHEALTHY
    assert r.detail.decode() == f"{FAKE_UUIDS[2]}: an HBM row remap is pending: the GPU needs a reset"
    r.status = 0
    r.health[1] = probe.nvml_health(FAKE_UUIDS[3])
    r.devices[1].uuid = FAKE_UUIDS[3].encode()
    assert check(ctypes.byref(r)) == _lib.GMP_ERR_UNHEALTHY and b"row remapping has failed" in r.detail
    # an older driver without the entry point (or a refusal): unknown, and not judged
    fake_nvml(fail=[("nvmlDeviceGetRemappedRows", 3)])
    h = probe.nvml_health(FAKE_UUIDS[0])
    assert h.remap_pending == 0xFFFFFFFF and h.remap_failure == 0xFFFFFFFF
    r.status, r.n = 0, 1
    r.health[0] = h
    assert check(ctypes.byref(r)) == 0


def test_health_reports_nvlink_error_totals(fake_nvml):
    fake_nvml(extra=[f"nvlinkerr {FAKE_UUIDS[4]} 7 0 12 1"])
    h = probe.nvml_health(FAKE_UUIDS[4])
    assert (h.nvlink_crc_flit_errors, h.nvlink_crc_data_errors, h.nvlink_replay_errors, h.nvlink_recovery_errors) == (7, 0, 12, 1)
    h = probe.nvml_health(FAKE_UUIDS[0])
    assert (h.nvlink_crc_flit_errors, h.nvlink_replay_errors) == (0, 0)
    fake_nvml(fail=[("nvmlDeviceGetFieldValues", 3)])
    assert probe.nvml_health(FAKE_UUIDS[0]).nvlink_replay_errors == (1 << 64) - 1     # unknown, not zero
