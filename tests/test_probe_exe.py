"""The standalone gpumounter-probe executable and gmp_probe_all's pre-flight checks, without a GPU."""

import json
import os
import subprocess

import pytest

from gpumounter_b200 import _lib, probe

EXE = probe.probe_executable()
HAS_GPU = os.path.exists("/dev/nvidiactl")


def run(*args):
    p = subprocess.run([EXE, *args], capture_output=True, text=True, timeout=120)
    line = p.stdout.strip().splitlines()[-1] if p.stdout.strip() else "{}"
    return p.returncode, json.loads(line), p.stderr


def test_executable_is_self_contained():
    out = subprocess.run(["ldd", EXE], capture_output=True, text=True).stdout
    libs = [l.split()[0] for l in out.splitlines() if "=>" in l or l.strip().startswith("/")]
    # only the C library family: no libstdc++, no libcudart, no libcuda at link time
    assert all(any(k in l for k in ("libc.so", "ld-linux", "libm.so", "libdl", "libpthread", "librt", "linux-vdso"))
               for l in libs), libs


def test_glibc_floor_is_2_34():
    """The probe is exec'ed inside arbitrary pods: it must not need a newer libc than 2.34 (see probe_main.cpp)."""
    import re

    out = subprocess.run(["objdump", "-T", EXE], capture_output=True, text=True).stdout
    versions = sorted({tuple(int(x) for x in v.split(".")) for v in re.findall(r"GLIBC_([0-9.]+)", out)})
    assert versions and versions[-1] <= (2, 34), versions


def test_missing_nodes_have_distinct_codes(tmp_path):
    d = tmp_path / "dev"
    d.mkdir()
    rc, j, _ = run("--dev-dir", str(d), "--uuid", "GPU-x:3")
    assert rc == -_lib.GMP_ERR_NODE_CTL and j["status"] == _lib.GMP_ERR_NODE_CTL and "nvidiactl" in j["detail"]
    (d / "nvidiactl").write_text("")
    rc, j, _ = run("--dev-dir", str(d), "--uuid", "GPU-x:3")
    assert rc == -_lib.GMP_ERR_NODE_UVM and "nvidia-uvm" in j["detail"]
    (d / "nvidia-uvm").write_text("")
    rc, j, _ = run("--dev-dir", str(d), "--uuid", "GPU-x:3")
    assert rc == -_lib.GMP_ERR_NODE_GPU and j["detail"].startswith(str(d / "nvidia3"))
    (d / "nvidia3").write_text("")
    os.chmod(d / "nvidia3", 0)
    rc, j, _ = run("--dev-dir", str(d), "--uuid", "GPU-x:3")
    if os.geteuid() != 0:  # root ignores file modes; the device cgroup is what refuses root in a real pod
        assert rc == -_lib.GMP_ERR_NODE_DENIED
    else:
        assert rc in (-_lib.GMP_ERR_DRIVER_LIB, -_lib.GMP_ERR_UNAVAILABLE, -_lib.GMP_ERR_DEVICE_SET, 0) or HAS_GPU


def test_nodes_only_stops_before_cuda(tmp_path):
    """--nodes-only is the in-pod half of the two-stage hook (nodes checked in the pod, kernels run from the worker's
    resident process): it never initialises CUDA, so its verdict does not depend on a driver being present."""
    d = tmp_path / "dev"
    d.mkdir()
    for name in ("nvidiactl", "nvidia-uvm", "nvidia3"):
        (d / name).write_text("")
    rc, j, _ = run("--dev-dir", str(d), "--uuid", "GPU-x:3", "--uuid", "GPU-y:4", "--nodes-only")
    assert rc == -_lib.GMP_ERR_NODE_GPU and j["nodes_status"] == _lib.GMP_ERR_NODE_GPU and "nvidia4" in j["detail"]
    (d / "nvidia4").write_text("")
    rc, j, _ = run("--dev-dir", str(d), "--uuid", "GPU-x:3", "--uuid", "GPU-y:4", "--nodes-only")
    if rc == -_lib.GMP_ERR_DRIVER_LIB:  # no libcuda.so.1 on this machine: reported as such, still without CUDA
        assert j["nodes_status"] == _lib.GMP_ERR_DRIVER_LIB and "libcuda.so.1" in j["detail"]
    else:
        assert rc == 0 and j["status"] == 0 and j["nodes_status"] == 0
        assert [x["uuid"] for x in j["devices"]] == ["GPU-x", "GPU-y"] and j["hbm"] == [] and j["ms_cuda_init"] == 0


@pytest.mark.skipif(HAS_GPU, reason="box has a GPU")
def test_no_driver_is_reported_not_faked(tmp_path):
    rc, j, _ = run("--skip-nodes")
    assert rc == -_lib.GMP_ERR_UNAVAILABLE and j["n_gpus"] == 0 and j["hbm"] == []
    r = probe.probe_all(flags=probe.PROBE_F_SKIP_NODES, raise_on_error=False)
    assert r.status == _lib.GMP_ERR_UNAVAILABLE
    assert json.loads(probe.report_json(r))["status"] == _lib.GMP_ERR_UNAVAILABLE


def test_bad_arguments():
    p = subprocess.run([EXE, "--nonsense"], capture_output=True, text=True)
    assert p.returncode == 64 and "usage" in p.stderr
    p = subprocess.run([EXE, "--bytes", "12Q"], capture_output=True, text=True)
    assert p.returncode == 64
    with pytest.raises(probe.ProbeError) as e:
        probe.probe_p2p([])
    assert e.value.code == _lib.GMP_ERR_INVALID
    with pytest.raises(probe.ProbeError) as e:
        probe.probe_p2p(["GPU-a", ""])
    assert e.value.code == _lib.GMP_ERR_INVALID


def test_report_json_is_always_valid_json():
    """Whatever ends up in a report - non-finite rates, hostile strings, every counter at its maximum - the line parses."""
    import ctypes
    import math

    r = probe.gmp_report()
    ctypes.memset(ctypes.byref(r), 0xFF, ctypes.sizeof(r))           # every integer field at its extreme
    r.status, r.n = -7, 2
    r.detail = b'quote " backslash \\ newline \n tab \t bell \x07 <script> ' + "é€😀".encode() + b"\xff\xfe\xc3"
    for i in range(2):
        r.devices[i].uuid = b'GPU-"x"\\'
        r.devices[i].pci_bus_id = b"0000:1B:00.0"
        r.hbm[i].uuid = b"GPU-y"
        r.hbm[i].write_gbs_best, r.hbm[i].read_gbs_best, r.hbm[i].copy_gbs_best = math.inf, -math.inf, math.nan
        r.hbm[i].h2d_gbs = 1e308
    r.p2p.n = 2
    for k in range(4):
        r.p2p.write_gbs[k] = math.nan
    text = probe.report_json(r)
    j = json.loads(text)
    assert "\n" not in text and j["status"] == -7 and j["n_gpus"] == 2
    assert j["hbm"][0]["write_gbs"] is None and j["hbm"][0]["read_gbs"] is None and j["hbm"][0]["copy_gbs"] is None
    assert j["detail"].startswith('quote " backslash \\ newline \n tab \t bell \x07 <script>')
    assert j["detail"].endswith("é€😀\ufffd\ufffd\ufffd")      # well-formed UTF-8 kept, stray bytes replaced
    assert j["devices"][0]["uuid"] == 'GPU-"x"\\'


def test_slow_p2p_pairs_are_named_in_the_report():
    """NVSwitch: every pair should measure the same.  The JSON names the pairs under 80 % of their kind's median."""
    r = probe.gmp_report()
    r.n, r.p2p.n = 4, 4
    for i in range(4):
        for k in range(4):
            if i != k:
                r.p2p.write_gbs[i * 4 + k] = 713.0
                r.p2p.read_gbs[i * 4 + k] = 778.0
                r.p2p.memcpy_gbs[i * 4 + k] = 777.0
    r.p2p.write_gbs[2 * 4 + 1] = 52.0      # GPU 2 -> GPU 1 fell back to PCIe
    r.p2p.read_gbs[3 * 4 + 0] = 600.0      # 77 % of the median: a degraded link
    r.p2p.read_gbs[0 * 4 + 3] = 640.0      # 82 %: within tolerance
    j = json.loads(probe.report_json(r))
    assert j["p2p"]["slow_pairs"] == [
        {"src": 2, "dst": 1, "kind": "store", "gbs": 52.0, "median_gbs": 713.0},
        {"src": 3, "dst": 0, "kind": "load", "gbs": 600.0, "median_gbs": 778.0},
    ]
    r.p2p.write_gbs[2 * 4 + 1] = 713.0
    r.p2p.read_gbs[3 * 4 + 0] = 778.0
    assert json.loads(probe.report_json(r))["p2p"]["slow_pairs"] == []


def test_thresholds_turn_a_clean_report_into_slow():
    """The operator's acceptance policy, applied to a finished report: rates under the threshold are GMP_ERR_SLOW with the
    offender named; unmeasured rates (0) and an already failed report are left alone."""
    import ctypes

    f = _lib.lib().gmp_report_apply_thresholds
    f.argtypes = [ctypes.POINTER(probe.gmp_report), ctypes.c_double, ctypes.c_double]
    r = probe.gmp_report()
    r.n, r.p2p.n = 2, 2
    for i, (w, rd, c) in enumerate([(7100.0, 6840.0, 6930.0), (7090.0, 5100.0, 0.0)]):   # GPU 1 reads slowly, copy not run
        r.hbm[i].uuid = f"GPU-{i}".encode()
        r.hbm[i].write_gbs_best, r.hbm[i].read_gbs_best, r.hbm[i].copy_gbs_best = w, rd, c
    r.p2p.write_gbs[1], r.p2p.write_gbs[2], r.p2p.read_gbs[1], r.p2p.read_gbs[2] = 713.0, 712.0, 778.0, 300.0
    assert f(ctypes.byref(r), 0.0, 0.0) == 0 and f(ctypes.byref(r), 5000.0, 250.0) == 0
    assert f(ctypes.byref(r), 5500.0, 0.0) == _lib.GMP_ERR_SLOW
    assert r.status == _lib.GMP_ERR_SLOW and r.detail == b"GPU-1: read-verify 5100.0 GB/s is under the threshold of 5500.0"
    j = json.loads(probe.report_json(r))
    assert j["status"] == _lib.GMP_ERR_SLOW and j["verdict"] == "a measured rate is under the threshold"
    assert f(ctypes.byref(r), 1.0, 1.0) == _lib.GMP_ERR_SLOW           # a failed report stays failed
    r.status = 0
    assert f(ctypes.byref(r), 0.0, 700.0) == _lib.GMP_ERR_SLOW
    assert r.detail == b"GPU 1 -> GPU 0: peer load 300.0 GB/s is under the threshold of 700.0"
    p = subprocess.run([EXE, "--min-hbm-gbs"], capture_output=True, text=True)
    assert p.returncode == 64 and "needs a value" in p.stderr


def test_health_policy_on_a_finished_report():
    import ctypes

    f = _lib.lib().gmp_report_check_health
    f.argtypes = [ctypes.POINTER(probe.gmp_report)]
    r = probe.gmp_report()
    r.n = 3
    for i in range(3):
        r.devices[i].uuid = f"GPU-{i}".encode()
        ctypes.memset(ctypes.byref(r.health[i]), 0xFF, ctypes.sizeof(r.health[i]))      # NVML could not say: not judged
    assert f(ctypes.byref(r)) == 0
    r.health[0].throttle_reasons, r.health[0].ecc_uncorrected_volatile = 0x4 | 0x1, 0    # SW power cap + idle: fine
    r.health[1].throttle_reasons, r.health[1].ecc_uncorrected_volatile = 0, 0
    assert f(ctypes.byref(r)) == 0
    r.health[2].throttle_reasons, r.health[2].ecc_uncorrected_volatile = 0x40, 0
    assert f(ctypes.byref(r)) == _lib.GMP_ERR_UNHEALTHY and r.detail == b"GPU-2: hardware thermal slowdown is active"
    r.status = 0
    r.health[1].ecc_uncorrected_volatile = 3
    assert f(ctypes.byref(r)) == _lib.GMP_ERR_UNHEALTHY
    assert r.detail == b"GPU-1: 3 uncorrected ECC errors since the last driver load"
    assert json.loads(probe.report_json(r))["verdict"] == "NVML reports uncorrected ECC errors or a hardware slowdown"
