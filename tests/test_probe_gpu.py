"""The one-shot probe and the device enumeration on a real box (GPU only)."""

import json
import os
import subprocess
import sys

import pytest

from gpumounter_b200 import ProbeError, _lib, probe
from oracle import gmo_ctypes as gmo

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cuda_enumeration_matches_torch(cuda):
    torch = cuda
    devs = probe.enumerate_cuda()
    assert len(devs) == torch.cuda.device_count() >= 1
    for d in devs:
        assert d.uuid == "GPU-" + str(torch.cuda.get_device_properties(d.cuda_ordinal).uuid)


def test_nvml_enumeration_equals_cuda_device_set(cuda):
    nv = probe.enumerate_nvml()
    cu = probe.enumerate_cuda()
    assert [d.nvml_index for d in nv] == list(range(len(nv)))
    for d in nv:
        assert d.uuid.startswith("GPU-") and len(d.uuid) == 40 and d.minor is not None
    visible = os.environ.get("CUDA_VISIBLE_DEVICES")
    if visible is None and len(nv) == len(cu):
        assert probe.device_sets_equal(nv, cu)
    else:  # CUDA may see a subset of the node
        assert {d.uuid for d in cu} <= {d.uuid for d in nv}
    assert not probe.device_sets_equal(nv, cu[:-1] if len(cu) == len(nv) else nv[:-1])


def test_probe_hbm_in_process_keep_device(cuda):
    nbytes = (256 << 20) + 48
    mask = (1 << probe.PAT_COUNT) - 1
    r = probe.probe_hbm(nbytes=nbytes, iters=2, pattern_mask=mask, seed=77, dma_bytes=8 << 20,
                        flags=probe.HBM_F_KEEP_DEVICE | probe.HBM_F_LIB_BASELINES)
    assert r.lib_memset_gbs > 100 and r.lib_memcpy_gbs > 100
    assert r.status == 0 and r.mismatch_lanes == 0 and r.first_bad_pattern == -1
    assert r.patterns_run == mask and r.bytes == nbytes and r.sm_count == 148
    assert r.uuid.decode() == probe.enumerate_cuda()[0].uuid
    for f in ("write_gbs_best", "read_gbs_best", "copy_gbs_best", "h2d_gbs", "d2h_gbs"):
        assert getattr(r, f) > 1.0, f
    assert r.h2d_bytes == 8 << 20 and r.d2h_bytes >= 8 << 20
    assert r.ns_total >= r.ns_kernels > 0
    # the context must still be usable by the caller (KEEP_DEVICE)
    t = cuda.ones(16, device="cuda")
    assert float(t.sum()) == 16.0


def test_probe_hbm_checksum_is_the_oracle_checksum(cuda):
    nbytes = (64 << 20) + 5
    r = probe.probe_hbm(nbytes=nbytes, iters=3, seed=1000, dma_bytes=0, flags=probe.HBM_F_KEEP_DEVICE)
    # last verify pass is the copy of the buffer written with seed + iters - 1
    assert (r.xor64, r.sum64) == gmo.pattern_checksum(nbytes, probe.PAT_SPLITMIX, 1002)


def test_probe_hbm_by_uuid_and_unknown_uuid(cuda):
    uuid = probe.enumerate_cuda()[-1].uuid
    r = probe.probe_hbm(uuid, nbytes=32 << 20, iters=1, dma_bytes=0, flags=probe.HBM_F_KEEP_DEVICE)
    assert r.uuid.decode() == uuid and r.cuda_ordinal == len(probe.enumerate_cuda()) - 1
    with pytest.raises(ProbeError) as e:
        probe.probe_hbm("GPU-00000000-0000-0000-0000-000000000000", nbytes=1 << 20, flags=probe.HBM_F_KEEP_DEVICE)
    assert e.value.code == _lib.GMP_ERR_NO_DEVICE


def test_probe_hbm_one_shot_subprocess_resets_the_device():
    """Default flags: the call ends with cudaDeviceReset, as the leak-free contract wants.
    Run in a child so that this process's torch context is not torn down."""
    code = (
        "import json,sys; sys.path.insert(0, %r)\n"
        "from gpumounter_b200 import probe\n"
        "r = probe.probe_hbm(nbytes=1<<30, iters=2)\n"
        "r2 = probe.probe_hbm(nbytes=1<<28, iters=1)\n"
        "print(json.dumps([r.as_dict(), r2.as_dict()]))\n" % ROOT
    )
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    r, r2 = json.loads(out.stdout.strip().splitlines()[-1])
    for x in (r, r2):
        assert x["status"] == 0 and x["mismatch_lanes"] == 0
        assert x["write_gbs_best"] > 100 and x["read_gbs_best"] > 100 and x["copy_gbs_best"] > 100
    assert r["h2d_bytes"] == 16 << 20


def test_probe_p2p_single_gpu_is_a_trivial_matrix(cuda):
    uuid = probe.enumerate_cuda()[0].uuid
    r = probe.probe_p2p([uuid], nbytes=64 << 20, flags=probe.P2P_F_KEEP_DEVICE)
    assert r.status == 0 and r.n == 1 and r.matrix("can_access") == [[1]] and r.mismatch_lanes == 0
    assert r.nvml_available == 1 and r.matrix_match == 1 and r.matrix("nvml_read") == [[0]]
    with pytest.raises(ProbeError) as e:
        probe.probe_p2p([uuid, uuid], nbytes=1 << 20, flags=probe.P2P_F_KEEP_DEVICE)
    assert e.value.code == _lib.GMP_ERR_INVALID


def test_probe_p2p_pairs(cuda):
    devs = probe.enumerate_cuda()
    if len(devs) < 2:
        pytest.skip("needs at least 2 GPUs")
    uuids = [d.uuid for d in devs]
    r = probe.probe_p2p(uuids, nbytes=256 << 20, iters=2, flags=probe.P2P_F_KEEP_DEVICE)
    n = len(uuids)
    assert r.status == 0 and r.mismatch_lanes == 0 and r.matrix_match == 1
    ca, wr, rd = r.matrix("can_access"), r.matrix("write_gbs"), r.matrix("read_gbs")
    for i in range(n):
        for j in range(n):
            if i == j:
                assert ca[i][j] == 1 and wr[i][j] == 0
            else:
                assert ca[i][j] == 1, "HGX B200: every pair is NVLink-connected"
                assert r.matrix("nvml_nvlink")[i][j] == 0
                assert wr[i][j] > 300 and rd[i][j] > 300, (i, j, wr[i][j], rd[i][j])
                assert r.matrix("memcpy_gbs")[i][j] > 300
                assert r.matrix("atomics_ok")[i][j] == 1 and r.matrix("nvml_atomics")[i][j] == 0
        assert r.bidir_gbs[i] > 500
        assert r.nvlink_active[i] == 18
    assert r.alltoall_gbs > 300 * n


def test_probe_all_and_json_in_a_child_process():
    code = (
        "import json,sys; sys.path.insert(0, %r)\n"
        "from gpumounter_b200 import probe\n"
        "nv = {d.uuid: d for d in probe.enumerate_nvml()}\n"
        "exp = [probe.Device(d.uuid, nv[d.uuid].minor, '', -1, -1) for d in probe.enumerate_cuda()]\n"
        "r = probe.probe_all(exp, hbm_bytes=1<<30, hbm_iters=1, p2p_bytes=128<<20, dma_bytes=16<<20)\n"
        "print(probe.report_json(r))\n" % ROOT
    )
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    j = json.loads(out.stdout.strip().splitlines()[-1])
    assert j["status"] == 0 and j["nodes_status"] == 0 and j["device_set_status"] == 0 and j["nvml_status"] == 0
    assert j["n_gpus"] == len(j["devices"]) == len(j["hbm"]) >= 1
    for d, h in zip(j["devices"], j["hbm"]):
        assert d["minor"] >= 0 and d["nvml_index"] >= 0 and d["cuda_ordinal"] >= 0 and h["uuid"] == d["uuid"]
        assert h["status"] == 0 and h["mismatch_lanes"] == 0 and h["write_gbs"] > 1000 and h["read_gbs"] > 1000
    if j["n_gpus"] >= 2:
        assert j["p2p"]["matrix_match"] == 1 and j["p2p"]["mismatch_lanes"] == 0
    for h in j["health"]:   # clocks and throttle reasons are recorded next to the bandwidths
        assert 1000 <= h["sm_mhz"] <= h["sm_max_mhz"] and h["mem_mhz"] > 1000 and h["ecc_uncorrected_volatile"] <= 0
        assert h["throttle_reasons"] & (0x8 | 0x40 | 0x20) == 0, "hardware / thermal slowdown during the probe"


def test_probe_all_rejects_a_wrong_device_set():
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "from gpumounter_b200 import probe\n"
        "cu = probe.enumerate_cuda()\n"
        "bad = [probe.Device('GPU-00000000-0000-0000-0000-000000000000', None, '', -1, -1)]\n"
        "r = probe.probe_all(bad, hbm_bytes=1<<20, raise_on_error=False, flags=probe.PROBE_F_SKIP_NODES)\n"
        "print(r.status, r.device_set_status)\n"
        "nv = {d.uuid: d for d in probe.enumerate_nvml()}\n"
        "wrong_minor = [probe.Device(d.uuid, (nv[d.uuid].minor + 1) %% 200, '', -1, -1) for d in cu]\n"
        "r = probe.probe_all(wrong_minor, hbm_bytes=1<<20, raise_on_error=False, flags=probe.PROBE_F_SKIP_NODES)\n"
        "print(r.status, r.device_set_status, r.detail.decode())\n" % ROOT
    )
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    l1, l2 = out.stdout.strip().splitlines()[-2:]
    assert l1.split() == [str(_lib.GMP_ERR_DEVICE_SET)] * 2
    assert l2.split()[:2] == [str(_lib.GMP_ERR_DEVICE_SET)] * 2 and "NVML says" in l2


def test_host_view_accepts_a_subset_and_a_resident_process_gets_warm():
    """The worker on the host sees every GPU of the node.  HOST_VIEW: probing a subset is fine, an unknown UUID is still
    refused; with KEEP_DEVICE the second probe of a resident process pays neither cuInit nor a context."""
    code = (
        "import json,sys; sys.path.insert(0, %r)\n"
        "from gpumounter_b200 import probe\n"
        "nv = {d.uuid: d for d in probe.enumerate_nvml()}\n"
        "cu = probe.enumerate_cuda()\n"
        "one = [probe.Device(cu[0].uuid, nv[cu[0].uuid].minor, '', -1, -1)]\n"
        "F = probe.PROBE_F_HOST_VIEW | probe.PROBE_F_SKIP_NODES | probe.PROBE_F_KEEP_DEVICE\n"
        "out = []\n"
        "for i in range(3):\n"
        "    r = probe.probe_all(one, hbm_bytes=256<<20, hbm_iters=1, dma_bytes=1<<20, flags=F, raise_on_error=False)\n"
        "    j = json.loads(probe.report_json(r))\n"
        "    out.append([r.status, j['n_gpus'], j['ms_to_usable'], j['hbm'][0]['ms_context'], j['hbm'][0]['mismatch_lanes']])\n"
        "bad = one + [probe.Device('GPU-00000000-0000-0000-0000-000000000000', None, '', -1, -1)]\n"
        "r = probe.probe_all(bad, hbm_bytes=1<<20, flags=F, raise_on_error=False)\n"
        "strict = probe.probe_all(one, hbm_bytes=1<<20, raise_on_error=False, flags=probe.PROBE_F_SKIP_NODES)\n"
        "print(json.dumps({'runs': out, 'n_cuda': len(cu), 'bad': r.status, 'strict': strict.status}))\n" % ROOT
    )
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    j = json.loads(out.stdout.strip().splitlines()[-1])
    assert all(r[0] == 0 and r[1] == 1 and r[4] == 0 for r in j["runs"]), j
    assert j["runs"][2][2] < 100 and j["runs"][2][3] < 20, j   # warm: ms to usable, ms for the context
    assert j["bad"] == _lib.GMP_ERR_DEVICE_SET
    # without HOST_VIEW a subset is a device-set mismatch (the in-pod rule); with one visible GPU the sets are equal
    assert j["strict"] == (0 if j["n_cuda"] == 1 else _lib.GMP_ERR_DEVICE_SET)


def test_probe_executable_on_the_box():
    exe = probe.probe_executable()
    p = subprocess.run([exe, "--bytes", "1G", "--iters", "1", "--p2p-bytes", "128M"], capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    j = json.loads(p.stdout.strip().splitlines()[-1])
    assert j["status"] == 0 and j["verdict"] == "ok" and j["n_gpus"] >= 1
    assert j["ms_total"] < 20000


def test_p2p_matrix_mismatch_against_nvml_is_reported(cuda, tmp_path, monkeypatch):
    """Fault injection: a fake NVML that lists the REAL devices but claims READ is not supported from GPU 0 to GPU 1,
    while CUDA says the pair is accessible -> GMP_ERR_P2P_MATRIX (bandwidth still measured and verified)."""
    devs = probe.enumerate_cuda()
    if len(devs) < 2:
        pytest.skip("needs at least 2 GPUs")
    real = {d.uuid: d for d in probe.enumerate_nvml()}
    fake = os.path.join(ROOT, "tests", "fake_nvml", "libfake_nvml.so")
    if not os.path.exists(fake):
        subprocess.run(["make", "-s", "-C", os.path.dirname(fake)], check=True)
    cfg = tmp_path / "nvml.cfg"
    lines = [f"gpu {d.uuid} {real[d.uuid].minor} {d.pci_bus_id}" for d in devs[:2]]
    lines += ["p2p 0 1 0 5", f"nvlink {devs[0].uuid} 18", f"nvlink {devs[1].uuid} 18"]   # READ 0->1: NOT_SUPPORTED
    cfg.write_text("\n".join(lines) + "\n")
    monkeypatch.setenv("GPUMOUNTER_NVML_LIB", fake)
    monkeypatch.setenv("GMP_FAKE_NVML_CONFIG", str(cfg))
    r = probe.probe_p2p([d.uuid for d in devs[:2]], nbytes=64 << 20, iters=1, flags=probe.P2P_F_KEEP_DEVICE,
                        raise_on_error=False)
    assert r.status == _lib.GMP_ERR_P2P_MATRIX and r.matrix_match == 0 and r.nvml_available == 1
    assert r.matrix("nvml_read") == [[0, 5], [0, 0]] and r.matrix("can_access") == [[1, 1], [1, 1]]
    assert r.mismatch_lanes == 0 and r.matrix("write_gbs")[0][1] > 100


def test_probe_executable_reads_environment_defaults():
    exe = probe.probe_executable()
    env = dict(os.environ, GPUMOUNTER_PROBE_BYTES="64M", GPUMOUNTER_PROBE_ITERS="1", GPUMOUNTER_PROBE_DMA_BYTES="0",
               GPUMOUNTER_PROBE_SKIP_P2P="1", GPUMOUNTER_PROBE_PATTERNS="all")
    p = subprocess.run([exe], capture_output=True, text=True, timeout=300, env=env)
    assert p.returncode == 0, p.stdout[-1000:] + p.stderr[-1000:]
    j = json.loads(p.stdout.strip().splitlines()[-1])
    assert j["status"] == 0 and all(h["bytes"] == 64 << 20 and h["h2d_gbs"] == 0 for h in j["hbm"]) and "p2p" not in j


def test_impossible_allocation_is_a_clean_nomem():
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "from gpumounter_b200 import probe\n"
        "r = probe.probe_hbm(nbytes=400 << 30, iters=1, dma_bytes=0, raise_on_error=False)\n"
        "print(r.status)\n"
        "r = probe.probe_hbm(nbytes=64 << 20, iters=1, dma_bytes=0)\n"
        "print(r.status)\n" % ROOT
    )
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    assert out.stdout.split() == [str(_lib.GMP_ERR_NOMEM), "0"]     # and the device is still usable afterwards


def test_repeated_one_shot_probes_leave_nothing_behind():
    """Config 5's leak-free contract on the probe side: after each call (default flags: cudaDeviceReset) the device's free
    memory is back where it started, and after KEEP_DEVICE calls only the context itself remains."""
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "from cuda.bindings import runtime as rt\n"
        "from gpumounter_b200 import probe\n"
        "def free():\n"
        "    err, f, t = rt.cudaMemGetInfo(); assert int(err) == 0; return f\n"
        "probe.probe_hbm(nbytes=256 << 20, iters=1, dma_bytes=1 << 20, flags=probe.HBM_F_KEEP_DEVICE)\n"
        "base = free()\n"
        "for i in range(12):\n"
        "    r = probe.probe_hbm(nbytes=(512 << 20) + 16 * i, iters=1, dma_bytes=4 << 20, flags=probe.HBM_F_KEEP_DEVICE)\n"
        "    assert r.status == 0\n"
        "kept = free()\n"
        "for i in range(4):\n"
        "    assert probe.probe_hbm(nbytes=512 << 20, iters=1, dma_bytes=4 << 20).status == 0   # resets the device\n"
        "uuid = probe.enumerate_cuda()[0].uuid\n"
        "assert probe.probe_p2p([uuid], nbytes=64 << 20).status == 0\n"
        "print(base, kept)\n" % ROOT
    )
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    base, kept = (int(x) for x in out.stdout.split()[-2:])
    assert abs(base - kept) <= 64 << 20, (base, kept)   # nothing accumulates over a dozen calls (allocator granularity aside)


def test_probe_hbm_retained_buffers_are_reused_and_released(cuda):
    """GMP_HBM_F_RETAIN (a resident worker probing again and again): the second call allocates nothing, answers the same
    bit-exact checksums, and the memory comes back on gmp_probe_hbm_release() - or on the next call without the flag."""
    torch = cuda
    uuid = probe.enumerate_cuda()[0].uuid
    nbytes = (512 << 20) + 16
    flags = probe.HBM_F_KEEP_DEVICE | probe.HBM_F_RETAIN
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    r1 = probe.probe_hbm(uuid, nbytes=nbytes, iters=1, seed=41, dma_bytes=4 << 20, flags=flags)
    free1, _ = torch.cuda.mem_get_info()
    assert r1.status == 0 and free0 - free1 >= 2 * nbytes                       # both buffers are still there
    r2 = probe.probe_hbm(uuid, nbytes=nbytes, iters=1, seed=41, dma_bytes=4 << 20, flags=flags)
    free2, _ = torch.cuda.mem_get_info()
    assert r2.status == 0 and free2 == free1                                    # nothing allocated, nothing freed
    assert (r2.xor64, r2.sum64) == (r1.xor64, r1.sum64) == gmo.pattern_checksum(nbytes, probe.PAT_SPLITMIX, 41)
    assert r2.ns_alloc < 1_000_000 and r2.ns_teardown < 1_000_000 and r2.ns_alloc < r1.ns_alloc
    assert r2.h2d_bytes == 4 << 20 and r2.mismatch_lanes == 0
    # a smaller request fits in what is retained; a larger one grows it
    r3 = probe.probe_hbm(uuid, nbytes=nbytes // 2, iters=1, seed=42, dma_bytes=0, flags=flags)
    assert r3.status == 0 and torch.cuda.mem_get_info()[0] == free1
    assert (r3.xor64, r3.sum64) == gmo.pattern_checksum(nbytes // 2, probe.PAT_SPLITMIX, 42)
    probe.probe_hbm_release(uuid)
    assert torch.cuda.mem_get_info()[0] >= free0 - (8 << 20)                    # given back (modulo allocator granularity)
    # a call WITHOUT the flag gives back what earlier RETAIN calls left
    probe.probe_hbm(uuid, nbytes=nbytes, iters=1, dma_bytes=0, flags=flags)
    assert free0 - torch.cuda.mem_get_info()[0] >= 2 * nbytes
    r4 = probe.probe_hbm(uuid, nbytes=64 << 20, iters=1, dma_bytes=0, flags=probe.HBM_F_KEEP_DEVICE)
    assert r4.status == 0 and torch.cuda.mem_get_info()[0] >= free0 - (8 << 20)
    probe.probe_hbm_release()                                                   # nothing left: still GMP_OK
