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
                        flags=probe.HBM_F_KEEP_DEVICE)
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
    assert r["h2d_bytes"] == 64 << 20
