"""The allocation-free verify entry point against the allocating one, through the C ABI (GPU box only).

gmp_hbm_read_verify_noalloc exists for callers that cannot allocate device memory: a health-check sidecar on a GPU
whose memory the tenant has filled, a monitor that must not cause cudaFree's device-wide synchronisation, a caller
whose own rules forbid allocator calls at that point.  Its scratch is one fixed array inside the library that every call
on a device shares, so this file checks the thing that sharing could break: two calls in a row, on different data,
each have to agree bit for bit with gmp_hbm_read_verify on the same buffer.

The file name sorts last on purpose: if this test fails, `pytest -x` has already run everything else.  It makes no
assumption about what ran before it and passes alone in a fresh process as well.
"""

import pytest

from gpumounter_b200 import ProbeError, _lib, probe
from oracle import gmo_ctypes as gmo  # the checker

pytestmark = pytest.mark.gpu

SEED = 0x9E3779B97F4A7C15


def as_tuple(r):
    return (r.mismatch_lanes, r.first_bad_lane, r.xor64, r.sum64)


def test_two_calls_in_a_row_agree_with_the_allocating_entry_point(cuda):
    torch = cuda
    s = torch.cuda.current_stream().cuda_stream
    # first call: a clean buffer spanning several 4 MiB segments, ragged end
    n1 = (19 << 20) + 7
    a = torch.empty(n1 + 64, dtype=torch.uint8, device="cuda")
    probe.hbm_write_pattern(a.data_ptr(), n1, probe.PAT_SPLITMIX, SEED, s)
    want1 = probe.hbm_read_verify(a.data_ptr(), n1, probe.PAT_SPLITMIX, SEED, s)
    got1 = probe.hbm_read_verify_noalloc(a.data_ptr(), n1, probe.PAT_SPLITMIX, SEED, s)
    assert as_tuple(got1) == as_tuple(want1) and got1.ok
    # second call straight after, on other data with faults in three different segments and in the ragged tail:
    # the array still holds the first call's slots and result, none of which may leak into this answer
    n2 = (9 << 20) + 3
    b = torch.empty(n2 + 64, dtype=torch.uint8, device="cuda")
    probe.hbm_write_pattern(b.data_ptr(), n2, probe.PAT_ADDRESS, b.data_ptr(), s)
    for off in (5, (4 << 20) + 1234, (8 << 20) + 99, n2 - 1):
        probe.hbm_corrupt(b.data_ptr(), n2, off, 0x04, s)
    want2 = probe.hbm_read_verify(b.data_ptr(), n2, probe.PAT_ADDRESS, b.data_ptr(), s)
    got2 = probe.hbm_read_verify_noalloc(b.data_ptr(), n2, probe.PAT_ADDRESS, b.data_ptr(), s)
    assert as_tuple(got2) == as_tuple(want2)
    assert (got2.mismatch_lanes, got2.first_bad_lane) == (4, 0)
    # and the first buffer again: still the first answer
    assert as_tuple(probe.hbm_read_verify_noalloc(a.data_ptr(), n1, probe.PAT_SPLITMIX, SEED, s)) == as_tuple(want1)


@pytest.mark.parametrize("nbytes", [0, 1, 15, 16, 17, 4 << 20, (4 << 20) + 5, (4 << 20) - 16, (64 << 20) + 48])
@pytest.mark.parametrize("pattern", range(probe.PAT_COUNT))
def test_sizes_and_patterns_agree(cuda, nbytes, pattern):
    torch = cuda
    s = torch.cuda.current_stream().cuda_stream
    t = torch.empty(nbytes + 64, dtype=torch.uint8, device="cuda")
    seed = SEED + pattern
    probe.hbm_write_pattern(t.data_ptr(), nbytes, pattern, seed, s)
    assert as_tuple(probe.hbm_read_verify_noalloc(t.data_ptr(), nbytes, pattern, seed, s)) == as_tuple(
        probe.hbm_read_verify(t.data_ptr(), nbytes, pattern, seed, s))
    # a wrong seed makes (nearly) every lane differ; both entry points must count the same lanes
    if pattern in (probe.PAT_SPLITMIX, probe.PAT_ADDRESS) and nbytes:
        assert as_tuple(probe.hbm_read_verify_noalloc(t.data_ptr(), nbytes, pattern, seed + 1, s)) == as_tuple(
            probe.hbm_read_verify(t.data_ptr(), nbytes, pattern, seed + 1, s))


def test_size_beyond_the_scratch_is_refused_with_a_clear_message(cuda):
    torch = cuda
    t = torch.empty(1 << 20, dtype=torch.uint8, device="cuda")
    with pytest.raises(ProbeError) as e:
        # only the size is looked at before the refusal; nothing is launched on the too-small tensor
        probe.hbm_read_verify_noalloc(t.data_ptr(), probe.NOALLOC_MAX_BYTES + 16)
    assert e.value.code == _lib.GMP_ERR_INVALID
    assert "exceed" in e.value.detail and "2048 segments" in e.value.detail


def test_the_largest_allowed_size(cuda):
    torch = cuda
    s = torch.cuda.current_stream().cuda_stream
    n = probe.NOALLOC_MAX_BYTES
    t = torch.empty(n, dtype=torch.uint8, device="cuda")
    probe.hbm_write_pattern(t.data_ptr(), n, probe.PAT_SPLITMIX, SEED + 9, s)
    probe.hbm_corrupt(t.data_ptr(), n, n - 7, 0x80, s)
    assert as_tuple(probe.hbm_read_verify_noalloc(t.data_ptr(), n, probe.PAT_SPLITMIX, SEED + 9, s)) == as_tuple(
        probe.hbm_read_verify(t.data_ptr(), n, probe.PAT_SPLITMIX, SEED + 9, s))


def test_survives_a_device_reset_done_by_the_library():
    """gmp_probe_hbm (default flags) ends in cudaDeviceReset; the scratch address looked up before it is then stale and
    the library has to look it up again.  Run in a child so this process's torch context is not torn down."""
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "from gpumounter_b200 import probe\n"
        "from cuda.bindings import runtime as rt   # plain device memory without torch: a reset would orphan torch's pool\n"
        "def check(nbytes, seed):\n"
        "    err, p = rt.cudaMalloc(nbytes); assert int(err) == 0, err\n"
        "    p = int(p)\n"
        "    probe.hbm_write_pattern(p, nbytes, 0, seed)\n"
        "    a = probe.hbm_read_verify(p, nbytes, 0, seed); b = probe.hbm_read_verify_noalloc(p, nbytes, 0, seed)\n"
        "    assert (a.mismatch_lanes, a.xor64, a.sum64) == (b.mismatch_lanes, b.xor64, b.sum64) == (0, a.xor64, a.sum64)\n"
        "    assert int(rt.cudaFree(p)[0]) == 0\n"
        "check((9 << 20) + 5, 1)\n"
        "r = probe.probe_hbm(nbytes=64 << 20, iters=1, dma_bytes=0)   # resets the device at the end\n"
        "assert r.status == 0\n"
        "check((9 << 20) + 5, 2)\n"
        "print('ok')\n" % root
    )
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and out.stdout.strip().endswith("ok"), out.stdout[-1000:] + out.stderr[-2000:]


@pytest.mark.parametrize("pattern", range(probe.PAT_COUNT))
def test_agrees_with_the_cpu_oracle(cuda, pattern):
    """Pinned to oracle/gmo_pattern.c, not only to the sibling CUDA entry point: closed-form checksum of a clean buffer,
    and the oracle's own verify pass over the bytes of a buffer with faults in three segments and in the ragged tail."""
    torch = cuda
    s = torch.cuda.current_stream().cuda_stream
    n = (9 << 20) + 11
    t = torch.empty(n + 64, dtype=torch.uint8, device="cuda")
    seed = t.data_ptr() if pattern == probe.PAT_ADDRESS else SEED + 3 * pattern
    probe.hbm_write_pattern(t.data_ptr(), n, pattern, seed, s)
    clean = probe.hbm_read_verify_noalloc(t.data_ptr(), n, pattern, seed, s)
    assert clean.ok and (clean.xor64, clean.sum64) == gmo.pattern_checksum(n, pattern, seed)
    for off in (17, (4 << 20) + 4321, (8 << 20) + 16, n - 2):
        probe.hbm_corrupt(t.data_ptr(), n, off, 0x21, s)
    got = probe.hbm_read_verify_noalloc(t.data_ptr(), n, pattern, seed, s)
    want = gmo.pattern_verify(bytearray(t[:n].cpu().numpy().tobytes()), pattern, seed)
    assert as_tuple(got) == tuple(want)
    assert got.mismatch_lanes == 4 and got.first_bad_lane == 1


def test_after_one_call_no_pattern_allocates(cuda):
    """The point of the entry point: after ONE call on the device (whatever its pattern) further calls - every pattern,
    several sizes - leave the device's free memory exactly where it was: no lazily loaded kernel, no hidden scratch."""
    torch = cuda
    s = torch.cuda.current_stream().cuda_stream
    n = (12 << 20) + 5
    t = torch.empty(n + 64, dtype=torch.uint8, device="cuda")
    probe.hbm_write_pattern(t.data_ptr(), n, probe.PAT_ZERO, 0, s)
    assert probe.hbm_read_verify_noalloc(t.data_ptr(), 0, probe.PAT_ZERO, 0, s).ok  # the first call; it launches nothing
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for pattern in range(probe.PAT_COUNT):
        for size in (n, 1 << 20, 16):
            probe.hbm_read_verify_noalloc(t.data_ptr(), size, pattern, SEED, s)
            free, _ = torch.cuda.mem_get_info()
            assert free == free0, (pattern, size, free0 - free)


def test_survives_a_device_reset_done_by_the_caller():
    """A cudaDeviceReset() of the CALLER's, which the library cannot see: the next call sets itself up again (the array's
    address is looked up per call, nothing stale is kept).  In a child: the reset would orphan this process's torch pool."""
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "from gpumounter_b200 import probe\n"
        "from cuda.bindings import runtime as rt\n"
        "def check(nbytes, seed):\n"
        "    err, p = rt.cudaMalloc(nbytes); assert int(err) == 0, err\n"
        "    p = int(p)\n"
        "    probe.hbm_write_pattern(p, nbytes, 0, seed)\n"
        "    a = probe.hbm_read_verify(p, nbytes, 0, seed); b = probe.hbm_read_verify_noalloc(p, nbytes, 0, seed)\n"
        "    assert (a.mismatch_lanes, a.xor64, a.sum64) == (b.mismatch_lanes, b.xor64, b.sum64) == (0, a.xor64, a.sum64)\n"
        "    assert int(rt.cudaFree(p)[0]) == 0\n"
        "check((9 << 20) + 5, 1)\n"
        "assert int(rt.cudaDeviceReset()[0]) == 0          # the caller's own reset\n"
        "check((9 << 20) + 5, 2)\n"
        "check((5 << 20) + 9, 3)\n"
        "print('ok')\n" % root
    )
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and out.stdout.strip().endswith("ok"), out.stdout[-1000:] + out.stderr[-2000:]
