"""Parity of the sm_100a HBM kernels against the CPU oracle, through the C ABI (GPU box only).

Integer/byte work: every comparison is bit-exact.
"""

import numpy as np
import pytest

from gpumounter_b200 import ProbeError, _lib, probe
from oracle import gmo_ctypes as gmo

pytestmark = pytest.mark.gpu

SEED = 0x9E3779B97F4A7C15
SIZES = [0, 1, 7, 8, 15, 16, 17, 255, 4096, 4096 + 9, 16 * 1024 * 4 - 16, 16 * 1024 * 4, 16 * 1024 * 4 + 16,
         (1 << 20) + 3, (64 << 20) + 5]


def dev_buffer(torch, nbytes, fill=0xEE):
    # over-allocate so that a write past the end would be visible
    t = torch.full((nbytes + 64,), fill, dtype=torch.uint8, device="cuda")
    assert t.data_ptr() % 16 == 0
    return t


def stream_of(torch):
    return torch.cuda.current_stream().cuda_stream


@pytest.mark.parametrize("pattern", range(probe.PAT_COUNT))
@pytest.mark.parametrize("nbytes", SIZES)
def test_write_pattern_bytes_equal_oracle(cuda, pattern, nbytes):
    torch = cuda
    t = dev_buffer(torch, nbytes)
    seed = SEED + pattern if pattern != probe.PAT_ADDRESS else t.data_ptr()
    probe.hbm_write_pattern(t.data_ptr(), nbytes, pattern, seed, stream_of(torch))
    torch.cuda.synchronize()
    got = t.cpu().numpy()
    want = np.frombuffer(bytes(gmo.pattern_fill(nbytes, pattern, seed)), dtype=np.uint8)
    assert np.array_equal(got[:nbytes], want)
    assert (got[nbytes:] == 0xEE).all(), "kernel wrote past the end of the buffer"


@pytest.mark.parametrize("pattern", range(probe.PAT_COUNT))
@pytest.mark.parametrize("nbytes", SIZES)
def test_read_verify_clean_buffer_matches_oracle_checksums(cuda, pattern, nbytes):
    torch = cuda
    t = dev_buffer(torch, nbytes)
    seed = SEED ^ (pattern * 0x1234567)
    probe.hbm_write_pattern(t.data_ptr(), nbytes, pattern, seed, stream_of(torch))
    r = probe.hbm_read_verify(t.data_ptr(), nbytes, pattern, seed, stream_of(torch))
    assert (r.mismatch_lanes, r.first_bad_lane) == (0, probe.UINT64_MAX)
    assert (r.xor64, r.sum64) == gmo.pattern_checksum(nbytes, pattern, seed)


@pytest.mark.parametrize("nbytes", [16, 4096 + 9, (1 << 20) + 3, (64 << 20) + 5])
def test_read_verify_finds_injected_faults_like_the_oracle(cuda, nbytes):
    torch = cuda
    rng = np.random.default_rng(nbytes)
    t = dev_buffer(torch, nbytes)
    s = stream_of(torch)
    probe.hbm_write_pattern(t.data_ptr(), nbytes, probe.PAT_SPLITMIX, SEED, s)
    offsets = sorted(set(int(o) for o in rng.integers(0, nbytes, size=min(37, nbytes))))
    offsets.append(nbytes - 1)  # last byte: ragged lane when nbytes % 16 != 0
    offsets = sorted(set(offsets))
    for off in offsets:
        probe.hbm_corrupt(t.data_ptr(), nbytes, off, 0x10, s)
    r = probe.hbm_read_verify(t.data_ptr(), nbytes, probe.PAT_SPLITMIX, SEED, s)
    host = bytearray(t.cpu().numpy()[:nbytes].tobytes())
    bad, first, x, ssum = gmo.pattern_verify(host, probe.PAT_SPLITMIX, SEED)
    assert bad == len({o // 16 for o in offsets})
    assert (r.mismatch_lanes, r.first_bad_lane, r.xor64, r.sum64) == (bad, first, x, ssum)


def test_read_verify_with_wrong_seed_flags_every_lane(cuda):
    torch = cuda
    nbytes = (1 << 20) + 3
    t = dev_buffer(torch, nbytes)
    probe.hbm_write_pattern(t.data_ptr(), nbytes, probe.PAT_SPLITMIX, 1, stream_of(torch))
    r = probe.hbm_read_verify(t.data_ptr(), nbytes, probe.PAT_SPLITMIX, 2, stream_of(torch))
    assert r.mismatch_lanes == (nbytes + 15) // 16 and r.first_bad_lane == 0


@pytest.mark.parametrize("nbytes", SIZES)
def test_copy_is_byte_exact(cuda, nbytes):
    torch = cuda
    src = torch.randint(0, 256, (nbytes + 64,), dtype=torch.uint8, device="cuda")
    dst = dev_buffer(torch, nbytes, fill=0x33)
    probe.hbm_copy(dst.data_ptr(), src.data_ptr(), nbytes, stream_of(torch))
    torch.cuda.synchronize()
    assert torch.equal(dst[:nbytes], src[:nbytes])
    assert bool((dst[nbytes:] == 0x33).all())


def test_async_verify_and_fetch(cuda):
    torch = cuda
    nbytes = (8 << 20) + 16
    t = dev_buffer(torch, nbytes)
    acc = torch.empty(4, dtype=torch.int64, device="cuda")
    s = stream_of(torch)
    probe.hbm_write_pattern(t.data_ptr(), nbytes, probe.PAT_SPLITMIX, 9, s)
    probe.hbm_read_verify_async(t.data_ptr(), nbytes, probe.PAT_SPLITMIX, 9, acc.data_ptr(), s)
    r = probe.hbm_verify_fetch(acc.data_ptr(), s)
    assert r.ok and (r.xor64, r.sum64) == gmo.pattern_checksum(nbytes, probe.PAT_SPLITMIX, 9)


@pytest.mark.parametrize("knob", ["write", "read", "copy"])
def test_every_launch_variant_is_bit_exact(cuda, knob):
    torch = cuda
    nbytes = (32 << 20) + 7
    a = dev_buffer(torch, nbytes)
    b = dev_buffer(torch, nbytes, fill=0x11)
    s = stream_of(torch)
    want = gmo.pattern_checksum(nbytes, probe.PAT_SPLITMIX, 123)
    keep_v, keep_c = probe.tune(f"{knob}_variant"), probe.tune(f"{knob}_ctas_per_sm")
    try:
        for variant in range(24 if knob == "copy" else 16):
            for ctas in (0, 1, 5):
                probe.tune(f"{knob}_variant", variant)
                probe.tune(f"{knob}_ctas_per_sm", ctas)
                a.fill_(0xEE)
                b.fill_(0x11)
                probe.hbm_write_pattern(a.data_ptr(), nbytes, probe.PAT_SPLITMIX, 123, s)
                probe.hbm_copy(b.data_ptr(), a.data_ptr(), nbytes, s)
                r = probe.hbm_read_verify(b.data_ptr(), nbytes, probe.PAT_SPLITMIX, 123, s)
                assert r.ok and (r.xor64, r.sum64) == want, (knob, variant, ctas)
                assert bool((a[nbytes:] == 0xEE).all()) and bool((b[nbytes:] == 0x11).all())
    finally:
        probe.tune(f"{knob}_variant", keep_v)
        probe.tune(f"{knob}_ctas_per_sm", keep_c)


def test_full_size_buffer_properties(cuda):
    """BASELINE size (8 GiB per buffer): checksum equals the oracle's closed-form pattern checksum,
    copy preserves it, and a single flipped bit anywhere is found at the right lane."""
    torch = cuda
    nbytes = 8 << 30
    a = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    b = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    s = stream_of(torch)
    seed = SEED + 5
    probe.hbm_write_pattern(a.data_ptr(), nbytes, probe.PAT_SPLITMIX, seed, s)
    r = probe.hbm_read_verify(a.data_ptr(), nbytes, probe.PAT_SPLITMIX, seed, s)
    want = gmo.pattern_checksum(nbytes, probe.PAT_SPLITMIX, seed)
    assert r.ok and (r.xor64, r.sum64) == want
    probe.hbm_copy(b.data_ptr(), a.data_ptr(), nbytes, s)
    r = probe.hbm_read_verify(b.data_ptr(), nbytes, probe.PAT_SPLITMIX, seed, s)
    assert r.ok and (r.xor64, r.sum64) == want
    for off in (0, nbytes // 2 + 5, nbytes - 1):
        probe.hbm_corrupt(b.data_ptr(), nbytes, off, 0x01, s)
        r = probe.hbm_read_verify(b.data_ptr(), nbytes, probe.PAT_SPLITMIX, seed, s)
        assert (r.mismatch_lanes, r.first_bad_lane) == (1, off // 16)
        assert r.xor64 == want[0] ^ (1 << (8 * (off % 8)))
        probe.hbm_corrupt(b.data_ptr(), nbytes, off, 0x01, s)  # undo
    # sampled lanes of the big buffer equal the oracle's lane words
    rng = np.random.default_rng(42)
    for L in [0, 1, (nbytes // 16) - 1] + [int(x) for x in rng.integers(0, nbytes // 16, size=64)]:
        got = a[16 * L:16 * L + 16].cpu().numpy().tobytes()
        w0, w1 = gmo.lane_words(probe.PAT_SPLITMIX, seed, L)
        assert got == w0.to_bytes(8, "little") + w1.to_bytes(8, "little")


def test_bad_arguments_are_rejected_on_the_gpu_box(cuda):
    torch = cuda
    t = dev_buffer(torch, 64)
    with pytest.raises(ProbeError) as e:
        probe.hbm_write_pattern(t.data_ptr() + 8, 16)
    assert e.value.code == _lib.GMP_ERR_INVALID
    with pytest.raises(ProbeError) as e:
        probe.hbm_read_verify(t.data_ptr(), 16, pattern=-1)
    assert e.value.code == _lib.GMP_ERR_INVALID
    with pytest.raises(ProbeError) as e:
        probe.hbm_corrupt(t.data_ptr(), 16, 16, 1)
    assert e.value.code == _lib.GMP_ERR_INVALID
