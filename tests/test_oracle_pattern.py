"""The pattern oracle against published SplitMix64 vectors and its own invariants (CPU only)."""

import random

import pytest

from oracle import gmo_ctypes as gmo

# Vigna's splitmix64.c with x = 1234567, first five outputs (widely published test vector).
SPLITMIX64_SEED_1234567 = [
    6457827717110365317,
    3203168211198807973,
    9817491932198370423,
    4593380528125082431,
    16408922859458223821,
]
M64 = (1 << 64) - 1


def test_splitmix64_known_answers():
    assert gmo.splitmix64_stream(1234567, 5) == SPLITMIX64_SEED_1234567


def test_lane_word0_is_the_splitmix_stream():
    # header definition: w0 of lane L is the L-th output of the stream seeded with `seed`
    for L, expect in enumerate(SPLITMIX64_SEED_1234567):
        w0, w1 = gmo.lane_words(gmo_pat("splitmix"), 1234567, L)
        assert w0 == expect
        swapped = ((w0 << 32) | (w0 >> 32)) & M64
        assert w1 == (~swapped & M64) ^ L


def gmo_pat(name):
    return {"splitmix": 0, "zero": 1, "ones": 2, "checker": 3, "address": 4}[name]


def test_fixed_patterns():
    assert gmo.lane_words(1, 99, 7) == (0, 0)
    assert gmo.lane_words(2, 99, 7) == (M64, M64)
    assert gmo.lane_words(3, 0, 0) == (0xAAAAAAAAAAAAAAAA,) * 2
    assert gmo.lane_words(3, 0, 1) == (0x5555555555555555,) * 2
    assert gmo.lane_words(3, 1, 0) == (0x5555555555555555,) * 2
    assert gmo.lane_words(4, 0x7F0000000000, 3) == (0x7F0000000000 + 48, 0x7F0000000000 + 56)
    with pytest.raises(ValueError):
        gmo.lane_words(5, 0, 0)


@pytest.mark.parametrize("pattern", range(5))
@pytest.mark.parametrize("nbytes", [0, 1, 7, 8, 9, 15, 16, 17, 31, 32, 4096, 4099, 65536 + 13])
def test_fill_verify_round_trip_and_checksum(pattern, nbytes):
    seed = 0x9E3779B97F4A7C15 + pattern
    buf = gmo.pattern_fill(nbytes, pattern, seed)
    bad, first, x, s = gmo.pattern_verify(buf, pattern, seed)
    assert (bad, first) == (0, gmo.UINT64_MAX)
    assert (x, s) == gmo.pattern_checksum(nbytes, pattern, seed)
    # checksum definition: little-endian u64 words, last one zero padded
    padded = bytes(buf) + b"\0" * (-nbytes % 8)
    words = [int.from_bytes(padded[i:i + 8], "little") for i in range(0, len(padded), 8)]
    xx = 0
    for w in words:
        xx ^= w
    assert x == xx and s == sum(words) & M64


def test_fill_matches_lane_words_bytewise():
    seed, n = 42, 16 * 10 + 11
    buf = gmo.pattern_fill(n, 0, seed)
    for L in range(11):
        w0, w1 = gmo.lane_words(0, seed, L)
        lane = (w0.to_bytes(8, "little") + w1.to_bytes(8, "little"))
        assert bytes(buf[16 * L:16 * L + 16]) == lane[:min(16, n - 16 * L)]


def test_verify_counts_lanes_and_finds_first():
    rng = random.Random(7)
    n = 16 * 1000 + 5
    buf = gmo.pattern_fill(n, 0, 5)
    offsets = sorted(rng.sample(range(n), 20))
    for off in offsets:
        buf[off] ^= 0x40
    lanes = sorted({o // 16 for o in offsets})
    bad, first, _, _ = gmo.pattern_verify(buf, 0, 5)
    assert bad == len(lanes) and first == lanes[0]
    # a flip in the ragged tail lane is found too
    buf2 = gmo.pattern_fill(n, 0, 5)
    buf2[n - 1] ^= 1
    assert gmo.pattern_verify(buf2, 0, 5)[:2] == (1, n // 16)


def test_wrong_seed_mismatches_everything():
    n = 16 * 64
    buf = gmo.pattern_fill(n, 0, 1)
    bad, first, _, _ = gmo.pattern_verify(buf, 0, 2)
    assert bad == 64 and first == 0


def test_cpu_probe_threads_agree():
    n = (1 << 20) + 3
    for threads in (1, 3, 8):
        secs, (bad, first, x, s) = gmo.cpu_probe(n, threads, 77)
        assert bad == 0 and first == gmo.UINT64_MAX
        assert (x, s) == gmo.pattern_checksum(n, 0, 77)
        assert all(t > 0 for t in secs)
