/*
 * oracle/gmo_pattern.h — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU statement (plain C) of the probe's pattern generator, verifier and
 * checksums, used only as the checker by tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs.  Nothing in gpumounter_b200/
 * may call it.
 *
 * The reference (jason-gideon/GPUMounter) has no device-memory code at all
 * (SURVEY.md 2.3), so these functions have no reference file:line to follow; they
 * restate the definition in include/gpumounter_probe.h section 3.  What pins them:
 * lane word w0 of GMO_PAT_SPLITMIX is, by definition, the L-th output of the
 * canonical SplitMix64 generator (Steele/Lea/Flood 2014; Vigna's splitmix64.c),
 * whose published outputs for seed 1234567 are checked in
 * tests/test_oracle_pattern.py.  Beyond that vector the pattern parity is
 * self-defined ("parity unpinned" does not apply to a definition we own).
 */
#ifndef GMO_PATTERN_H_
#define GMO_PATTERN_H_
#include <stddef.h>
#include <stdint.h>

enum { GMO_PAT_SPLITMIX = 0, GMO_PAT_ZERO = 1, GMO_PAT_ONES = 2, GMO_PAT_CHECKER = 3, GMO_PAT_ADDRESS = 4, GMO_PAT_COUNT = 5 };

typedef struct gmo_verify_result {
    uint64_t mismatch_lanes;
    uint64_t first_bad_lane; /* UINT64_MAX when none */
    uint64_t xor64;
    uint64_t sum64;
} gmo_verify_result;

/* Canonical SplitMix64: advances *state by the golden gamma and returns the mixed value. */
uint64_t gmo_splitmix64_next(uint64_t* state);

/* The two little-endian u64 words of 16-byte lane L. Returns 0, or -1 for an unknown pattern. */
int gmo_lane_words(int pattern, uint64_t seed, uint64_t lane, uint64_t* w0, uint64_t* w1);

/* Fill `bytes` bytes (any size, ragged last lane allowed) with the pattern. */
int gmo_pattern_fill(void* buf, size_t bytes, int pattern, uint64_t seed);

/* Compare `bytes` bytes against the pattern and checksum what was read. */
int gmo_pattern_verify(const void* buf, size_t bytes, int pattern, uint64_t seed, gmo_verify_result* out);

/* Checksums of the pattern itself, computed without materialising a buffer
 * (lets tests check multi-GiB device buffers against the oracle). */
int gmo_pattern_checksum(size_t bytes, int pattern, uint64_t seed, uint64_t* xor64, uint64_t* sum64);

/* Same three passes on `threads` host threads over host DRAM, timed:
 * fill(a) -> verify(a) -> memcpy-style copy(a->b).  The CPU baseline of bench.py.
 * seconds[0..2] receive wall seconds of each pass; returns 0 or -1. */
int gmo_cpu_probe(size_t bytes, int threads, uint64_t seed, double seconds[3], gmo_verify_result* out);

#endif
