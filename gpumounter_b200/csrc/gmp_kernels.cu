// HBM probe kernels for NVIDIA B200 (sm_100a): pattern write, read-verify, copy.
//
// All three are pure streaming kernels bound by HBM3e bandwidth (no reuse, a few integer ops per 16 bytes), so the
// design rules are the memory ones:
//   * every register-path access is one vector instruction per thread: 16 bytes (LDG/STG.E.128) or, Blackwell only,
//     32 bytes (LDG/STG.E.ENL2.256); a warp instruction covers 512 B or 1 KiB of contiguous memory;
//   * each thread issues all UNROLL loads of a tile before it touches any of them (memory-level parallelism hides the
//     ~600-cycle DRAM latency); __launch_bounds__(T, 1024/T) keeps ptxas from chasing 32 registers and serialising them;
//   * CTAs walk contiguous tiles in grid-stride order, so the whole chip sweeps one narrow window of memory at a time;
//   * the grid is sized per kernel from a sweep on silicon (profiles/r01_sweep.md): many small waves for the write
//     kernel (the hardware scheduler evens out the ~10 % speed spread between SMs), exactly one resident wave for the
//     verify kernel (its accumulators live in registers; four atomics per CTA at the end);
//   * streaming cache hints (.cs) keep the 126 MB L2 from being polluted by data that is never re-read;
//   * the copy does not use the register path at all: cp.async.bulk moves 112 KiB chunks global -> shared -> global
//     through the TMA engine, one issuing thread per CTA (k_hbm_copy_bulk); the vector kernel k_hbm_copy remains for
//     the last partial chunk, for misaligned pointers and for peer copies;
//   * the verify fast path only ORs difference bits into one word per tile and re-walks a dirty tile in a
//     __noinline__ slow path: 33 instead of 47 instructions per lane, which moved the kernel off the ALU roof.
//
// Algorithmic bytes per launch (used for every GB/s figure): write = N, verify = N, copy = 2N for an N-byte buffer.
//
// There is no reference counterpart (the reference has no device code); the CPU statement of the same arithmetic used
// by the tests is oracle/gmo_pattern.c.

#include "gmp_internal.h"

#include <atomic>
#include <stdlib.h>
#include <string_view>

namespace gmp {

// ---- cache-hinted vector accesses ------------------------------------------------
// A "chunk" is VEC adjacent 16-byte lanes moved by ONE instruction: VEC=1 is the
// classic 128-bit LDG/STG, VEC=2 is Blackwell's 256-bit LDG.E.256 / STG.E.256
// (sm_100+, needs 32-byte alignment), which halves the LSU instruction count.
enum { HINT_DEFAULT = 0, HINT_STREAM = 1, HINT_NOALLOC = 2, HINT_L2EF = 3 };

// Loads are deliberately NOT volatile: the scheduler must be free to hoist all
// UNROLL loads of a tile above the first use (front-batched LDGs = memory-level
// parallelism).  Stores are volatile + "memory" so they are never elided.
#define GMP_LD128(QUAL)                                                                       \
    asm("ld.global" QUAL ".v4.u32 {%0, %1, %2, %3}, [%4];"                                    \
        : "=r"(v[0].x), "=r"(v[0].y), "=r"(v[0].z), "=r"(v[0].w)                             \
        : "l"(p))
#define GMP_ST128(QUAL)                                                                       \
    asm volatile("st.global" QUAL ".v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v[0].x),   \
                 "r"(v[0].y), "r"(v[0].z), "r"(v[0].w)                                        \
                 : "memory")
#define GMP_LD256(QUAL)                                                                       \
    do {                                                                                      \
        uint64_t a__, b__, c__, d__;                                                          \
        asm("ld.global" QUAL ".v4.u64 {%0, %1, %2, %3}, [%4];"                                \
            : "=l"(a__), "=l"(b__), "=l"(c__), "=l"(d__)                                      \
            : "l"(p));                                                                        \
        v[0] = make_uint4((uint32_t)a__, (uint32_t)(a__ >> 32), (uint32_t)b__, (uint32_t)(b__ >> 32)); \
        v[1] = make_uint4((uint32_t)c__, (uint32_t)(c__ >> 32), (uint32_t)d__, (uint32_t)(d__ >> 32)); \
    } while (0)
#define GMP_ST256(QUAL)                                                                       \
    do {                                                                                      \
        const uint64_t a__ = ((uint64_t)v[0].y << 32) | v[0].x, b__ = ((uint64_t)v[0].w << 32) | v[0].z; \
        const uint64_t c__ = ((uint64_t)v[1].y << 32) | v[1].x, d__ = ((uint64_t)v[1].w << 32) | v[1].z; \
        asm volatile("st.global" QUAL ".v4.u64 [%0], {%1, %2, %3, %4};" ::"l"(p), "l"(a__), "l"(b__), \
                     "l"(c__), "l"(d__)                                                       \
                     : "memory");                                                             \
    } while (0)

template <int VEC, int HINT>
__device__ __forceinline__ void ld_chunk(const uint4* p, uint4 (&v)[VEC]) {
    static_assert(VEC == 1 || VEC == 2, "chunk is 16 or 32 bytes");
    if (VEC == 1) {
        if (HINT == HINT_STREAM) GMP_LD128(".cs");
        else if (HINT == HINT_NOALLOC) GMP_LD128(".nc.L1::no_allocate");
        else if (HINT == HINT_L2EF) GMP_LD128(".cs");  // L2::evict_first exists for 256-bit accesses only
        else GMP_LD128("");
    } else {
        if (HINT == HINT_STREAM) GMP_LD256(".cs");
        else if (HINT == HINT_NOALLOC) GMP_LD256(".nc.L1::no_allocate");
        else if (HINT == HINT_L2EF) GMP_LD256(".L2::evict_first");
        else GMP_LD256("");
    }
}

template <int VEC, int HINT>
__device__ __forceinline__ void st_chunk(uint4* p, const uint4 (&v)[VEC]) {
    if (VEC == 1) {
        if (HINT == HINT_STREAM) GMP_ST128(".cs");
        else if (HINT == HINT_NOALLOC) GMP_ST128(".L1::no_allocate");
        else if (HINT == HINT_L2EF) GMP_ST128(".cs");
        else GMP_ST128("");
    } else {
        if (HINT == HINT_STREAM) GMP_ST256(".cs");
        else if (HINT == HINT_NOALLOC) GMP_ST256(".L1::no_allocate");
        else if (HINT == HINT_L2EF) GMP_ST256(".L2::evict_first");
        else GMP_ST256("");
    }
}

__device__ __forceinline__ uint4 pack(uint64_t w0, uint64_t w1) {
    return make_uint4((uint32_t)w0, (uint32_t)(w0 >> 32), (uint32_t)w1, (uint32_t)(w1 >> 32));
}

template <int PAT>
__device__ __forceinline__ void write_tail(uint4* dst, uint64_t nlanes, uint32_t tail_bytes, uint64_t seed) {
    // ragged last lane: its leading tail_bytes bytes, written by one thread
    uint64_t w0, w1;
    lane_words<PAT>(seed, nlanes, w0, w1);
    unsigned char* p = reinterpret_cast<unsigned char*>(dst + nlanes);
    for (uint32_t b = 0; b < tail_bytes; ++b) {
        const uint64_t w = (b < 8) ? w0 : w1;
        p[b] = (unsigned char)(w >> (8 * (b & 7)));
    }
}

// Tile walk shared by the three kernels.  A tile is THREADS*UNROLL chunks; thread t
// owns chunks t, t+THREADS, ... of the tile, so every warp instruction touches
// 32*VEC*16 contiguous bytes.  Lanes past the last full chunk row are handled one
// 16-byte lane at a time by the slow path.
#define GMP_TILE_PROLOGUE                                                          \
    constexpr uint64_t kTileLanes = (uint64_t)THREADS * UNROLL * VEC;              \
    const uint64_t ntiles = (nlanes + kTileLanes - 1) / kTileLanes;

// ---- write --------------------------------------------------------------------
template <int PAT, int THREADS, int UNROLL, int VEC, int HINT>
__global__ void __launch_bounds__(THREADS, 1024 / THREADS)
    k_hbm_write_pattern(uint4* __restrict__ dst, uint64_t nlanes, uint32_t tail_bytes, uint64_t seed) {
    GMP_TILE_PROLOGUE
    for (uint64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const uint64_t base = t * kTileLanes + (uint64_t)threadIdx.x * VEC;
        if ((t + 1) * kTileLanes <= nlanes) {
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) {
                const uint64_t L = base + (uint64_t)j * THREADS * VEC;
                uint4 v[VEC];
#pragma unroll
                for (int k = 0; k < VEC; ++k) {
                    uint64_t w0, w1;
                    lane_words<PAT>(seed, L + k, w0, w1);
                    v[k] = pack(w0, w1);
                }
                st_chunk<VEC, HINT>(dst + L, v);
            }
        } else {
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) {
#pragma unroll
                for (int k = 0; k < VEC; ++k) {
                    const uint64_t L = base + (uint64_t)j * THREADS * VEC + k;
                    if (L < nlanes) {
                        uint64_t w0, w1;
                        lane_words<PAT>(seed, L, w0, w1);
                        uint4 v[1] = {pack(w0, w1)};
                        st_chunk<1, HINT>(dst + L, v);
                    }
                }
            }
        }
    }
    if (tail_bytes != 0 && blockIdx.x == 0 && threadIdx.x == 0) write_tail<PAT>(dst, nlanes, tail_bytes, seed);
}

// ---- read + verify --------------------------------------------------------------
// acc[0] += mismatching lanes; acc[1] = max(UINT64_MAX - bad lane) (0 = none);
// acc[2] ^= words; acc[3] += words.  All four combine commutatively, so the
// result is independent of CTA scheduling and bit-exact against the oracle.
__device__ __forceinline__ uint64_t shfl_xor64(uint64_t v, int m) {
    uint32_t lo = __shfl_xor_sync(0xffffffffu, (uint32_t)v, m);
    uint32_t hi = __shfl_xor_sync(0xffffffffu, (uint32_t)(v >> 32), m);
    return ((uint64_t)hi << 32) | lo;
}

template <int THREADS>
__device__ __forceinline__ void block_fold_and_publish(uint64_t bad, uint64_t inv_first, uint64_t x, uint64_t s,
                                                       unsigned long long* __restrict__ acc) {
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) {
        bad += shfl_xor64(bad, m);
        const uint64_t o = shfl_xor64(inv_first, m);
        inv_first = o > inv_first ? o : inv_first;
        x ^= shfl_xor64(x, m);
        s += shfl_xor64(s, m);
    }
    constexpr int kWarps = THREADS / 32;
    __shared__ uint64_t sh[4][kWarps];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
        sh[0][warp] = bad;
        sh[1][warp] = inv_first;
        sh[2][warp] = x;
        sh[3][warp] = s;
    }
    __syncthreads();
    if (warp == 0) {
        bad = lane < kWarps ? sh[0][lane] : 0;
        inv_first = lane < kWarps ? sh[1][lane] : 0;
        x = lane < kWarps ? sh[2][lane] : 0;
        s = lane < kWarps ? sh[3][lane] : 0;
#pragma unroll
        for (int m = 16; m > 0; m >>= 1) {
            bad += shfl_xor64(bad, m);
            const uint64_t o = shfl_xor64(inv_first, m);
            inv_first = o > inv_first ? o : inv_first;
            x ^= shfl_xor64(x, m);
            s += shfl_xor64(s, m);
        }
        if (lane == 0) {
            if (bad) atomicAdd(acc + 0, (unsigned long long)bad);
            if (inv_first) atomicMax(acc + 1, (unsigned long long)inv_first);
            atomicXor(acc + 2, (unsigned long long)x);
            atomicAdd(acc + 3, (unsigned long long)s);
        }
    }
}

// Slow, exact accounting of one lane (used for ragged edges and after a tile has
// been found dirty).
template <int PAT>
__device__ __forceinline__ void check_lane(const uint4& v, uint64_t seed, uint64_t L, uint64_t& bad,
                                           uint64_t& inv_first, uint64_t& x, uint64_t& s) {
    uint64_t e0, e1;
    lane_words<PAT>(seed, L, e0, e1);
    const uint64_t g0 = ((uint64_t)v.y << 32) | v.x;
    const uint64_t g1 = ((uint64_t)v.w << 32) | v.z;
    x ^= g0 ^ g1;
    s += g0 + g1;
    if ((g0 ^ e0) | (g1 ^ e1)) {
        bad += 1;
        const uint64_t inv = ~0ull - L;
        inv_first = inv > inv_first ? inv : inv_first;
    }
}

// Fast path of one lane: checksums, and OR the lane's difference bits into `dirty`.
// No per-lane predicate or counter: a clean buffer (the overwhelmingly common case)
// pays four LOP3s here instead of a compare + predicated count + first-lane select;
// a tile whose `dirty` ends up non-zero is re-walked by count_bad_lane().
template <int PAT>
__device__ __forceinline__ void fold_lane(const uint4& v, uint64_t seed, uint64_t L, uint32_t& dirty, uint64_t& x,
                                          uint64_t& s) {
    uint64_t e0, e1;
    lane_words<PAT>(seed, L, e0, e1);
    const uint64_t g0 = ((uint64_t)v.y << 32) | v.x;
    const uint64_t g1 = ((uint64_t)v.w << 32) | v.z;
    x ^= g0 ^ g1;
    s += g0 + g1;
    const uint64_t d = (g0 ^ e0) | (g1 ^ e1);
    dirty |= (uint32_t)d | (uint32_t)(d >> 32);
}

template <int PAT>
__device__ __noinline__ void count_bad_lane(const uint4 v, uint64_t seed, uint64_t L, uint64_t* bad,
                                            uint64_t* inv_first) {
    uint64_t e0, e1;
    lane_words<PAT>(seed, L, e0, e1);
    const uint64_t g0 = ((uint64_t)v.y << 32) | v.x;
    const uint64_t g1 = ((uint64_t)v.w << 32) | v.z;
    if ((g0 ^ e0) | (g1 ^ e1)) {
        *bad += 1;
        const uint64_t inv = ~0ull - L;
        *inv_first = inv > *inv_first ? inv : *inv_first;
    }
}

template <int PAT, int THREADS, int UNROLL, int VEC, int HINT>
__global__ void __launch_bounds__(THREADS, 1024 / THREADS)
    k_hbm_read_verify(const uint4* __restrict__ src, uint64_t nlanes, uint32_t tail_bytes, uint64_t seed,
                      unsigned long long* __restrict__ acc) {
    GMP_TILE_PROLOGUE
    uint64_t bad = 0, inv_first = 0, x = 0, s = 0;
    for (uint64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const uint64_t base = t * kTileLanes + (uint64_t)threadIdx.x * VEC;
        if ((t + 1) * kTileLanes <= nlanes) {
            uint4 v[UNROLL][VEC];
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) ld_chunk<VEC, HINT>(src + base + (uint64_t)j * THREADS * VEC, v[j]);
            uint32_t dirty = 0;
#pragma unroll
            for (int j = 0; j < UNROLL; ++j)
#pragma unroll
                for (int k = 0; k < VEC; ++k)
                    fold_lane<PAT>(v[j][k], seed, base + (uint64_t)j * THREADS * VEC + k, dirty, x, s);
            if (dirty != 0) {  // rare: some lane of this thread's tile differs, account for it exactly
#pragma unroll
                for (int j = 0; j < UNROLL; ++j)
#pragma unroll
                    for (int k = 0; k < VEC; ++k)
                        count_bad_lane<PAT>(v[j][k], seed, base + (uint64_t)j * THREADS * VEC + k, &bad, &inv_first);
            }
        } else {
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) {
#pragma unroll
                for (int k = 0; k < VEC; ++k) {
                    const uint64_t L = base + (uint64_t)j * THREADS * VEC + k;
                    if (L < nlanes) {
                        uint4 v[1];
                        ld_chunk<1, HINT>(src + L, v);
                        check_lane<PAT>(v[0], seed, L, bad, inv_first, x, s);
                    }
                }
            }
        }
    }
    if (tail_bytes != 0 && blockIdx.x == 0 && threadIdx.x == 0) {
        // ragged last lane: compare its leading bytes, checksum them zero-padded
        uint64_t e0, e1;
        lane_words<PAT>(seed, nlanes, e0, e1);
        const unsigned char* p = reinterpret_cast<const unsigned char*>(src + nlanes);
        uint64_t g0 = 0, g1 = 0;
        for (uint32_t b = 0; b < tail_bytes; ++b) {
            const uint64_t byte = p[b];
            if (b < 8)
                g0 |= byte << (8 * b);
            else
                g1 |= byte << (8 * (b - 8));
        }
        const uint64_t m0 = tail_bytes >= 8 ? ~0ull : ((1ull << (8 * tail_bytes)) - 1);
        const uint64_t m1 = tail_bytes <= 8 ? 0ull : ((1ull << (8 * (tail_bytes - 8))) - 1);
        x ^= g0 ^ g1;
        s += g0 + g1;
        if (((g0 ^ e0) & m0) | ((g1 ^ e1) & m1)) {
            bad += 1;
            const uint64_t inv = ~0ull - nlanes;
            inv_first = inv > inv_first ? inv : inv_first;
        }
    }
    block_fold_and_publish<THREADS>(bad, inv_first, x, s, acc);
}

// ---- copy -----------------------------------------------------------------------
template <int THREADS, int UNROLL, int VEC, int HINT>
__global__ void __launch_bounds__(THREADS, 1024 / THREADS)
    k_hbm_copy(uint4* __restrict__ dst, const uint4* __restrict__ src, uint64_t nlanes, uint32_t tail_bytes) {
    GMP_TILE_PROLOGUE
    for (uint64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const uint64_t base = t * kTileLanes + (uint64_t)threadIdx.x * VEC;
        if ((t + 1) * kTileLanes <= nlanes) {
            uint4 v[UNROLL][VEC];
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) ld_chunk<VEC, HINT>(src + base + (uint64_t)j * THREADS * VEC, v[j]);
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) st_chunk<VEC, HINT>(dst + base + (uint64_t)j * THREADS * VEC, v[j]);
        } else {
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) {
#pragma unroll
                for (int k = 0; k < VEC; ++k) {
                    const uint64_t L = base + (uint64_t)j * THREADS * VEC + k;
                    if (L < nlanes) {
                        uint4 v[1];
                        ld_chunk<1, HINT>(src + L, v);
                        st_chunk<1, HINT>(dst + L, v);
                    }
                }
            }
        }
    }
    if (tail_bytes != 0 && blockIdx.x == 0 && threadIdx.x == 0) {
        const unsigned char* sp = reinterpret_cast<const unsigned char*>(src + nlanes);
        unsigned char* dp = reinterpret_cast<unsigned char*>(dst + nlanes);
        for (uint32_t b = 0; b < tail_bytes; ++b) dp[b] = sp[b];
    }
}

// ---- copy through the TMA engine -------------------------------------------------------------
// cp.async.bulk global -> shared (completion on an mbarrier) and shared -> global (bulk group), driven by ONE
// thread per CTA: no data ever touches a register, the LSU or the ALU; the bytes in flight per SM are
// (STAGES-1) x CHUNK of shared memory instead of registers.  SASS: UBLKCP.S.G / UBLKCP.G.S, SYNCS.
namespace bulk {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void store(void* gdst, const void* smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

}  // namespace bulk

template <int STAGES, int CHUNK>
__global__ void __launch_bounds__(32) k_hbm_copy_bulk(unsigned char* __restrict__ dst, const unsigned char* __restrict__ src,
                                                      uint64_t nchunks) {
    extern __shared__ __align__(128) unsigned char stage_mem[];
    __shared__ __align__(8) uint64_t full[STAGES];
    if (threadIdx.x != 0) return;  // the engine needs one issuing thread; the warp's other lanes leave
    for (int i = 0; i < STAGES; ++i) bulk::mbar_init(&full[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");

    const uint64_t first = blockIdx.x, stride = gridDim.x;
    const uint64_t mine = nchunks > first ? (nchunks - first + stride - 1) / stride : 0;
    auto gsrc = [&](uint64_t k) { return src + (first + k * stride) * (uint64_t)CHUNK; };
    auto gdst = [&](uint64_t k) { return dst + (first + k * stride) * (uint64_t)CHUNK; };
    auto issue_load = [&](uint64_t k) {
        const int st = (int)(k % STAGES);
        bulk::mbar_expect_tx(&full[st], CHUNK);
        bulk::load(stage_mem + (size_t)st * CHUNK, gsrc(k), CHUNK, &full[st]);
    };
    const uint64_t prologue = mine < (uint64_t)STAGES ? mine : (uint64_t)STAGES;
    for (uint64_t k = 0; k < prologue; ++k) issue_load(k);
    for (uint64_t k = 0; k < mine; ++k) {
        const int st = (int)(k % STAGES);
        bulk::mbar_wait(&full[st], (uint32_t)((k / STAGES) & 1));
        bulk::store(gdst(k), stage_mem + (size_t)st * CHUNK, CHUNK);
        bulk::commit();
        // refill the stage of chunk k-1: its store must have finished READING shared memory, which is the case once
        // at most one bulk group (the store just issued) is still pending
        if (k >= 1 && k - 1 + STAGES < mine) {
            bulk::wait_read<1>();
            issue_load(k - 1 + STAGES);
        }
    }
    bulk::wait_all();  // writes complete before the CTA (and its shared memory) goes away
}

// ---- pattern write through the TMA engine (peer stores) -------------------------------------------
// Vector stores into PEER memory plateau at ~713 GB/s per direction whatever the geometry (profiles/r01_sweep_p2p.json)
// while the copy engines move 777 GB/s over the same NVLink pair.  Here the CTA's threads only generate the pattern -
// into shared memory - and ONE thread hands each finished CHUNK to the bulk-copy engine (cp.async.bulk shared -> global),
// which issues the largest write requests the fabric takes; STAGES chunks are in flight per CTA.  The last chunk may be
// short (any multiple of 16 bytes).  SASS: STS.128, UBLKCP.G.S, no STG at all.
template <int PAT, int THREADS, int STAGES, int CHUNK>
__global__ void __launch_bounds__(THREADS) k_hbm_write_pattern_bulk(unsigned char* __restrict__ dst, uint64_t nlanes, uint64_t seed) {
    extern __shared__ __align__(128) unsigned char stage_mem[];
    constexpr uint64_t kChunkLanes = CHUNK / 16;
    const uint64_t nchunks = (nlanes + kChunkLanes - 1) / kChunkLanes;
    uint64_t k = 0;
    for (uint64_t c = blockIdx.x; c < nchunks; c += gridDim.x, ++k) {
        const int st = (int)(k % STAGES);
        if (k >= (uint64_t)STAGES) {
            // the bulk store that last used this stage must have finished READING it
            if (threadIdx.x == 0) bulk::wait_read<STAGES - 1>();
            __syncthreads();
        }
        uint4* stage = reinterpret_cast<uint4*>(stage_mem + (size_t)st * CHUNK);
        const uint64_t L0 = c * kChunkLanes;
        const uint32_t lanes = (uint32_t)(nlanes - L0 < kChunkLanes ? nlanes - L0 : kChunkLanes);
#pragma unroll 4
        for (uint32_t l = threadIdx.x; l < lanes; l += THREADS) {
            uint64_t w0, w1;
            lane_words<PAT>(seed, L0 + l, w0, w1);
            stage[l] = pack(w0, w1);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the bulk engine
        __syncthreads();
        if (threadIdx.x == 0) {
            bulk::store(dst + L0 * 16, stage, lanes * 16u);
            bulk::commit();
        }
    }
    if (threadIdx.x == 0) bulk::wait_all();  // writes complete before the CTA (and its shared memory) goes away
}

// ---- all-to-all pattern push: ONE launch per sender ------------------------------------------------
// Every GPU of an n-GPU pod writes its slice into each of the n-1 peers.  Done as n-1 launches in a row a sender feeds
// one peer at a time and pays a launch tail per peer (8 GPUs, 128 MiB slices: 622 GB/s per GPU, 458 with 32 MiB slices);
// here consecutive tiles go to consecutive peers, so all n-1 NVSwitch routes of the sender carry traffic for the whole
// kernel and there is one tail.  Peer p of sender i is GPU (i+1+p) mod n: at any moment the senders aim at different
// receivers.  Same tile body and store instructions as k_hbm_write_pattern (1024 threads x 32-byte stores, .cs).
struct AllToAllArgs {
    uint4* dst[GMP_MAX_DEVICES];
    uint64_t seed[GMP_MAX_DEVICES];
    int npeers;
};

template <int PAT, int THREADS, int UNROLL, int VEC, int HINT>
__global__ void __launch_bounds__(THREADS, 1024 / THREADS) k_peer_alltoall_write(const __grid_constant__ AllToAllArgs a, uint64_t nlanes) {
    GMP_TILE_PROLOGUE
    const uint64_t total = ntiles * (uint64_t)a.npeers;
    for (uint64_t w = blockIdx.x; w < total; w += gridDim.x) {
        const int p = (int)(w % (uint64_t)a.npeers);
        const uint64_t t = w / (uint64_t)a.npeers;
        uint4* __restrict__ dst = a.dst[p];
        const uint64_t seed = a.seed[p];
        const uint64_t base = t * kTileLanes + (uint64_t)threadIdx.x * VEC;
        if ((t + 1) * kTileLanes <= nlanes) {
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) {
                const uint64_t L = base + (uint64_t)j * THREADS * VEC;
                uint4 v[VEC];
#pragma unroll
                for (int k = 0; k < VEC; ++k) {
                    uint64_t w0, w1;
                    lane_words<PAT>(seed, L + k, w0, w1);
                    v[k] = pack(w0, w1);
                }
                st_chunk<VEC, HINT>(dst + L, v);
            }
        } else {
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) {
#pragma unroll
                for (int k = 0; k < VEC; ++k) {
                    const uint64_t L = base + (uint64_t)j * THREADS * VEC + k;
                    if (L < nlanes) {
                        uint64_t w0, w1;
                        lane_words<PAT>(seed, L, w0, w1);
                        uint4 v[1] = {pack(w0, w1)};
                        st_chunk<1, HINT>(dst + L, v);
                    }
                }
            }
        }
    }
}

__global__ void k_hbm_corrupt(unsigned char* p, size_t off, unsigned char mask) { p[off] ^= mask; }

// ---- launch geometry --------------------------------------------------------------
static std::atomic<unsigned long long> g_kernel_launches{0};
static inline void count_launch() { g_kernel_launches.fetch_add(1, std::memory_order_relaxed); }

Tunables& tunables() {
    // Picked by tools/sweep_hbm.py on a B200 (profiles/r01_sweep.md): many small CTAs (several waves, so the
    // hardware balances the ~10% speed spread between SMs) for the store-heavy kernels, exactly one resident
    // wave for the register-accumulating verify kernel.
    static Tunables t = {/*write_ctas_per_sm*/ 24, /*read*/ 4, /*copy*/ 128,
                         /*write_variant*/ 12, /*read_variant*/ 10, /*copy_variant*/ 21,
                         // pointers into PEER memory (NVLink), picked by tools/sweep_p2p.py:
                         /*peer_store_variant*/ 11, /*peer_store_ctas_per_sm*/ 8,  // 1024x1x32B .cs: ~713 GB/s whatever the geometry
                         /*peer_load_variant*/ 1, /*peer_load_ctas_per_sm*/ 32};   // 256x4x16B .cs x 32 CTAs/SM: 781 GB/s (cudaMemcpyPeer: 770)
    return t;
}

cudaError_t current_device_info(DeviceInfo* info) {
    static std::atomic<int> sm_cache[GMP_MAX_DEVICES * 4];
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    int sms = 0;
    const bool cacheable = dev >= 0 && dev < GMP_MAX_DEVICES * 4;
    if (cacheable) sms = sm_cache[dev].load(std::memory_order_relaxed);
    if (sms == 0) {
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return e;
        if (cacheable) sm_cache[dev].store(sms, std::memory_order_relaxed);
    }
    info->ordinal = dev;
    info->sm_count = sms;
    return cudaSuccess;
}

// CTAs of `kernel` that fit on one SM at once (occupancy API, cached per instantiation).
template <auto Kernel, int THREADS>
static int resident_ctas_per_sm() {
    static std::atomic<int> cache{0};
    int v = cache.load(std::memory_order_relaxed);
    if (v == 0) {
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, Kernel, THREADS, 0) != cudaSuccess || v < 1) {
            (void)cudaGetLastError();
            v = 1;
        }
        cache.store(v, std::memory_order_relaxed);
    }
    return v;
}

// ctas_per_sm == 0 means "exactly one resident wave" (persistent kernel); > 0 is an
// explicit multiple of the SM count (several waves of smaller CTAs).
static unsigned grid_for(uint64_t nlanes, uint64_t tile, int sm_count, int ctas_per_sm, int resident) {
    uint64_t ntiles = (nlanes + tile - 1) / tile;
    if (ntiles == 0) ntiles = 1;  // still one CTA for the ragged tail
    const uint64_t per_sm = ctas_per_sm > 0 ? (uint64_t)ctas_per_sm : (uint64_t)resident;
    const uint64_t g = (uint64_t)sm_count * per_sm;
    return (unsigned)(ntiles < g ? ntiles : g);
}

// Variant table: (THREADS, UNROLL, VEC, HINT).  The defaults in tunables() were picked
// by tools/sweep_hbm.py on a B200 (profiles/); the others stay selectable through
// gmp_tune() so that the sweep can be re-run on new silicon.  VEC=2 variants need
// 32-byte-aligned pointers; the launchers fall back to kFallbackVariant otherwise.
#ifdef GMP_LEAN
// The standalone probe carries only the geometries it uses (defaults, peer-memory configs, the alignment
// fallback): a tenth of the device code, so the module loads faster on every freshly mounted GPU.
#define GMP_VARIANTS(X)            \
    X(1, 256, 4, 1, HINT_STREAM)   \
    X(10, 256, 2, 2, HINT_DEFAULT) \
    X(11, 1024, 1, 2, HINT_STREAM) \
    X(12, 1024, 2, 1, HINT_STREAM)
#else
#define GMP_VARIANTS(X)            \
    X(0, 256, 4, 1, HINT_DEFAULT)  \
    X(1, 256, 4, 1, HINT_STREAM)   \
    X(2, 256, 8, 1, HINT_STREAM)   \
    X(3, 512, 4, 1, HINT_STREAM)   \
    X(4, 256, 4, 1, HINT_NOALLOC)  \
    X(5, 512, 2, 1, HINT_STREAM)   \
    X(6, 128, 8, 1, HINT_STREAM)   \
    X(7, 256, 2, 2, HINT_STREAM)   \
    X(8, 256, 4, 2, HINT_STREAM)   \
    X(9, 512, 2, 2, HINT_STREAM)   \
    X(10, 256, 2, 2, HINT_DEFAULT) \
    X(11, 1024, 1, 2, HINT_STREAM) \
    X(12, 1024, 2, 1, HINT_STREAM) \
    X(13, 256, 1, 2, HINT_STREAM)  \
    X(14, 512, 1, 2, HINT_STREAM)  \
    X(15, 256, 4, 2, HINT_L2EF)
#endif
constexpr int kVariantCount = 16;
constexpr int kFallbackVariant = 1;

static bool variant_is_wide(int id) {
    switch (id) {
#define X(ID, T, U, V, H) \
    case ID:              \
        return V == 2;
        GMP_VARIANTS(X)
#undef X
        default:
            return false;
    }
}

static int pick_variant(int wanted, const void* p0, const void* p1) {
    const uintptr_t bits = reinterpret_cast<uintptr_t>(p0) | reinterpret_cast<uintptr_t>(p1);
    if (variant_is_wide(wanted) && (bits & 31u) != 0) return kFallbackVariant;
    return wanted;
}

template <int PAT>
static cudaError_t launch_write_pat(uint4* dst, uint64_t nlanes, uint32_t tail, uint64_t seed, int sm_count,
                                    cudaStream_t s, const LaunchCfg* cfg) {
    Tunables tn = tunables();
    if (cfg) {
        tn.write_variant = cfg->variant;
        tn.write_ctas_per_sm = cfg->ctas_per_sm;
    }
    switch (pick_variant(tn.write_variant, dst, nullptr)) {
#define X(ID, T, U, V, H)                                                                            \
    case ID:                                                                                         \
        k_hbm_write_pattern<PAT, T, U, V, H>                                                         \
            <<<grid_for(nlanes, (uint64_t)T * U * V, sm_count, tn.write_ctas_per_sm,                 \
                        resident_ctas_per_sm<k_hbm_write_pattern<PAT, T, U, V, H>, T>()),            \
               T, 0, s>>>(dst, nlanes, tail, seed);                                                  \
        break;
        GMP_VARIANTS(X)
#undef X
        default:
            return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

template <int PAT>
static cudaError_t launch_verify_pat(const uint4* src, uint64_t nlanes, uint32_t tail, uint64_t seed,
                                     unsigned long long* acc, int sm_count, cudaStream_t s, const LaunchCfg* cfg) {
    Tunables tn = tunables();
    if (cfg) {
        tn.read_variant = cfg->variant;
        tn.read_ctas_per_sm = cfg->ctas_per_sm;
    }
    switch (pick_variant(tn.read_variant, src, nullptr)) {
#define X(ID, T, U, V, H)                                                                           \
    case ID:                                                                                        \
        k_hbm_read_verify<PAT, T, U, V, H>                                                          \
            <<<grid_for(nlanes, (uint64_t)T * U * V, sm_count, tn.read_ctas_per_sm,                 \
                        resident_ctas_per_sm<k_hbm_read_verify<PAT, T, U, V, H>, T>()),             \
               T, 0, s>>>(src, nlanes, tail, seed, acc);                                            \
        break;
        GMP_VARIANTS(X)
#undef X
        default:
            return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

#define GMP_PATTERN_SWITCH(pattern, CALL)        \
    switch (pattern) {                           \
        case GMP_PAT_SPLITMIX:                   \
            return CALL(GMP_PAT_SPLITMIX);       \
        case GMP_PAT_ZERO:                       \
            return CALL(GMP_PAT_ZERO);           \
        case GMP_PAT_ONES:                       \
            return CALL(GMP_PAT_ONES);           \
        case GMP_PAT_CHECKER:                    \
            return CALL(GMP_PAT_CHECKER);        \
        case GMP_PAT_ADDRESS:                    \
            return CALL(GMP_PAT_ADDRESS);        \
        default:                                 \
            return cudaErrorInvalidValue;        \
    }

LaunchCfg peer_store_cfg() { return {tunables().peer_store_variant, tunables().peer_store_ctas_per_sm}; }
LaunchCfg peer_load_cfg() { return {tunables().peer_load_variant, tunables().peer_load_ctas_per_sm}; }

// Write variants kBulkWriteFirst.. : pattern -> shared memory -> bulk store (k_hbm_write_pattern_bulk).
constexpr int kBulkWriteFirst = 32;
#ifdef GMP_LEAN
#define GMP_BULK_WRITE_VARIANTS(X) X(1, 256, 4, 32768)
#else
#define GMP_BULK_WRITE_VARIANTS(X) \
    X(0, 256, 4, 16384)           \
    X(1, 256, 4, 32768)           \
    X(2, 512, 4, 32768)           \
    X(3, 256, 2, 65536)           \
    X(4, 512, 3, 65536)           \
    X(5, 128, 8, 8192)            \
    X(6, 1024, 2, 98304)          \
    X(7, 256, 6, 32768)
#endif
constexpr int kBulkWriteCount = 8;

template <int PAT, int THREADS, int STAGES, int CHUNK>
static cudaError_t launch_write_bulk_cfg(unsigned char* dst, uint64_t nlanes, uint64_t seed, int sm_count, int ctas_per_sm,
                                         cudaStream_t s) {
    constexpr int kSmem = STAGES * CHUNK;
    // set on every launch: the attribute lives in the context, which callers may have reset since the last call
    cudaError_t e = cudaFuncSetAttribute(k_hbm_write_pattern_bulk<PAT, THREADS, STAGES, CHUNK>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
    if (e != cudaSuccess) return e;
    const uint64_t nchunks = (nlanes * 16 + CHUNK - 1) / CHUNK;
    const int fit = (227 * 1024) / (kSmem + 1024);
    const int per_sm = ctas_per_sm > 0 ? ctas_per_sm : (fit > 0 ? fit : 1);
    uint64_t grid = (uint64_t)sm_count * (uint64_t)per_sm;
    if (grid > nchunks) grid = nchunks;
    k_hbm_write_pattern_bulk<PAT, THREADS, STAGES, CHUNK><<<(unsigned)grid, THREADS, kSmem, s>>>(dst, nlanes, seed);
    return cudaGetLastError();
}

template <int PAT>
static cudaError_t launch_write_bulk_pat(int variant, unsigned char* dst, uint64_t nlanes, uint64_t seed, int sm_count,
                                         int ctas_per_sm, cudaStream_t s) {
    switch (variant - kBulkWriteFirst) {
#define X(ID, T, ST, CH) \
    case ID:             \
        return launch_write_bulk_cfg<PAT, T, ST, CH>(dst, nlanes, seed, sm_count, ctas_per_sm, s);
        GMP_BULK_WRITE_VARIANTS(X)
#undef X
        default:
            return cudaErrorNotSupported;
    }
}

cudaError_t launch_write(void* dptr, size_t bytes, int pattern, uint64_t seed, cudaStream_t s, const LaunchCfg* cfg) {
    if (bytes == 0) return cudaSuccess;
    DeviceInfo di;
    cudaError_t e = current_device_info(&di);
    if (e != cudaSuccess) return e;
    const uint64_t nlanes = bytes / 16;
    const uint32_t tail = (uint32_t)(bytes % 16);
    count_launch();
    const int variant = cfg ? cfg->variant : tunables().write_variant;
    LaunchCfg fallback = {kFallbackVariant, cfg ? cfg->ctas_per_sm : tunables().write_ctas_per_sm};
    if (variant >= kBulkWriteFirst) {
        // whole 16-byte lanes through the bulk engine; a ragged buffer (never the case for the P2P probe) or a pattern
        // / geometry this build does not carry goes through the vector kernel
        if (tail == 0 && nlanes != 0 && (reinterpret_cast<uintptr_t>(dptr) & 15u) == 0) {
            const int ctas = cfg ? cfg->ctas_per_sm : tunables().write_ctas_per_sm;
            cudaError_t be = cudaErrorNotSupported;
            switch (pattern) {
                case GMP_PAT_SPLITMIX:
                    be = launch_write_bulk_pat<GMP_PAT_SPLITMIX>(variant, static_cast<unsigned char*>(dptr), nlanes, seed, di.sm_count, ctas, s);
                    break;
#ifndef GMP_LEAN
                case GMP_PAT_ZERO:
                    be = launch_write_bulk_pat<GMP_PAT_ZERO>(variant, static_cast<unsigned char*>(dptr), nlanes, seed, di.sm_count, ctas, s);
                    break;
                case GMP_PAT_ONES:
                    be = launch_write_bulk_pat<GMP_PAT_ONES>(variant, static_cast<unsigned char*>(dptr), nlanes, seed, di.sm_count, ctas, s);
                    break;
                case GMP_PAT_CHECKER:
                    be = launch_write_bulk_pat<GMP_PAT_CHECKER>(variant, static_cast<unsigned char*>(dptr), nlanes, seed, di.sm_count, ctas, s);
                    break;
                case GMP_PAT_ADDRESS:
                    be = launch_write_bulk_pat<GMP_PAT_ADDRESS>(variant, static_cast<unsigned char*>(dptr), nlanes, seed, di.sm_count, ctas, s);
                    break;
#endif
                default:
                    break;
            }
            if (be != cudaErrorNotSupported) return be;
        }
        cfg = &fallback;
    }
#define CALL(P) launch_write_pat<P>(static_cast<uint4*>(dptr), nlanes, tail, seed, di.sm_count, s, cfg)
    GMP_PATTERN_SWITCH(pattern, CALL)
#undef CALL
}

cudaError_t launch_verify(const void* dptr, size_t bytes, int pattern, uint64_t seed, unsigned long long* d_acc,
                          cudaStream_t s, const LaunchCfg* cfg) {
    if (pattern < 0 || pattern >= GMP_PAT_COUNT) return cudaErrorInvalidValue;
    cudaError_t e = cudaMemsetAsync(d_acc, 0, 4 * sizeof(unsigned long long), s);
    if (e != cudaSuccess) return e;
    if (bytes == 0) return cudaSuccess;
    DeviceInfo di;
    e = current_device_info(&di);
    if (e != cudaSuccess) return e;
    const uint64_t nlanes = bytes / 16;
    const uint32_t tail = (uint32_t)(bytes % 16);
    count_launch();
#define CALL(P) launch_verify_pat<P>(static_cast<const uint4*>(dptr), nlanes, tail, seed, d_acc, di.sm_count, s, cfg)
    GMP_PATTERN_SWITCH(pattern, CALL)
#undef CALL
}

cudaError_t launch_alltoall_write(void* const* dsts, const uint64_t* seeds, int npeers, size_t slice_bytes, cudaStream_t s) {
    if (npeers < 1 || npeers > GMP_MAX_DEVICES || (slice_bytes & 31u) != 0) return cudaErrorInvalidValue;
    if (slice_bytes == 0) return cudaSuccess;
    DeviceInfo di;
    cudaError_t e = current_device_info(&di);
    if (e != cudaSuccess) return e;
    AllToAllArgs a;
    for (int p = 0; p < GMP_MAX_DEVICES; ++p) {
        a.dst[p] = p < npeers ? static_cast<uint4*>(dsts[p]) : nullptr;
        a.seed[p] = p < npeers ? seeds[p] : 0;
        if (p < npeers && (reinterpret_cast<uintptr_t>(dsts[p]) & 31u) != 0) return cudaErrorInvalidValue;
    }
    a.npeers = npeers;
    const uint64_t nlanes = slice_bytes / 16;
    constexpr int T = 1024, U = 1, V = 2;
    const uint64_t tiles = ((nlanes + (uint64_t)T * U * V - 1) / ((uint64_t)T * U * V)) * (uint64_t)npeers;
    const int per_sm = tunables().peer_store_ctas_per_sm > 0 ? tunables().peer_store_ctas_per_sm : 8;
    uint64_t grid = (uint64_t)di.sm_count * (uint64_t)per_sm;
    if (grid > tiles) grid = tiles;
    count_launch();
    k_peer_alltoall_write<GMP_PAT_SPLITMIX, T, U, V, HINT_STREAM><<<(unsigned)grid, T, 0, s>>>(a, nlanes);
    return cudaGetLastError();
}

constexpr int kBulkVariantFirst = kVariantCount;  // copy variants 16..23

template <int STAGES, int CHUNK>
static cudaError_t launch_copy_bulk_cfg(unsigned char* dst, const unsigned char* src, size_t bytes, int sm_count, int ctas_per_sm,
                                        cudaStream_t s) {
    constexpr int kSmem = STAGES * CHUNK;
    {
        // set on every launch: the attribute lives in the context, and callers may have reset the device since the
        // last call (gmp_probe_hbm does, by default); the call costs about a microsecond
        cudaError_t e = cudaFuncSetAttribute(k_hbm_copy_bulk<STAGES, CHUNK>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
        if (e != cudaSuccess) return e;
    }
    const uint64_t nchunks = bytes / CHUNK;
    if (nchunks > 0) {
        const int fit = (227 * 1024) / (kSmem + 1024);
        const int per_sm = ctas_per_sm > 0 ? ctas_per_sm : (fit > 0 ? fit : 1);
        uint64_t grid = (uint64_t)sm_count * (uint64_t)per_sm;
        if (grid > nchunks) grid = nchunks;
        k_hbm_copy_bulk<STAGES, CHUNK><<<(unsigned)grid, 32, kSmem, s>>>(dst, src, nchunks);
        count_launch();
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    const size_t done = (size_t)nchunks * CHUNK;
    if (done < bytes) {
        const LaunchCfg rest = {kFallbackVariant, 0};
        return launch_copy(dst + done, src + done, bytes - done, s, &rest);
    }
    return cudaSuccess;
}

static cudaError_t launch_copy_bulk(int variant, void* dst, const void* src, size_t bytes, int sm_count, int ctas_per_sm,
                                    cudaStream_t s) {
    if (((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src)) & 15u) != 0) return cudaErrorNotSupported;
    unsigned char* d = static_cast<unsigned char*>(dst);
    const unsigned char* c = static_cast<const unsigned char*>(src);
    switch (variant - kBulkVariantFirst) {
#ifndef GMP_LEAN
        case 0: return launch_copy_bulk_cfg<6, 16384>(d, c, bytes, sm_count, ctas_per_sm, s);
        case 1: return launch_copy_bulk_cfg<3, 32768>(d, c, bytes, sm_count, ctas_per_sm, s);
        case 2: return launch_copy_bulk_cfg<12, 8192>(d, c, bytes, sm_count, ctas_per_sm, s);
        case 3: return launch_copy_bulk_cfg<3, 65536>(d, c, bytes, sm_count, ctas_per_sm, s);
        case 4: return launch_copy_bulk_cfg<2, 98304>(d, c, bytes, sm_count, ctas_per_sm, s);
        case 6: return launch_copy_bulk_cfg<4, 49152>(d, c, bytes, sm_count, ctas_per_sm, s);
        case 7: return launch_copy_bulk_cfg<7, 32768>(d, c, bytes, sm_count, ctas_per_sm, s);
#endif
        case 5: return launch_copy_bulk_cfg<2, 114688>(d, c, bytes, sm_count, ctas_per_sm, s);
        default: return cudaErrorNotSupported;
    }
}

cudaError_t launch_copy(void* dst, const void* src, size_t bytes, cudaStream_t s, const LaunchCfg* cfg) {
    if (bytes == 0) return cudaSuccess;
    DeviceInfo di;
    cudaError_t e = current_device_info(&di);
    if (e != cudaSuccess) return e;
    const uint64_t nlanes = bytes / 16;
    const uint32_t tail = (uint32_t)(bytes % 16);
    Tunables tn = tunables();
    if (cfg) {
        tn.copy_variant = cfg->variant;
        tn.copy_ctas_per_sm = cfg->ctas_per_sm;
    }
    if (tn.copy_variant >= kVariantCount) {
        // TMA-bulk variants: whole chunks through the copy engine, the remainder through the vector kernel
        cudaError_t be = launch_copy_bulk(tn.copy_variant, dst, src, bytes, di.sm_count, tn.copy_ctas_per_sm, s);
        if (be != cudaErrorNotSupported) return be;
        tn.copy_variant = kFallbackVariant;
    }
    count_launch();
    switch (pick_variant(tn.copy_variant, dst, src)) {
#define X(ID, T, U, V, H)                                                                          \
    case ID:                                                                                       \
        k_hbm_copy<T, U, V, H><<<grid_for(nlanes, (uint64_t)T * U * V, di.sm_count, tn.copy_ctas_per_sm,       \
                                          resident_ctas_per_sm<k_hbm_copy<T, U, V, H>, T>()),     \
                                 T, 0, s>>>(static_cast<uint4*>(dst), static_cast<const uint4*>(src), nlanes, tail); \
        break;
        GMP_VARIANTS(X)
#undef X
        default:
            return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

void decode_acc(const unsigned long long raw[4], gmp_verify_result* out) {
    out->mismatch_lanes = raw[0];
    out->first_bad_lane = raw[1] ? (~0ull - raw[1]) : ~0ull;
    out->xor64 = raw[2];
    out->sum64 = raw[3];
}

void host_lane_words(int pattern, uint64_t seed, uint64_t L, uint64_t& w0, uint64_t& w1) {
    switch (pattern) {
        case GMP_PAT_SPLITMIX: lane_words<GMP_PAT_SPLITMIX>(seed, L, w0, w1); break;
        case GMP_PAT_ZERO: lane_words<GMP_PAT_ZERO>(seed, L, w0, w1); break;
        case GMP_PAT_ONES: lane_words<GMP_PAT_ONES>(seed, L, w0, w1); break;
        case GMP_PAT_CHECKER: lane_words<GMP_PAT_CHECKER>(seed, L, w0, w1); break;
        default: lane_words<GMP_PAT_ADDRESS>(seed, L, w0, w1); break;
    }
}

}  // namespace gmp

// ---- C ABI: section 3 of gpumounter_probe.h ----------------------------------------
using namespace gmp;

static int check_buffer(const void* p, size_t bytes, int pattern) {
    if (pattern < 0 || pattern >= GMP_PAT_COUNT) {
        set_error("unknown pattern %d", pattern);
        return GMP_ERR_INVALID;
    }
    if (bytes != 0 && p == nullptr) {
        set_error("NULL device pointer with bytes=%zu", bytes);
        return GMP_ERR_INVALID;
    }
    if ((reinterpret_cast<uintptr_t>(p) & 15u) != 0) {
        set_error("device pointer %p is not 16-byte aligned", p);
        return GMP_ERR_INVALID;
    }
    return GMP_OK;
}

extern "C" int gmp_hbm_write_pattern(void* dptr, size_t bytes, int pattern, uint64_t seed, void* stream) {
    clear_error();
    int rc = check_buffer(dptr, bytes, pattern);
    if (rc != GMP_OK) return rc;
    GMP_CUDA(launch_write(dptr, bytes, pattern, seed, static_cast<cudaStream_t>(stream)));
    return GMP_OK;
}

extern "C" int gmp_hbm_read_verify(const void* dptr, size_t bytes, int pattern, uint64_t seed, gmp_verify_result* out,
                                   void* stream) try {
    clear_error();
    if (out == nullptr) {
        set_error("out is NULL");
        return GMP_ERR_INVALID;
    }
    int rc = check_buffer(dptr, bytes, pattern);
    if (rc != GMP_OK) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    // scratch: the 32-byte device accumulator the kernel reduces into
    unsigned long long* d_acc = nullptr;
    GMP_CUDA(cudaMalloc(&d_acc, 4 * sizeof(unsigned long long)));
    unsigned long long raw[4] = {0, 0, 0, 0};
    cudaError_t e = launch_verify(dptr, bytes, pattern, seed, d_acc, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(raw, d_acc, sizeof(raw), cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    cudaError_t ef = cudaFree(d_acc);
    if (e != cudaSuccess) return cuda_fail(e, "gmp_hbm_read_verify");
    if (ef != cudaSuccess) return cuda_fail(ef, "cudaFree(accumulator)");
    decode_acc(raw, out);
    return GMP_OK;
} catch (...) {
    return ::gmp::caught_exception();
}

extern "C" int gmp_hbm_copy(void* dst, const void* src, size_t bytes, void* stream) {
    clear_error();
    int rc = check_buffer(dst, bytes, 0);
    if (rc == GMP_OK) rc = check_buffer(src, bytes, 0);
    if (rc != GMP_OK) return rc;
    GMP_CUDA(launch_copy(dst, src, bytes, static_cast<cudaStream_t>(stream)));
    return GMP_OK;
}

extern "C" int gmp_hbm_corrupt(void* dptr, size_t bytes, size_t byte_offset, uint8_t mask, void* stream) {
    clear_error();
    if (dptr == nullptr || byte_offset >= bytes) {
        set_error("corrupt: offset %zu outside buffer of %zu bytes", byte_offset, bytes);
        return GMP_ERR_INVALID;
    }
    k_hbm_corrupt<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<unsigned char*>(dptr), byte_offset, mask);
    GMP_CUDA(cudaGetLastError());
    return GMP_OK;
}

extern "C" int gmp_hbm_read_verify_async(const void* dptr, size_t bytes, int pattern, uint64_t seed, void* d_acc,
                                         void* stream) {
    clear_error();
    if (d_acc == nullptr || (reinterpret_cast<uintptr_t>(d_acc) & 7u) != 0) {
        set_error("d_acc must be 32 bytes of 8-byte-aligned device memory");
        return GMP_ERR_INVALID;
    }
    int rc = check_buffer(dptr, bytes, pattern);
    if (rc != GMP_OK) return rc;
    GMP_CUDA(launch_verify(dptr, bytes, pattern, seed, static_cast<unsigned long long*>(d_acc),
                           static_cast<cudaStream_t>(stream)));
    return GMP_OK;
}

extern "C" int gmp_hbm_verify_fetch(const void* d_acc, gmp_verify_result* out, void* stream) {
    clear_error();
    if (d_acc == nullptr || out == nullptr) {
        set_error("NULL argument");
        return GMP_ERR_INVALID;
    }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    unsigned long long raw[4];
    GMP_CUDA(cudaMemcpyAsync(raw, d_acc, sizeof(raw), cudaMemcpyDeviceToHost, s));
    GMP_CUDA(cudaStreamSynchronize(s));
    decode_acc(raw, out);
    return GMP_OK;
}

extern "C" unsigned long long gmp_kernel_launches(void) { return g_kernel_launches.load(std::memory_order_relaxed); }

extern "C" int gmp_tune(const char* knob, int value) try {
    clear_error();
    if (knob == nullptr) return GMP_ERR_INVALID;
    Tunables& t = tunables();
    int* slot = nullptr;
    int limit = 256;
    const std::string_view k(knob);
    if (k == "write_ctas_per_sm") slot = &t.write_ctas_per_sm;
    else if (k == "read_ctas_per_sm") slot = &t.read_ctas_per_sm;
    else if (k == "copy_ctas_per_sm") slot = &t.copy_ctas_per_sm;
    else if (k == "write_variant") { slot = &t.write_variant; limit = kBulkWriteFirst + kBulkWriteCount - 1; }
    else if (k == "read_variant") { slot = &t.read_variant; limit = kVariantCount - 1; }
    else if (k == "copy_variant") { slot = &t.copy_variant; limit = kVariantCount + 8 - 1; }
    else if (k == "peer_store_variant") { slot = &t.peer_store_variant; limit = kBulkWriteFirst + kBulkWriteCount - 1; }
    else if (k == "peer_store_ctas_per_sm") slot = &t.peer_store_ctas_per_sm;
    else if (k == "peer_load_variant") { slot = &t.peer_load_variant; limit = kVariantCount - 1; }
    else if (k == "peer_load_ctas_per_sm") slot = &t.peer_load_ctas_per_sm;
    if (slot == nullptr) {
        set_error("unknown knob '%s'", knob);
        return GMP_ERR_INVALID;
    }
    if (value >= 0) {
        if (value > limit) {
            set_error("knob '%s': value %d out of range", knob, value);
            return GMP_ERR_INVALID;
        }
        *slot = value;
    }
    return *slot;
} catch (...) {
    return ::gmp::caught_exception();
}

// ---- read-verify without any allocation -----------------------------------------------------
// gmp_hbm_read_verify() above obtains its 32-byte accumulator with cudaMalloc and returns it with
// cudaFree on every call.  Three kinds of caller of this library cannot afford that:
//   * a health-check sidecar re-verifying a buffer on a GPU whose memory the tenant has filled: the
//     allocator may simply answer out-of-memory, and a probe must not compete with the workload for it;
//   * a monitor that must not disturb the tenant: cudaFree synchronises the WHOLE device, so a periodic
//     check through the allocating entry point stalls every stream of the workload once per call;
//   * a caller whose own rules forbid allocator calls at that point (a latency-critical section of the worker, a
//     process whose device memory is budgeted up front).
// For them the scratch lives in device memory that exists from the moment the library is loaded: a
// file-scope array that every GPU using the library holds until the process exits (65 568 bytes).  The
// buffer is cut into 4 MiB segments, one CTA per segment writes that segment's partial verdict into its
// own slot (plain stores, no atomics), and a one-CTA fold kernel combines the slots into the result slot.
// 2048 slots => at most 8 GiB per call.
namespace gmp {

constexpr size_t kNoAllocSegmentBytes = (size_t)4 << 20;
constexpr int kNoAllocMaxSegments = 2048;
constexpr int kNoAllocResultSlot = kNoAllocMaxSegments;  // index (in 4-word slots) of the folded result
static_assert((unsigned long long)kNoAllocSegmentBytes * kNoAllocMaxSegments == GMP_NOALLOC_MAX_BYTES,
              "header limit and scratch size must agree");

// In HOST code this name is only a handle for cudaGetSymbolAddress(); it is not a device address.
static __device__ __align__(256) unsigned long long g_noalloc_scratch[(kNoAllocMaxSegments + 1) * 4];

template <int PAT>
__global__ void __launch_bounds__(256, 4)
    k_hbm_read_verify_segments(const uint4* __restrict__ src, uint64_t nlanes, uint32_t tail_bytes, uint64_t seed,
                               unsigned long long* __restrict__ slots) {
    constexpr int THREADS = 256, UNROLL = 4;
    constexpr uint64_t kSegLanes = kNoAllocSegmentBytes / 16;
    const uint64_t seg_begin = (uint64_t)blockIdx.x * kSegLanes;
    const uint64_t seg_end = seg_begin + kSegLanes < nlanes ? seg_begin + kSegLanes : nlanes;
    uint64_t bad = 0, inv_first = 0, x = 0, s = 0;
    for (uint64_t tile = seg_begin; tile < seg_end; tile += (uint64_t)THREADS * UNROLL) {
        const uint64_t base = tile + threadIdx.x;
        if (tile + (uint64_t)THREADS * UNROLL <= seg_end) {
            uint4 v[UNROLL][1];
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) ld_chunk<1, HINT_DEFAULT>(src + base + (uint64_t)j * THREADS, v[j]);
            uint32_t dirty = 0;
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) fold_lane<PAT>(v[j][0], seed, base + (uint64_t)j * THREADS, dirty, x, s);
            if (dirty != 0) {
#pragma unroll
                for (int j = 0; j < UNROLL; ++j)
                    count_bad_lane<PAT>(v[j][0], seed, base + (uint64_t)j * THREADS, &bad, &inv_first);
            }
        } else {
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) {
                const uint64_t L = base + (uint64_t)j * THREADS;
                if (L < seg_end) {
                    uint4 v[1];
                    ld_chunk<1, HINT_DEFAULT>(src + L, v);
                    check_lane<PAT>(v[0], seed, L, bad, inv_first, x, s);
                }
            }
        }
    }
    // the ragged last lane belongs to the last segment's CTA
    if (tail_bytes != 0 && blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) {
        uint64_t e0, e1;
        lane_words<PAT>(seed, nlanes, e0, e1);
        const unsigned char* p = reinterpret_cast<const unsigned char*>(src + nlanes);
        uint64_t g0 = 0, g1 = 0;
        for (uint32_t b = 0; b < tail_bytes; ++b) {
            const uint64_t byte = p[b];
            if (b < 8) g0 |= byte << (8 * b);
            else g1 |= byte << (8 * (b - 8));
        }
        const uint64_t m0 = tail_bytes >= 8 ? ~0ull : ((1ull << (8 * tail_bytes)) - 1);
        const uint64_t m1 = tail_bytes <= 8 ? 0ull : ((1ull << (8 * (tail_bytes - 8))) - 1);
        x ^= g0 ^ g1;
        s += g0 + g1;
        if (((g0 ^ e0) & m0) | ((g1 ^ e1) & m1)) {
            bad += 1;
            const uint64_t inv = ~0ull - nlanes;
            inv_first = inv > inv_first ? inv : inv_first;
        }
    }
    // fold the CTA and store its verdict into this segment's own slot (no atomics: the slot is ours alone)
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) {
        bad += shfl_xor64(bad, m);
        const uint64_t o = shfl_xor64(inv_first, m);
        inv_first = o > inv_first ? o : inv_first;
        x ^= shfl_xor64(x, m);
        s += shfl_xor64(s, m);
    }
    __shared__ uint64_t sh[4][THREADS / 32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
        sh[0][warp] = bad;
        sh[1][warp] = inv_first;
        sh[2][warp] = x;
        sh[3][warp] = s;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        bad = 0, inv_first = 0, x = 0, s = 0;
        for (int w = 0; w < THREADS / 32; ++w) {
            bad += sh[0][w];
            inv_first = sh[1][w] > inv_first ? sh[1][w] : inv_first;
            x ^= sh[2][w];
            s += sh[3][w];
        }
        unsigned long long* slot = slots + (size_t)blockIdx.x * 4;
        slot[0] = bad;
        slot[1] = inv_first;
        slot[2] = x;
        slot[3] = s;
    }
}

// One CTA: combine the first `nsegs` slots into the result slot, which the host cleared beforehand.
__global__ void __launch_bounds__(256) k_noalloc_fold(unsigned long long* __restrict__ slots, int nsegs) {
    uint64_t bad = 0, inv_first = 0, x = 0, s = 0;
    for (int k = threadIdx.x; k < nsegs; k += 256) {
        const unsigned long long* slot = slots + (size_t)k * 4;
        bad += slot[0];
        inv_first = slot[1] > inv_first ? slot[1] : inv_first;
        x ^= slot[2];
        s += slot[3];
    }
    block_fold_and_publish<256>(bad, inv_first, x, s, slots + (size_t)kNoAllocResultSlot * 4);
}

// Set-up for the CURRENT device, done by every call (a few microseconds once the module is resident):
//  * the array's device address.  Looked up each time instead of cached per device: the address belongs to the device's
//    current primary context, and a cudaDeviceReset() made by the caller - which this library cannot see - would turn a
//    cached pointer into a dangling one for the rest of the process;
//  * all six kernels of this entry point are made resident (cudaFuncGetAttributes loads a kernel's module without
//    launching it).  CUDA loads a kernel lazily at its first launch, and that load allocates device memory for the
//    code: without this step a first call with pattern A would leave the kernels of patterns B..E to be loaded - and
//    allocated - by some later call, on a GPU that may be full by then.  After ONE call on a device, whatever its
//    pattern or size, no later call on that device allocates anything.
static cudaError_t noalloc_setup_on_current_device(unsigned long long** out) {
    void* addr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&addr, g_noalloc_scratch);  // the current device's copy of the array
    if (e != cudaSuccess) return e;
    cudaFuncAttributes fa;
    if ((e = cudaFuncGetAttributes(&fa, k_hbm_read_verify_segments<GMP_PAT_SPLITMIX>)) != cudaSuccess) return e;
    if ((e = cudaFuncGetAttributes(&fa, k_hbm_read_verify_segments<GMP_PAT_ZERO>)) != cudaSuccess) return e;
    if ((e = cudaFuncGetAttributes(&fa, k_hbm_read_verify_segments<GMP_PAT_ONES>)) != cudaSuccess) return e;
    if ((e = cudaFuncGetAttributes(&fa, k_hbm_read_verify_segments<GMP_PAT_CHECKER>)) != cudaSuccess) return e;
    if ((e = cudaFuncGetAttributes(&fa, k_hbm_read_verify_segments<GMP_PAT_ADDRESS>)) != cudaSuccess) return e;
    if ((e = cudaFuncGetAttributes(&fa, k_noalloc_fold)) != cudaSuccess) return e;
    *out = static_cast<unsigned long long*>(addr);
    return cudaSuccess;
}

template <int PAT>
static cudaError_t launch_segments(const uint4* src, uint64_t nlanes, uint32_t tail, uint64_t seed,
                                   unsigned long long* slots, int nsegs, cudaStream_t s) {
    k_hbm_read_verify_segments<PAT><<<(unsigned)nsegs, 256, 0, s>>>(src, nlanes, tail, seed, slots);
    return cudaGetLastError();
}

}  // namespace gmp

extern "C" int gmp_hbm_read_verify_noalloc(const void* dptr, size_t bytes, int pattern, uint64_t seed,
                                           gmp_verify_result* out, void* stream) try {
    clear_error();
    if (out == nullptr) {
        set_error("out is NULL");
        return GMP_ERR_INVALID;
    }
    int rc = check_buffer(dptr, bytes, pattern);
    if (rc != GMP_OK) return rc;
    if (bytes > GMP_NOALLOC_MAX_BYTES) {
        set_error("%zu bytes exceed the %llu-byte limit of the allocation-free verify (static scratch holds %d segments "
                  "of %zu MiB); use gmp_hbm_read_verify or split the buffer",
                  bytes, (unsigned long long)GMP_NOALLOC_MAX_BYTES, kNoAllocMaxSegments, kNoAllocSegmentBytes >> 20);
        return GMP_ERR_INVALID;
    }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    unsigned long long* scratch = nullptr;
    GMP_CUDA(noalloc_setup_on_current_device(&scratch));
    unsigned long long* result = scratch + (size_t)kNoAllocResultSlot * 4;
    // The array still holds the previous call's slots and result.  Slots are fully rewritten by the CTAs that
    // own them and only the first nsegs are read back; the result slot is accumulated into, so it must be zero.
    GMP_CUDA(cudaMemsetAsync(result, 0, 4 * sizeof(unsigned long long), s));
    const uint64_t nlanes = bytes / 16;
    const uint32_t tail = (uint32_t)(bytes % 16);
    const int nsegs = (int)((bytes + kNoAllocSegmentBytes - 1) / kNoAllocSegmentBytes);
    if (nsegs > 0) {
        const uint4* src = static_cast<const uint4*>(dptr);
        cudaError_t e = cudaErrorInvalidValue;
        switch (pattern) {
            case GMP_PAT_SPLITMIX: e = launch_segments<GMP_PAT_SPLITMIX>(src, nlanes, tail, seed, scratch, nsegs, s); break;
            case GMP_PAT_ZERO: e = launch_segments<GMP_PAT_ZERO>(src, nlanes, tail, seed, scratch, nsegs, s); break;
            case GMP_PAT_ONES: e = launch_segments<GMP_PAT_ONES>(src, nlanes, tail, seed, scratch, nsegs, s); break;
            case GMP_PAT_CHECKER: e = launch_segments<GMP_PAT_CHECKER>(src, nlanes, tail, seed, scratch, nsegs, s); break;
            case GMP_PAT_ADDRESS: e = launch_segments<GMP_PAT_ADDRESS>(src, nlanes, tail, seed, scratch, nsegs, s); break;
        }
        GMP_CUDA(e);
        k_noalloc_fold<<<1, 256, 0, s>>>(scratch, nsegs);
        GMP_CUDA(cudaGetLastError());
    }
    unsigned long long raw[4] = {0, 0, 0, 0};
    GMP_CUDA(cudaMemcpyAsync(raw, result, sizeof(raw), cudaMemcpyDeviceToHost, s));
    GMP_CUDA(cudaStreamSynchronize(s));
    decode_acc(raw, out);
    return GMP_OK;
} catch (...) {
    return ::gmp::caught_exception();
}
