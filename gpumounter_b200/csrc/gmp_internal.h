// Internal declarations shared by the translation units of libgpumounter_probe.so.
// Nothing here is part of the ABI; the ABI is include/gpumounter_probe.h.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <stdarg.h>
#include <time.h>

#include "gpumounter_probe.h"

namespace gmp {

// ---- error plumbing ---------------------------------------------------------
void set_error(const char* fmt, ...) __attribute__((format(printf, 1, 2)));
void clear_error();
// For `extern "C" int f(...) try { ... } catch (...) { return ::gmp::caught_exception(); }` (gmp_nvml.cpp)
int caught_exception() noexcept;
// Maps a CUDA failure to a GMP code and records "what: cudaGetErrorString".
int cuda_fail(cudaError_t e, const char* what);

#define GMP_CUDA(call)                                          \
    do {                                                        \
        cudaError_t e__ = (call);                               \
        if (e__ != cudaSuccess) return ::gmp::cuda_fail(e__, #call); \
    } while (0)

static inline uint64_t now_ns() {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (uint64_t)ts.tv_sec * 1000000000ull + (uint64_t)ts.tv_nsec;
}

// ---- pattern v1 (see gpumounter_probe.h section 3) --------------------------
constexpr uint64_t kGamma = 0x9E3779B97F4A7C15ull;

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__host__ __device__ __forceinline__ uint64_t rotl32of64(uint64_t v) {
    return (v << 32) | (v >> 32);
}

// The two words of lane L.  PAT is a compile-time constant in the kernels and a
// runtime value in the (small) host-side generator used for the DMA check.
template <int PAT>
__host__ __device__ __forceinline__ void lane_words(uint64_t seed, uint64_t L, uint64_t& w0, uint64_t& w1) {
    if (PAT == GMP_PAT_SPLITMIX) {
        w0 = mix64(seed + (L + 1) * kGamma);
        w1 = ~rotl32of64(w0) ^ L;
    } else if (PAT == GMP_PAT_ZERO) {
        w0 = w1 = 0;
    } else if (PAT == GMP_PAT_ONES) {
        w0 = w1 = ~0ull;
    } else if (PAT == GMP_PAT_CHECKER) {
        uint64_t v = (L & 1) ? 0x5555555555555555ull : 0xAAAAAAAAAAAAAAAAull;
        if (seed & 1) v = ~v;
        w0 = w1 = v;
    } else {  // GMP_PAT_ADDRESS
        w0 = seed + 16 * L;
        w1 = w0 + 8;
    }
}

void host_lane_words(int pattern, uint64_t seed, uint64_t L, uint64_t& w0, uint64_t& w1);

// ---- kernel launchers (gmp_kernels.cu) ---------------------------------------
// All return cudaGetLastError() of the launch; none synchronise.
struct DeviceInfo {
    int ordinal;
    int sm_count;
};
cudaError_t current_device_info(DeviceInfo* info);

// cfg == nullptr: the process-wide tunables for local HBM.
struct LaunchCfg {
    int variant;
    int ctas_per_sm;
};
cudaError_t launch_write(void* dptr, size_t bytes, int pattern, uint64_t seed, cudaStream_t s,
                         const LaunchCfg* cfg = nullptr);
// d_acc: 4 x u64 of device memory, zeroed by this call before the kernel.
cudaError_t launch_verify(const void* dptr, size_t bytes, int pattern, uint64_t seed,
                          unsigned long long* d_acc, cudaStream_t s, const LaunchCfg* cfg = nullptr);
cudaError_t launch_copy(void* dst, const void* src, size_t bytes, cudaStream_t s, const LaunchCfg* cfg = nullptr);
// One launch that writes SPLITMIX slices of slice_bytes (a multiple of 32) into npeers destinations at once, tile by
// tile round-robin over the destinations; dsts[p] gets the pattern of seeds[p].  The all-to-all round of the P2P probe.
cudaError_t launch_alltoall_write(void* const* dsts, const uint64_t* seeds, int npeers, size_t slice_bytes, cudaStream_t s);
// Geometry for kernels whose pointer is PEER memory (NVLink): picked by tools/sweep_p2p.py, held in tunables().
LaunchCfg peer_store_cfg();
LaunchCfg peer_load_cfg();
void decode_acc(const unsigned long long raw[4], gmp_verify_result* out);

// Launch geometry knobs (benchmarking only; see gmp_tune()).
struct Tunables {
    int write_ctas_per_sm;
    int read_ctas_per_sm;
    int copy_ctas_per_sm;
    int write_variant;
    int read_variant;
    int copy_variant;
    int peer_store_variant;
    int peer_store_ctas_per_sm;
    int peer_load_variant;
    int peer_load_ctas_per_sm;
};
Tunables& tunables();

// ---- NVML helpers (gmp_nvml.cpp) ------------------------------------------------
// READ/WRITE/NVLINK/ATOMICS status matrices (n x n each) and active NVLink counts in one NVML session.
int nvml_p2p_everything(const char* const* uuids, int n, int32_t* caps_out[4], int32_t* nvlink_active);

// NVML's (minor, index, bus id) of the devices named by devs[i].uuid, looked up one by one; entries NVML cannot
// answer for get minor = UINT32_MAX.  GMP_OK when all n were found.
int nvml_lookup_by_uuid(gmp_device* devs, int n, int* n_found);

// Drops per-device state cached across calls (what GMP_HBM_F_RETAIN calls left behind); call after cudaDeviceReset().
void forget_device_state(int ordinal);

// ---- misc --------------------------------------------------------------------
void format_uuid(const unsigned char bytes[16], char out[GMP_UUID_LEN]);
// GMP_OK and *ordinal, or GMP_ERR_UNAVAILABLE / GMP_ERR_NO_DEVICE.
int find_cuda_device(const char* uuid, int* ordinal, int* device_count);

}  // namespace gmp
