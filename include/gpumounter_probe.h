/*
 * gpumounter_probe.h — C ABI of libgpumounter_probe.so
 *
 * The post-mount GPU probe and the native node-local mount path for GPUMounter,
 * built for NVIDIA B200 (sm_100a).  Everything behind this header is C++/CUDA;
 * nothing in the signatures is a C++ or torch type, so the library binds from
 * Go (cgo), Python (ctypes) or C alike.  INTEGRATION.md shows the cgo stub and
 * the hook in the worker.
 *
 * "ref:" citations are paths inside the GPUMounter reference tree
 * (jason-gideon/GPUMounter) naming the interface an entry point replaces or
 * extends.  The reference has no GPU compute path; the probe entry points
 * (sections 3-5) are additive and hang off the worker's AddGPU handler right
 * after its mount loop (ref: pkg/server/gpu-mount/server.go:80-98).
 *
 * Conventions (same shape as the reference's cgo->NVML seam,
 * ref: pkg/util/gpu/collector/nvml/bindings.go:19-57, nvml.go:17-119):
 *   - every function returns GMP_OK (0) or a negative GMP_ERR_* code; nothing
 *     throws or aborts across the ABI; gmp_strerror() names a code and
 *     gmp_last_error() gives this thread's detail string;
 *   - all buffers are caller-owned and fixed-size; the library never hands out
 *     memory the caller has to free;
 *   - the library loads on a box with no NVIDIA driver (CUDA runtime is linked
 *     statically and opens libcuda.so.1 lazily; NVML is dlopen'ed at call time,
 *     ref: pkg/util/gpu/collector/nvml/nvml_dl.go:29-36).  On such a box every
 *     GPU entry point returns GMP_ERR_UNAVAILABLE — there is NO CPU fallback.
 *   - thread-safe for calls on distinct devices; calls that name the same
 *     device serialise on the device like any CUDA work.
 */
#ifndef GPUMOUNTER_PROBE_H_
#define GPUMOUNTER_PROBE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GMP_ABI_VERSION 1

/* The library is built with -fvisibility=hidden; only GMP_API symbols are exported. */
#if defined(__GNUC__)
#define GMP_API __attribute__((visibility("default")))
#else
#define GMP_API
#endif

/* ------------------------------------------------------------------------- */
/* 1. Status codes                                                            */
/* ------------------------------------------------------------------------- */
enum {
    GMP_OK                   = 0,
    GMP_ERR_INVALID          = -1,  /* bad argument (NULL, misaligned, unknown pattern, size over a stated limit) */
    GMP_ERR_UNAVAILABLE      = -2,  /* no CUDA driver / no usable GPU on this box */
    GMP_ERR_NO_DEVICE        = -3,  /* UUID / ordinal does not name a visible device */
    GMP_ERR_CUDA             = -4,  /* a CUDA call failed; see gmp_last_error() */
    GMP_ERR_NVML             = -5,  /* an NVML call failed; message is "nvml: <nvmlErrorString>" (ref: bindings.go:43-49) */
    GMP_ERR_NVML_NOT_FOUND   = -6,  /* "could not load NVML library" (ref: bindings.go:51-57) */
    GMP_ERR_VERIFY           = -7,  /* the probe ran but read back data that differs from the pattern */
    GMP_ERR_NOMEM            = -8,  /* device or host allocation failed */
    GMP_ERR_SYS              = -9,  /* a system call failed (errno in gmp_last_error()) */
    GMP_ERR_BUFFER_TOO_SMALL = -10, /* caller buffer/capacity too small; required size is reported where possible */
    GMP_ERR_DEVICE_SET       = -11, /* CUDA-visible (UUID) set differs from the expected / NVML set */
    GMP_ERR_P2P_MATRIX       = -12, /* cudaDeviceCanAccessPeer matrix differs from NVML's P2P status matrix */
    GMP_ERR_CGROUP_DRIVER    = -13, /* "unsupported cgroup driver: <x>" (ref: pkg/util/cgroup/cgroup.go:78-84) */
    GMP_ERR_NAME             = -14, /* invalid character in a cgroup name component (ref: cgroup.go:37-39, a Go panic there) */
    GMP_ERR_EMPTY_CGROUP     = -15, /* cgroup.procs has no complete line (ref: util.go:50 indexes pids[0] and panics) */
    GMP_ERR_TARGET           = -16, /* "Target must be specified" (ref: pkg/util/namespace/namespace.go:71-73) */
    GMP_ERR_BUSY             = -17, /* GPU has processes of the pod and force is off (ref: util.go:106-109) */
    GMP_ERR_NODE_CTL         = -18, /* /dev/nvidiactl missing or not openable from the probing context */
    GMP_ERR_NODE_UVM         = -19, /* /dev/nvidia-uvm missing or not openable */
    GMP_ERR_NODE_GPU         = -20, /* /dev/nvidiaN missing, or present with no device behind it (ENXIO / ENODEV) */
    GMP_ERR_NODE_DENIED      = -21, /* /dev/nvidiaN exists but open() is refused (device cgroup) */
    GMP_ERR_DRIVER_LIB       = -22, /* libcuda.so.1 not resolvable from the probing context */
    GMP_ERR_STILL_MOUNTED    = -23, /* after an unmount: the device node is still there (or still opens) in the pod */
    GMP_ERR_UNSUPPORTED      = -24, /* section 7: a cgroup-v2 device program that can be neither decoded nor re-loaded */
    GMP_ERR_SLOW             = -25, /* everything verified, but a measured rate is under the caller's threshold */
    GMP_ERR_UNHEALTHY        = -26, /* everything verified, but NVML reports uncorrected ECC errors or a hardware slowdown */
    GMP_ERR_TIMEOUT          = -27  /* the in-pod probe did not finish within gmp_mount_env.probe_timeout_ms and was killed */
};

GMP_API int         gmp_abi_version(void);
GMP_API const char* gmp_strerror(int code);
/* Detail text of the last failure on the calling thread (never NULL). */
GMP_API const char* gmp_last_error(void);

/* ------------------------------------------------------------------------- */
/* 2. Device enumeration                                                      */
/* ------------------------------------------------------------------------- */
#define GMP_UUID_LEN    96  /* = NVML_DEVICE_UUID_V2_BUFFER_SIZE, ref: bindings.go:38 */
#define GMP_PCI_LEN     32
#define GMP_MAX_DEVICES 16

typedef struct gmp_device {
    char     uuid[GMP_UUID_LEN];    /* exactly the string nvmlDeviceGetUUID returns, "GPU-xxxxxxxx-xxxx-..." */
    uint32_t minor;                 /* /dev/nvidia<minor>; UINT32_MAX when not known (CUDA-side enumeration) */
    char     pci_bus_id[GMP_PCI_LEN];
    int32_t  nvml_index;            /* -1 when not known */
    int32_t  cuda_ordinal;          /* -1 when not known */
} gmp_device;

/*
 * NVML enumeration in NVML index order: for i in [0,count) handle-by-index ->
 * minor number -> UUID.  Replaces GPUCollector.GetGPUInfo
 * (ref: pkg/util/gpu/collector/collector.go:40-79) and the cgo wrappers it
 * calls (ref: pkg/util/gpu/collector/nvml/nvml.go:17-31,83-110).  The NVML
 * library name is "libnvidia-ml.so.1" (ref: nvml_dl.go:30) unless the
 * environment variable GPUMOUNTER_NVML_LIB names another file (used by the
 * tests to load a fake).  NVML is initialised and shut down inside the call,
 * as the reference does (ref: collector.go:43-47).
 */
GMP_API int gmp_enumerate_nvml(gmp_device* out, int cap, int* n);

/* The devices CUDA sees in this process, ordinal order, UUID in NVML's format. */
GMP_API int gmp_enumerate_cuda(gmp_device* out, int cap, int* n);

/*
 * (UUID, minor) set equality between two enumerations, never by index
 * (CUDA ordinals and NVML indices differ).  minor is compared only where both
 * sides know it.  GMP_OK or GMP_ERR_DEVICE_SET.
 */
GMP_API int gmp_device_sets_equal(const gmp_device* a, int na, const gmp_device* b, int nb);

/* Driver and NVML version strings (ref: pkg/util/gpu/collector/nvml/nvml.go:83-100; 80-byte buffers there). */
GMP_API int gmp_nvml_versions(char* driver, size_t driver_cap, char* nvml, size_t nvml_cap);

/*
 * Keeps NVML attached between calls.  Every NVML-backed entry point brackets itself with nvmlInit_v2 / nvmlShutdown, as
 * the reference does per query (ref: pkg/device/nvidia.go:58-63); NVML counts those references, and a cold attach costs
 * 30-140 ms.  A caller about to make several such calls (RemoveGPU's busy check asks once per GPU, ref:
 * server.go:139-153, then again inside UnmountGPU, util.go:100-109) takes one reference first - acquire != 0 - and gives
 * it back - acquire == 0 - when done; the calls in between then cost microseconds.  No reference counterpart.
 * GMP_ERR_NVML_NOT_FOUND / GMP_ERR_NVML when NVML cannot be attached, GMP_ERR_INVALID for an unmatched release.
 */
GMP_API int gmp_nvml_hold(int acquire);

typedef struct gmp_process_info {
    uint32_t pid;              /* host PID (the worker runs with hostPID) */
    uint32_t reserved;
    uint64_t used_gpu_memory;  /* ref: nvml.go:12-15 ProcessInfo{Pid, UsedGPUMemory} */
} gmp_process_info;

/*
 * Processes running on one GPU: NVML init -> handle by UUID -> graphics
 * processes, then compute processes, appended in that order -> shutdown.
 * Replaces NvidiaGPU.GetRunningProcess (ref: pkg/device/nvidia.go:58-87) and
 * the two cgo wrappers (ref: nvml.go:33-73).  *n receives the total found;
 * GMP_ERR_BUFFER_TOO_SMALL when it exceeds cap.
 */
GMP_API int gmp_gpu_running_processes(const char* uuid, gmp_process_info* out, int cap, int* n);

/*
 * NVML's view of peer-to-peer capability between the n GPUs named by UUID:
 * status[i*n+j] = nvmlDeviceGetP2PStatus(i, j, caps_index) for i != j, 0 (OK)
 * on the diagonal.  caps_index: 0 READ, 1 WRITE, 2 NVLINK, 3 ATOMICS, 4 PCI;
 * status 0 = NVML_P2P_STATUS_OK.  No reference counterpart (its vendored
 * nvml.h declares the call, ref: lib/nvml.h:2663, but nothing uses it).
 */
GMP_API int gmp_nvml_p2p_status(const char* const* uuids, int n, int caps_index, int32_t* status);

/*
 * Clocks, throttle reasons, temperature, power and ECC counters of one GPU as NVML reports them at the moment of the
 * call (every bandwidth figure is recorded with the clocks it was taken at).  No reference counterpart.
 * A field NVML cannot answer is left at UINT32_MAX / UINT64_MAX.  throttle_reasons is NVML's clocks-event bit mask
 * (0x8 hw_slowdown, 0x40 hw_thermal, 0x20 sw_thermal, 0x4 sw_power_cap, 0x1 gpu_idle).
 */
typedef struct gmp_gpu_health {
    uint32_t sm_mhz, mem_mhz, sm_max_mhz, mem_max_mhz;
    uint32_t temperature_c;
    uint32_t power_mw;
    uint64_t throttle_reasons;
    uint64_t ecc_uncorrected_volatile;
    uint64_t ecc_corrected_volatile;
    /* HBM row remapping (nvmlDeviceGetRemappedRows): rows replaced after correctable / uncorrectable errors, whether a
     * remap is still waiting for the next GPU reset, and whether remapping has FAILED (no spare rows left in a bank) */
    uint32_t remapped_rows_correctable, remapped_rows_uncorrectable;
    uint32_t remap_pending, remap_failure;
    /* NVLink error counters, totals over all links since the driver was loaded (nvmlDeviceGetFieldValues,
     * NVML_FI_DEV_NVLINK_*_ERROR_COUNT_TOTAL): flow-control CRC, data CRC, replays, recoveries.  Reported, not judged:
     * what matters is whether they MOVE while the P2P probe runs, which two consecutive reports show. */
    uint64_t nvlink_crc_flit_errors, nvlink_crc_data_errors, nvlink_replay_errors, nvlink_recovery_errors;
} gmp_gpu_health;

GMP_API int gmp_nvml_health(const char* uuid, gmp_gpu_health* out);

/* Number of NVLink links NVML reports enabled on the GPU (18 on an HGX B200). */
GMP_API int gmp_nvml_nvlink_active(const char* uuid, int* active_links, int* max_links);

/* ------------------------------------------------------------------------- */
/* 3. HBM probe kernels on caller-owned DEVICE memory (current CUDA device)   */
/* ------------------------------------------------------------------------- */
/*
 * No reference counterpart: the reference never touches device memory
 * (SURVEY.md 2.3).  These are the sm_100a kernels of the post-mount probe.
 *
 * Pattern "v1", defined on 16-byte lanes L = 0,1,2,... of the buffer, two
 * little-endian u64 words per lane (w0 at byte 16L, w1 at byte 16L+8):
 *   GMP_PAT_SPLITMIX : w0 = the L-th output of the canonical SplitMix64 stream
 *                      seeded with `seed` = mix(seed + (L+1)*0x9E3779B97F4A7C15);
 *                      w1 = ~rotl64(w0,32) ^ L
 *   GMP_PAT_ZERO     : w0 = w1 = 0
 *   GMP_PAT_ONES     : w0 = w1 = ~0
 *   GMP_PAT_CHECKER  : even L: 0xAAAA...AA, odd L: 0x5555...55; inverted when seed&1
 *   GMP_PAT_ADDRESS  : w0 = seed + 16L, w1 = seed + 16L + 8 (address-in-address;
 *                      pass the buffer's device address as seed)
 * A buffer whose size is not a multiple of 16 ends in a ragged lane holding
 * the leading bytes of that lane's pattern.  The base pointer must be 16-byte
 * aligned; bytes == 0 is a successful no-op.
 *
 * `stream` is a cudaStream_t passed as void* (NULL = the legacy default
 * stream).  write/copy are asynchronous on that stream; read_verify returns
 * after its result has reached the host struct.
 */
enum {
    GMP_PAT_SPLITMIX = 0,
    GMP_PAT_ZERO     = 1,
    GMP_PAT_ONES     = 2,
    GMP_PAT_CHECKER  = 3,
    GMP_PAT_ADDRESS  = 4,
    GMP_PAT_COUNT    = 5
};

typedef struct gmp_verify_result {
    uint64_t mismatch_lanes;  /* 16-byte lanes (ragged last lane included) with any differing byte */
    uint64_t first_bad_lane;  /* smallest such lane index; UINT64_MAX when none */
    uint64_t xor64;           /* XOR of all little-endian u64 words READ (last word zero-padded) */
    uint64_t sum64;           /* their sum modulo 2^64 */
} gmp_verify_result;

GMP_API int gmp_hbm_write_pattern(void* dptr, size_t bytes, int pattern, uint64_t seed, void* stream);
GMP_API int gmp_hbm_read_verify(const void* dptr, size_t bytes, int pattern, uint64_t seed,
                        gmp_verify_result* out, void* stream);
GMP_API int gmp_hbm_copy(void* dst, const void* src, size_t bytes, void* stream);
/*
 * The same verify pass without the host round trip: `d_acc` is 32 bytes of
 * 8-byte-aligned DEVICE memory owned by the caller, cleared and then reduced
 * into on `stream`; gmp_hbm_verify_fetch() copies it back, waits for the
 * stream and decodes it.  Lets a caller queue write/verify/copy passes back to
 * back and read the verdicts once.
 */
GMP_API int gmp_hbm_read_verify_async(const void* dptr, size_t bytes, int pattern, uint64_t seed,
                              void* d_acc, void* stream);
GMP_API int gmp_hbm_verify_fetch(const void* d_acc, gmp_verify_result* out, void* stream);
/*
 * gmp_hbm_read_verify() for callers that must not allocate.  It has no counterpart in the reference
 * (nothing there touches device memory).
 *
 * Who needs it: a health-check sidecar on a GPU whose memory the tenant has already filled (cudaMalloc
 * can fail there, and a probe should not compete with the workload for memory); a monitor that must
 * not stall the tenant (the allocating entry point ends in cudaFree, which synchronises the whole
 * device); and a caller whose own rules forbid allocator calls at that point.  This call makes no
 * cudaMalloc / cudaFree / other allocating call: its scratch is a 65 568-byte array that is part of the
 * library image and sits in the memory of every GPU that uses the library, until the process exits.
 *
 * Same arguments, same result and the same bit-exact checksums as gmp_hbm_read_verify().  Limits and
 * rules that come with the fixed scratch:
 *   - bytes <= GMP_NOALLOC_MAX_BYTES (8 GiB: 2048 segments of 4 MiB, one scratch slot each); larger
 *     buffers are refused with GMP_ERR_INVALID and a message saying so — verify them in pieces;
 *   - every call on a device uses the same array, which still holds the previous call's data (the
 *     call clears the part that has to start at zero).  Two calls whose GPU work overlaps — different
 *     streams, different threads — race on it and return garbage: serialise them per device;
 *   - the FIRST call on a device is the one that costs something: it loads the library's device code
 *     there (the array comes with it) and makes all of this entry point's kernels resident, whatever
 *     pattern that first call uses.  Make it early — while the GPU still has room, and before starting
 *     any stream capture; whether later calls may run beside a capture is the caller's to establish.
 *     Every call looks the array's address up again (microseconds), so a cudaDeviceReset() made by the
 *     caller between two calls is harmless: the next call simply is a "first call" again.
 */
#define GMP_NOALLOC_MAX_BYTES (8ull << 30)
GMP_API int gmp_hbm_read_verify_noalloc(const void* dptr, size_t bytes, int pattern, uint64_t seed,
                                        gmp_verify_result* out, void* stream);
/* Fault injection for tests: XOR one byte of device memory with `mask`. */
GMP_API int gmp_hbm_corrupt(void* dptr, size_t bytes, size_t byte_offset, uint8_t mask, void* stream);
/*
 * Benchmarking knob, process-wide, not for production callers: selects the
 * launch geometry of the three kernels.  Knobs: "write_ctas_per_sm",
 * "read_ctas_per_sm", "copy_ctas_per_sm", "write_variant", "read_variant",
 * "copy_variant", and for pointers into peer memory "peer_store_variant",
 * "peer_store_ctas_per_sm", "peer_load_variant", "peer_load_ctas_per_sm".  A *_ctas_per_sm of 0 means one resident wave (persistent
 * grid sized by the occupancy API); N > 0 launches N x (SM count) CTAs.
 * value < 0 only queries.  Returns the value in force (>= 0) or
 * GMP_ERR_INVALID.
 */
GMP_API int gmp_tune(const char* knob, int value);
/* Number of write / verify / copy kernels this process has launched through the library so far (a copy whose size is
 * not a multiple of the TMA chunk is two launches).  Benchmarking aid: bench.py reports the difference over its timed
 * region as `gpu_launches`. */
GMP_API unsigned long long gmp_kernel_launches(void);

/* ------------------------------------------------------------------------- */
/* 4. One-shot HBM probe of one mounted GPU                                   */
/* ------------------------------------------------------------------------- */
/*
 * Called by the worker after util.MountGPU succeeded for the device
 * (ref: pkg/server/gpu-mount/server.go:80-98; the *NvidiaGPU it holds carries
 * the UUID, ref: pkg/device/nvidia.go:10-17).  Selects the device by UUID,
 * allocates two `bytes`-sized buffers, and per selected pattern runs
 * write -> read-verify (iters times), then one copy pass + verify of the
 * copy, then an optional host<->device DMA round trip, and frees everything.
 * Bandwidths use algorithmic bytes: write = bytes, verify = bytes,
 * copy = 2*bytes.
 */
#define GMP_HBM_DEFAULT_BYTES      (8ull << 30)
#define GMP_HBM_DEFAULT_DMA_BYTES  (16ull << 20)

typedef struct gmp_hbm_opts {
    uint64_t bytes;         /* per buffer; 0 -> GMP_HBM_DEFAULT_BYTES */
    uint32_t iters;         /* timed write+verify repetitions per pattern; 0 -> 3 */
    uint32_t pattern_mask;  /* bit k selects GMP_PAT_k; 0 -> SPLITMIX only */
    uint64_t seed;          /* 0 -> 0x9E3779B97F4A7C15; iteration i uses seed + i */
    uint64_t dma_bytes;     /* pinned host<->device round trip; 0 -> skip; UINT64_MAX -> default */
    uint32_t flags;         /* GMP_HBM_F_* */
    uint32_t reserved;
} gmp_hbm_opts;

enum {
    GMP_HBM_F_SKIP_COPY   = 1u << 0,
    GMP_HBM_F_KEEP_DEVICE = 1u << 1,  /* do not cudaDeviceReset() at the end (in-process callers that share the context) */
    GMP_HBM_F_LIB_BASELINES = 1u << 2, /* also time cudaMemsetAsync and cudaMemcpyAsync D2D on the same buffers */
    /* Implies KEEP_DEVICE.  For a resident worker that probes the same GPU again and again (the churn of BASELINE
     * configs[4]): the call leaves its stream, events, accumulators, the registered host staging buffer AND the two
     * device buffers in place for the next call on that device, so that a call is the passes and one synchronisation
     * (cudaMalloc/cudaFree of 2 x 8 GiB and the pinning of the staging buffer are 30 of the 36 ms of a cold call, and
     * they serialise inside the driver when several GPUs are probed at once).  The memory stays allocated until
     * gmp_probe_hbm_release() or a call on that device without this flag: a worker must release before it hands the GPU
     * to the tenant. */
    GMP_HBM_F_RETAIN      = 1u << 3
};

typedef struct gmp_hbm_result {
    char     uuid[GMP_UUID_LEN];
    int32_t  cuda_ordinal;
    int32_t  status;             /* same value the call returned */
    uint64_t bytes;              /* per buffer, as probed */
    uint32_t iters;
    uint32_t patterns_run;       /* mask */
    double   write_gbs_best,  write_gbs_median;
    double   read_gbs_best,   read_gbs_median;
    double   copy_gbs_best,   copy_gbs_median;
    double   h2d_gbs, d2h_gbs;   /* DMA round trip; 0 when skipped */
    uint64_t mismatch_lanes;     /* summed over every verify pass, DMA included */
    uint64_t first_bad_lane;     /* of the first failing pass; UINT64_MAX when none */
    int32_t  first_bad_pattern;  /* -1 when none */
    uint32_t sm_count;
    uint64_t xor64, sum64;       /* of the last SPLITMIX verify pass (or last pass run) */
    uint64_t h2d_bytes, d2h_bytes; /* bytes actually moved over PCIe by the call */
    /* phase wall times, ns, CLOCK_MONOTONIC */
    uint64_t ns_select_device;   /* cudaGetDeviceCount + UUID match + cudaSetDevice */
    uint64_t ns_context;         /* primary-context creation (cudaFree(0)) */
    uint64_t ns_alloc;
    uint64_t ns_first_kernel;    /* untimed first touch: write+verify of a 4 MiB prefix, launch to verified result */
    uint64_t ns_kernels;         /* all timed passes: first launch to the passes' stream drained (ONE synchronisation) */
    uint64_t ns_dma;             /* what the DMA round trip added after that; its set-up and transfers run on a second
                                    stream beside the passes, so this is only the part they did not hide */
    uint64_t ns_teardown;
    uint64_t ns_total;
    uint64_t mono_first_verified_ns; /* absolute CLOCK_MONOTONIC time at which the first kernel result verified
                                        (comparable across processes of one host: the "usable" instant) */
    double   lib_memset_gbs;         /* same buffer through cudaMemsetAsync (bytes written / time), best of iters */
    double   lib_memcpy_gbs;         /* A -> B through cudaMemcpyAsync device-to-device (2N bytes / time), best of iters;
                                        both 0 unless GMP_HBM_F_LIB_BASELINES is set */
    uint64_t mono_context_begin_ns;  /* absolute CLOCK_MONOTONIC around primary-context creation on this GPU: with one */
    uint64_t mono_context_end_ns;    /* host thread per GPU (gmp_probe_all) the intervals show whether the driver overlaps them */
} gmp_hbm_result;

/* uuid == NULL or "" probes CUDA ordinal 0. */
GMP_API int gmp_probe_hbm(const char* uuid, const gmp_hbm_opts* opts, gmp_hbm_result* out);
/* Frees what GMP_HBM_F_RETAIN calls left on the device named by uuid (NULL or "": on every device).  GMP_OK also when
 * nothing was retained.  No reference counterpart. */
GMP_API int gmp_probe_hbm_release(const char* uuid);

/* ------------------------------------------------------------------------- */
/* 5. Peer-to-peer probe across the GPUs mounted into one pod                 */
/* ------------------------------------------------------------------------- */
/*
 * No reference counterpart.  For the n GPUs named by UUID (n = 1, 2, 4, 8):
 *   - can_access[i*n+j]   = cudaDeviceCanAccessPeer(i, j)          (1 on the diagonal)
 *   - nvml_*[i*n+j]       = nvmlDeviceGetP2PStatus(i, j, READ/WRITE/NVLINK/ATOMICS), 0 = OK
 *   - matrix_match        = for all i != j: can_access == (nvml_read == OK && nvml_write == OK)
 *   - write_gbs[i*n+j]    = a kernel on GPU i STORES the pattern into GPU j's HBM over NVLink
 *                           (16/32-byte peer stores), GPU j then verifies its own memory;
 *   - read_gbs[i*n+j]     = a kernel on GPU i LOADS and verifies GPU j's buffer over NVLink;
 *   - memcpy_gbs[i*n+j]   = cudaMemcpyPeerAsync from GPU i's buffer to GPU j's (copy engines), verified by j;
 *   - atomics_ok[i*n+j]   = native atomics over NVLink work from i on j's memory (what NCCL/NVSHMEM-style
 *                           workloads in the pod will rely on); a wrong sum is counted as a verification failure;
 *   - bidir_gbs[i]        = GPU i's traffic (sent + received bytes / time) while all GPUs exchange
 *                           with a partner at once, worst round of the n-1 round schedule;
 *   - alltoall_gbs        = aggregate when every GPU writes to every peer at once.
 * Bytes: one ordered pair moves `bytes` (algorithmic = wire bytes).
 * Every transfer is verified; mismatch_lanes counts 16-byte lanes that differ.
 * The call enables peer access itself and, unless GMP_P2P_F_KEEP_DEVICE is set,
 * ends with cudaDeviceReset() on every device it touched.
 */
#define GMP_P2P_DEFAULT_BYTES (1ull << 30)
#define GMP_P2P_CELLS (GMP_MAX_DEVICES * GMP_MAX_DEVICES)

typedef struct gmp_p2p_opts {
    uint64_t bytes;   /* per ordered pair; 0 -> GMP_P2P_DEFAULT_BYTES */
    uint32_t iters;   /* timed repetitions per measurement; 0 -> 2 */
    uint32_t flags;   /* GMP_P2P_F_* */
    uint64_t seed;    /* 0 -> default */
} gmp_p2p_opts;

enum {
    GMP_P2P_F_KEEP_DEVICE   = 1u << 0,
    GMP_P2P_F_SKIP_BIDIR    = 1u << 1,
    GMP_P2P_F_SKIP_ALLTOALL = 1u << 2,
    GMP_P2P_F_MATRIX_ONLY   = 1u << 3   /* capability matrices only, no transfers */
};

typedef struct gmp_p2p_result {
    int32_t  n;
    int32_t  status;
    char     uuid[GMP_MAX_DEVICES][GMP_UUID_LEN];
    int32_t  cuda_ordinal[GMP_MAX_DEVICES];
    int32_t  nvlink_active[GMP_MAX_DEVICES];   /* -1 when NVML is not available */
    int32_t  nvml_available;
    int32_t  matrix_match;                      /* 1 match, 0 mismatch, -1 not checkable (no NVML) */
    int32_t  can_access[GMP_P2P_CELLS];
    int32_t  nvml_read[GMP_P2P_CELLS];
    int32_t  nvml_write[GMP_P2P_CELLS];
    int32_t  nvml_nvlink[GMP_P2P_CELLS];
    int32_t  nvml_atomics[GMP_P2P_CELLS];
    double   write_gbs[GMP_P2P_CELLS];          /* best of iters; 0 on the diagonal / inaccessible pairs */
    double   read_gbs[GMP_P2P_CELLS];
    double   memcpy_gbs[GMP_P2P_CELLS];         /* cudaMemcpyPeerAsync i -> j (copy engines), the library baseline */
    int32_t  atomics_ok[GMP_P2P_CELLS];         /* 65 536 atomicAdd's from GPU i on a counter in GPU j's memory summed
                                                   exactly: 1 yes, 0 no, -1 not run (diagonal, inaccessible pair) */
    double   bidir_gbs[GMP_MAX_DEVICES];
    double   alltoall_gbs;
    uint64_t bytes;
    uint64_t mismatch_lanes;
    int32_t  first_bad_src, first_bad_dst;      /* ordered pair of the first failing transfer; -1 when none */
    uint64_t ns_setup, ns_pairs, ns_bidir, ns_alltoall, ns_teardown, ns_total;
} gmp_p2p_result;

GMP_API int gmp_probe_p2p(const char* const* uuids, int n, const gmp_p2p_opts* opts, gmp_p2p_result* out);

/* Maps (enable != 0) or unmaps CUDA device `peer`'s memory into the context of CUDA device `device`, so that
 * section-3 kernels launched on `device` may be given pointers that live on `peer`.  Already-enabled /
 * already-disabled is not an error.  For callers that drive the peer kernels themselves; gmp_probe_p2p
 * does this on its own. */
GMP_API int gmp_peer_access(int device, int peer, int enable);

/* ------------------------------------------------------------------------- */
/* 5b. Whole-pod probe and its JSON line                                       */
/* ------------------------------------------------------------------------- */
/*
 * What the hook runs after AddGPU's mount loop (ref: server.go:97-98), inside the
 * pod's device cgroup and mount namespace:
 *   1. device nodes: /dev/nvidiactl, /dev/nvidia-uvm and every expected /dev/nvidia<minor>
 *      exist and open (distinct GMP_ERR_NODE_* codes);
 *   2. device set: the CUDA-visible UUID set equals the expected set (GMP_ERR_DEVICE_SET);
 *   3. HBM probe on every GPU, one host thread per GPU (the reference's mount loop is serial);
 *   4. NVML cross-check, loaded on its own thread while CUDA starts: (UUID, minor) equals NVML's
 *      enumeration (ref: collector.go:56-77) when NVML is reachable from the pod;
 *   5. P2P probe across them (n >= 2).
 * expected may be NULL/0: probe everything CUDA sees.
 */
typedef struct gmp_probe_opts {
    gmp_hbm_opts hbm;
    gmp_p2p_opts p2p;
    uint32_t     flags;       /* GMP_PROBE_F_* */
    uint32_t     reserved;
    const char*  dev_dir;     /* where the device nodes are looked up; NULL -> "/dev" */
    /* GPUs the pod holds ALREADY (earlier AddGPU calls, or natively): they must be visible too - the CUDA-visible set
     * has to equal expected + also_visible exactly - but they are not probed: they may be running the tenant's work.
     * gpumounter-probe: --also-mounted UUID[:minor].  NULL / 0: the pod holds nothing but `expected`. */
    const gmp_device* also_visible;
    int32_t      n_also_visible;
    int32_t      reserved2;
} gmp_probe_opts;

enum {
    GMP_PROBE_F_SKIP_NODES = 1u << 0,
    GMP_PROBE_F_SKIP_HBM   = 1u << 1,
    GMP_PROBE_F_SKIP_P2P   = 1u << 2,
    GMP_PROBE_F_SKIP_NVML  = 1u << 3,
    GMP_PROBE_F_KEEP_DEVICE = 1u << 4,
    /* Stop after step 1 (device nodes + driver library): no CUDA, no NVML, a few ms.  What the worker runs INSIDE the
     * pod when the kernels are run from its own resident process instead (GMP_PROBE_F_HOST_VIEW). */
    GMP_PROBE_F_NODES_ONLY = 1u << 5,
    /* The caller is not confined to the pod's device set - the worker itself, on the host, where CUDA sees every GPU
     * of the node: each expected UUID must be visible to CUDA, but CUDA seeing MORE devices than expected is not an
     * error.  With GMP_PROBE_F_KEEP_DEVICE in a long-lived process, cuInit and the primary contexts are paid once. */
    GMP_PROBE_F_HOST_VIEW = 1u << 6
};

typedef struct gmp_report {
    int32_t        status;          /* first failing check, or GMP_OK */
    int32_t        n;
    int32_t        nodes_status;    /* GMP_OK or a GMP_ERR_NODE_* / GMP_ERR_DRIVER_LIB code */
    int32_t        device_set_status;
    int32_t        nvml_status;     /* GMP_OK, or why NVML could not be consulted */
    int32_t        reserved;
    gmp_device     devices[GMP_MAX_DEVICES];   /* CUDA view joined with NVML (minor, nvml_index) by UUID */
    gmp_hbm_result hbm[GMP_MAX_DEVICES];
    gmp_p2p_result p2p;
    gmp_gpu_health health[GMP_MAX_DEVICES];    /* sampled right after the HBM passes, while the GPUs are still warm */
    uint64_t       ns_nodes, ns_enumerate, ns_hbm, ns_p2p, ns_total;
    uint64_t       ns_cuda_init;    /* cuInit + device enumeration (inside ns_enumerate) */
    uint64_t       ns_nvml;         /* NVML load + init + enumeration, overlapped with CUDA start-up on its own thread */
    uint64_t       mono_start_ns;   /* absolute CLOCK_MONOTONIC at entry */
    uint64_t       mono_usable_ns;  /* ... when the LAST GPU's first kernel result verified; 0 if none did */
    char           detail[256];     /* gmp_last_error() of the first failing check */
} gmp_report;

GMP_API int gmp_probe_all(const gmp_device* expected, int n_expected, const gmp_probe_opts* opts, gmp_report* out);
/*
 * The RemoveGPU-side check (no reference counterpart; the reference returns Success as soon as `rm` exited 0,
 * ref: pkg/util/util.go:131-134, server.go:155-168): seen from the calling context — run it in the pod, e.g. through
 * gmp_probe_in_pod with "--gone <uuid>:<minor>" — none of the listed devices' nodes <dev_dir>/nvidia<minor> may exist any
 * more.  A node that is still present is GMP_ERR_STILL_MOUNTED, whether or not the device cgroup would still let it
 * open (the detail string says which).  dev_dir NULL = "/dev".
 */
GMP_API int gmp_verify_gone(const gmp_device* gone, int n, const char* dev_dir);
/*
 * Acceptance thresholds on a finished report (the operator's policy, e.g. "85 % of this node's usual HBM rate"): when the
 * report is otherwise GMP_OK and any GPU's best write, read-verify or copy rate is under min_hbm_gbs, or any measured
 * peer store / load rate is under min_p2p_gbs, status becomes GMP_ERR_SLOW and detail names the first offender.  A
 * threshold of 0 is not applied; rates that were not measured (0) are not judged.  Returns the report's status.
 * gpumounter-probe: --min-hbm-gbs N, --min-p2p-gbs N.  No reference counterpart.
 */
GMP_API int gmp_report_apply_thresholds(gmp_report* report, double min_hbm_gbs, double min_p2p_gbs);

/*
 * Health policy on a finished report: when the report is otherwise GMP_OK and a GPU's NVML sample (health[], taken
 * right after the passes) shows volatile UNCORRECTED ECC errors, or a throttle reason that means the hardware is
 * protecting itself - HW slowdown (0x8), HW thermal slowdown (0x40), HW power brake (0x80) - or an HBM row remap that
 * failed or is still pending (the GPU needs a reset before it is handed to a tenant), status becomes
 * GMP_ERR_UNHEALTHY and detail names the GPU and the reason.  Values NVML could not supply (all ones) are not judged.
 * gpumounter-probe: --strict-health.  No reference counterpart.
 */
GMP_API int gmp_report_check_health(gmp_report* report);

/* One JSON object (single line, no trailing newline) describing the report.  GMP_ERR_BUFFER_TOO_SMALL
 * when cap is not enough (32 KiB always is). */
GMP_API int gmp_report_json(const gmp_report* report, char* buf, size_t cap);

/* ------------------------------------------------------------------------- */
/* 6. Node-local mount path (native restatement of the worker's L2/L3 logic)  */
/* ------------------------------------------------------------------------- */
/*
 * Same functions, argument meaning and error behaviour as the reference's Go
 * packages, so the worker can call them through cgo one for one; what changes
 * is the mechanism: the reference shells out (`sh -c "echo ... > devices.allow"`,
 * `nsenter --target P --mount sh -c "mknod ..."`, 2-3 fork/execs per GPU and an
 * image that must contain sh/mknod/rm/kill, ref: docs/guide/FAQ.md:3-4); here
 * the permission line is written with write(2) and the device node is made by
 * a forked child that joins the target's mount namespace with setns(2) and
 * calls mknod(2)/unlink(2) itself — no shell, nothing needed inside the image.
 * Where Go would panic the functions return an error code instead.
 */

/* The device record, ref: pkg/device/nvidia.go:10-17 (same fields, same order). */
typedef struct gmp_gpu {
    int32_t minor_number;
    char    device_file_path[64];  /* "/dev/nvidia" + minor, ref: nvidia.go:29,40 */
    char    uuid[GMP_UUID_LEN];
    char    state[24];             /* "GPU_FREE_STATE" | "GPU_ALLOCATED_STATE", ref: nvidia.go:20-23 */
    char    pod_name[256];
    char    ns[256];
} gmp_gpu;

#define GMP_NVIDIA_MAJOR 195  /* ref: nvidia.go:37 */

GMP_API int gmp_device_new(int minor, const char* uuid, gmp_gpu* out);        /* ref: nvidia.go:25-34 */
GMP_API int gmp_device_string(const gmp_gpu* gpu, char* buf, size_t cap);     /* ref: nvidia.go:43-50, Go json.Marshal output */
GMP_API int gmp_device_reset_state(gmp_gpu* gpu);                             /* ref: nvidia.go:52-56 */

/* The fields of corev1.Pod the path reads.  Quantities are the strings of the
 * pod spec ("500m", "1Gi"); NULL = key absent.  Only cpu and memory take part
 * in QoS (ref: pkg/util/cgroup/cgroup.go:171-175). */
typedef struct gmp_container_resources {
    const char* cpu_request;
    const char* memory_request;
    const char* cpu_limit;
    const char* memory_limit;
} gmp_container_resources;

typedef struct gmp_pod {
    const char* name;
    const char* ns;
    const char* uid;
    const char* node_name;
    const char* container_id;  /* Status.ContainerStatuses[0].ContainerID, "docker://<id>" (ref: util.go:22-23) */
    int32_t     n_containers;
    const gmp_container_resources* containers;
} gmp_pod;

enum { GMP_QOS_GUARANTEED = 0, GMP_QOS_BURSTABLE = 1, GMP_QOS_BESTEFFORT = 2 };
enum { GMP_MOUNT_ENTIRE = 0, GMP_MOUNT_SINGLE = 1, GMP_MOUNT_NONE = 2, GMP_MOUNT_UNKNOWN = 3 }; /* ref: pkg/util/gpu/types.go:22-27 */

/* ref: cgroup.go:177-237 GetPodQOS */
GMP_API int gmp_get_pod_qos(const gmp_pod* pod, int* qos);
/* ref: cgroup.go:32-45 NewCgroupName's check; GMP_ERR_NAME where Go panics */
GMP_API int gmp_cgroup_name_check(const char* component);
/* ref: cgroup.go:52-68 CgroupName.ToSystemd (with runc's ExpandSlice, go.mod:8) */
GMP_API int gmp_cgroup_name_to_systemd(const char* const* parts, int n, char* out, size_t cap);
/* ref: cgroup.go:74-76 CgroupName.ToCgroupfs */
GMP_API int gmp_cgroup_name_to_cgroupfs(const char* const* parts, int n, char* out, size_t cap);
/* ref: cgroup.go:78-84 GetCgroupDriver; reads $CGROUP_DRIVER; GMP_ERR_CGROUP_DRIVER with
 * "unsupported cgroup driver: <value>" as the detail */
GMP_API int gmp_get_cgroup_driver(char* out, size_t cap);
/* ref: cgroup.go:86-113 GetCgroupName */
GMP_API int gmp_get_cgroup_name(const char* cgroup_driver, const gmp_pod* pod, const char* container_id,
                                char* out, size_t cap);

/* Where the node's files live.  NULL (or NULL members) = the reference's constants.
 * Tests and the cluster-less harness point these at private directories. */
typedef struct gmp_mount_env {
    const char* devices_root;   /* "/sys/fs/cgroup/devices", ref: cgroup.go:116 */
    const char* dev_prefix;     /* "/dev/nvidia", ref: nvidia.go:40 */
    const char* cgroup_driver;  /* overrides $CGROUP_DRIVER when not NULL */
    const char* proc_root;      /* "/proc": where <pid>/ns/mnt is looked up */
    const char* cgroup2_root;   /* "/sys/fs/cgroup": the unified hierarchy, used with GMP_EXT_CGROUP_V2 (section 7) */
    uint32_t    extensions;     /* GMP_EXT_* bits (section 7); 0 = exactly the reference's behaviour */
    uint32_t    probe_timeout_ms; /* gmp_probe_in_pod: kill the probe and return GMP_ERR_TIMEOUT after this long; 0 = wait
                                     for ever (a GPU that hangs must not hang the worker's AddGPU with it) */
} gmp_mount_env;

/* ref: util.go:21-34 (the first steps of MountGPU / UnmountGPU): container id of ContainerStatuses[0] with
 * "docker://" removed, $CGROUP_DRIVER, then GetCgroupName.  The one place where gmp_mount_env.extensions
 * (GMP_EXT_CRI_RUNTIMES, section 7) changes the answer. */
GMP_API int gmp_get_pod_cgroup_path(const gmp_mount_env* env, const gmp_pod* pod, char* out, size_t cap);
/* ref: cgroup.go:115-118 GetDeviceGroupPath */
GMP_API int gmp_get_device_group_path(const gmp_mount_env* env, const char* cgroup_path, char* out, size_t cap);
/* ref: cgroup.go:120-141 GetCgroupPIDs: the lines of cgroup.procs that end in '\n' (a last
 * line without newline is dropped, as there).  pids: caller array of char[16]; *n = lines found. */
GMP_API int gmp_get_cgroup_pids(const gmp_mount_env* env, const char* cgroup_path, char (*pids)[16], int cap, int* n);
/* ref: cgroup.go:143-155 / :157-169: writes "c 195:<minor> rw" to devices.allow / devices.deny */
GMP_API int gmp_add_gpu_device_permission(const gmp_mount_env* env, const char* cgroup_path, const gmp_gpu* gpu);
GMP_API int gmp_remove_gpu_device_permission(const gmp_mount_env* env, const char* cgroup_path, const gmp_gpu* gpu);

/* nsenter configuration, ref: pkg/util/namespace/namespace.go:16-39 (same fields). */
typedef struct gmp_ns_config {
    int32_t cgroup;  const char* cgroup_file;
    int32_t follow_context;
    int32_t gid;
    int32_t ipc;     const char* ipc_file;
    int32_t mount;   const char* mount_file;
    int32_t net;     const char* net_file;
    int32_t no_fork;
    int32_t pid;     const char* pid_file;
    int32_t preserve_credentials;
    const char* root_directory;
    int32_t target;
    int32_t uid;
    int32_t user;    const char* user_file;
    int32_t uts;     const char* uts_file;
    const char* working_directory;
} gmp_ns_config;

/* ref: namespace.go:70-165 buildCommand: the nsenter argument vector, NUL-separated in `out`
 * ("nsenter\0--target\0<pid>\0..."), *argc = number of strings.  GMP_ERR_TARGET when target == 0. */
GMP_API int gmp_ns_build_command(const gmp_ns_config* cfg, char* out, size_t cap, int* argc);
/* ref: namespace.go:42-68 Execute: runs `nsenter <flags> program args...` like the reference and
 * captures stdout / stderr.  Kept for arbitrary programs; the three calls below do not use it. */
GMP_API int gmp_ns_execute(const gmp_ns_config* cfg, const char* program, const char* const* args, int nargs,
                           char* out, size_t out_cap, char* err, size_t err_cap);
/* ref: namespace.go:167-177 / :179-189 / :191-201 — native: setns + mknod(2) / unlink(2); kill(2). */
GMP_API int gmp_ns_add_gpu_device_file(const gmp_mount_env* env, const gmp_ns_config* cfg, const gmp_gpu* gpu);
GMP_API int gmp_ns_remove_gpu_device_file(const gmp_mount_env* env, const gmp_ns_config* cfg, const gmp_gpu* gpu);
GMP_API int gmp_ns_kill_running_gpu_processes(const gmp_ns_config* cfg, const char* const* pids, int n);

/* ref: pkg/util/util.go:207-226 CanMount; returns 1 / 0 */
GMP_API int gmp_can_mount(int mount_type, int request_is_entire_mount);
/* ref: util.go:152-196 GetPodGPUProcesses: cgroup.procs ∩ NVML(graphics then compute) PIDs, in
 * NVML order.  *n = 0 is the reference's nil. */
GMP_API int gmp_get_pod_gpu_processes(const gmp_mount_env* env, const gmp_pod* pod, const gmp_gpu* gpu,
                                      char (*pids)[16], int cap, int* n);
/* ref: util.go:17-71 MountGPU and util.go:73-147 UnmountGPU (GMP_ERR_BUSY = the GPUBusy sentinel). */
GMP_API int gmp_mount_gpu(const gmp_mount_env* env, const gmp_pod* pod, const gmp_gpu* gpu);
GMP_API int gmp_unmount_gpu(const gmp_mount_env* env, const gmp_pod* pod, const gmp_gpu* gpu, int force);
/*
 * Batched forms (no reference counterpart): the loop of
 * ref: pkg/server/gpu-mount/server.go:80-95 / :155-168 done with one cgroup
 * path resolution, one cgroup.procs read and ONE namespace-entering child
 * for all n GPUs.  All-or-nothing is NOT promised, as in the reference: on an
 * error the GPUs already handled stay handled; *done counts them.
 */
GMP_API int gmp_mount_gpus(const gmp_mount_env* env, const gmp_pod* pod, const gmp_gpu* gpus, int n, int* done);
GMP_API int gmp_unmount_gpus(const gmp_mount_env* env, const gmp_pod* pod, const gmp_gpu* gpus, int n, int force,
                             int* done);

/*
 * Runs the standalone probe (gpumounter-probe, built next to the library) INSIDE the pod and returns its
 * JSON line.  No reference counterpart; hook point ref: pkg/server/gpu-mount/server.go:97-98, after the
 * mount loop.  A forked child opens `probe_exe` in the caller's namespace, joins the pod's device cgroup
 * (the reference's `nsenter --mount` never does, so a check run its way would be gated by the WORKER's
 * cgroup), setns()es into the pod's mount namespace and fexecve()s the binary with
 * "--uuid <uuid>:<minor>" for each GPU plus `extra_args`.  Returns the probe's verdict (GMP_OK or the
 * GMP_ERR_* it exited with), or GMP_ERR_SYS when the child could not be set up; *exit_status receives the
 * raw exit status.
 */
GMP_API int gmp_probe_in_pod(const gmp_mount_env* env, const gmp_pod* pod, const gmp_gpu* gpus, int n,
                             const char* probe_exe, const char* const* extra_args, int n_extra,
                             char* json, size_t cap, int* exit_status);

/*
 * The same check with ONE probe process per GPU, started together, each in a child of the pod's cgroup-v1 devices group
 * that may open only its own GPU (a child group can only narrow what the pod holds, so what verifies there is allowed
 * in the pod).  No reference counterpart.  Why: CUDA start-up, not the kernels, is the mount-to-usable latency, and it
 * does not scale inside one process - on an 8-GPU HGX B200 one process needs 1.3-1.5 s until all eight freshly mounted
 * GPUs have run a verified kernel (cuInit attaches every GPU it can open, then eight primary contexts contend in the
 * driver); eight processes that each attach one GPU need 0.86 s (profiles/r02_ctx_timeline_8gpu.json).  After the last of
 * them has finished, one more probe process runs in the pod's own group with --skip-hbm: it checks that CUDA in the
 * pod enumerates exactly the mounted set and runs the P2P step.  `json` receives
 *   {"status","verdict","detail","fanout":true,"n_gpus","mono_start_ns","mono_usable_ns","reports":[<one gpumounter-probe
 *    line per GPU, in the order of gpus>],"whole_pod":<that last process's line or null>}
 * where mono_usable_ns is the instant the LAST GPU's first kernel result verified (0 if one never did).  The child groups
 * are removed before the call returns.  A pod whose devices group holds a wildcard over the NVIDIA major (a privileged pod:
 * "a *:* rwm") cannot be narrowed: its per-GPU probes are then started with --host-view and each attaches every GPU the pod
 * sees (correct, but no faster than one process).  GMP_ERR_UNSUPPORTED with GMP_EXT_CGROUP_V2 (use gmp_probe_in_pod there).
 * gmp_mount_env.probe_timeout_ms bounds the per-GPU phase and the whole-pod phase separately.
 */
GMP_API int gmp_probe_in_pod_fanout(const gmp_mount_env* env, const gmp_pod* pod, const gmp_gpu* gpus, int n,
                                    const char* probe_exe, const char* const* extra_args, int n_extra,
                                    char* json, size_t cap, int* exit_status);

/*
 * The pod's view of its device nodes without running anything of the pod's image: a helper process joins the pod's
 * device cgroup and mount namespace and open(2)s /dev/nvidiactl, /dev/nvidia-uvm and the node of every GPU in `gpus`.
 * GMP_OK, or GMP_ERR_NODE_CTL / GMP_ERR_NODE_UVM / GMP_ERR_NODE_GPU (absent, or no device behind it) /
 * GMP_ERR_NODE_DENIED (the device cgroup refuses: EPERM) with *failed_index = the position in `gpus`, or GMP_ERR_SYS
 * when the namespace or cgroup could not be entered.  No reference counterpart; stage 1 of the resident hook
 *  for images in which gpumounter-probe cannot start (musl, distroless, scratch).
 */
GMP_API int gmp_check_nodes_in_pod(const gmp_mount_env* env, const gmp_pod* pod, const gmp_gpu* gpus, int n,
                                   int* failed_index);

/* The RemoveGPU-side counterpart: GMP_OK when none of the GPUs' nodes exists in the pod's mount namespace any more,
 * GMP_ERR_STILL_MOUNTED with *failed_index otherwise (the reference returns Success once `rm` exited 0, util.go:131-134).
 * Same helper, nothing executed in the pod; gmp_verify_gone is the check run from inside by gpumounter-probe --gone. */
GMP_API int gmp_verify_gone_in_pod(const gmp_mount_env* env, const gmp_pod* pod, const gmp_gpu* gpus, int n,
                                   int* failed_index);

/* kubelet PodResources listing, ref: collector.go:114-122 (v1alpha1 messages). */
typedef struct gmp_container_devices { const char* resource_name; int32_t n_ids; const char* const* device_ids; } gmp_container_devices;
typedef struct gmp_container_entry   { const char* name; int32_t n_devices; const gmp_container_devices* devices; } gmp_container_entry;
typedef struct gmp_pod_resources     { const char* name; const char* ns; int32_t n_containers; const gmp_container_entry* containers; } gmp_pod_resources;

/* ref: collector.go:81-88; *index = -1 and GMP_ERR_NO_DEVICE when absent */
GMP_API int gmp_get_gpu_by_uuid(const gmp_gpu* list, int n, const char* uuid, int* index);
/* ref: collector.go:113-135 (+ resetGPUStatus :140-144), applied to an already fetched listing */
GMP_API int gmp_update_gpu_status(gmp_gpu* list, int n, const gmp_pod_resources* pods, int n_pods);
/* ref: collector.go:155-161: the pod's own GPUs and those of its "<pod>-slave-pod-" pods in gpu-pool */
GMP_API int gmp_get_pod_gpu_resources(const gmp_gpu* list, int n, const char* pod_name, const char* ns,
                                      int32_t* indices, int* count);
/* ref: allocator.go:158-187 GetMountType on a refreshed list */
GMP_API int gmp_get_mount_type(const gmp_gpu* list, int n, const char* pod_name, const char* ns, int* mount_type);
/* ref: allocator.go:101-126 GetRemoveGPU */
GMP_API int gmp_get_remove_gpu(const gmp_gpu* list, int n, const char* owner_name, const char* owner_ns,
                               const char* const* uuids, int n_uuids, int32_t* indices, int* count);
/* ref: allocator.go:191-199: "<owner>-slave-pod-<6 hex>" from three random bytes */
GMP_API int gmp_slave_pod_name(const char* owner, const unsigned char rand3[3], char* out, size_t cap);

/* ---- 7. Extensions beyond the reference -----------------------------------
 *
 * Nothing in this section has a reference counterpart: the reference only knows dockershim container ids and the
 * cgroup-v1 `devices` controller (ref: pkg/util/cgroup/cgroup.go:104-108,115-118,143-169), so on a node with
 * containerd / CRI-O or the unified cgroup hierarchy its mount path resolves a directory that does not exist.
 * Everything here is OFF unless asked for through gmp_mount_env.extensions; with extensions == 0 the mount path
 * behaves exactly as the reference does (including for container ids it does not understand).
 */
enum {
    /* "containerd://<id>" and "cri-o://<id>" container ids: scheme stripped like "docker://"; systemd scope
     * "cri-containerd-<id>.scope" / "crio-<id>.scope", cgroupfs leaf "<id>" / "crio-<id>". */
    GMP_EXT_CRI_RUNTIMES = 1u << 0,
    /* Unified hierarchy: the pod's cgroup lives under gmp_mount_env.cgroup2_root ("/sys/fs/cgroup") and device
     * access is decided by BPF_PROG_TYPE_CGROUP_DEVICE programs, not devices.allow / devices.deny files. */
    GMP_EXT_CGROUP_V2    = 1u << 1,
    /* gmp_probe_in_pod: run the probe with the WORKER's credentials (root, every capability), as round 1 did.  Off by
     * default: the probe then runs as the user and group of the pod's first process, with an empty capability bounding,
     * permitted and effective set, no_new_privs set, and in the pod's net / ipc / uts namespaces as well as its mount
     * namespace - what it loads from the pod's root (ld.so, libc, libcuda.so.1) is the tenant's to choose, so it must
     * not run with more than the tenant already has. */
    GMP_EXT_PROBE_KEEP_PRIVILEGES = 1u << 2
};

/* Which hierarchy decides device access on this node: *version = 1 when <devices_root> (default
 * /sys/fs/cgroup/devices) is a directory - the v1 controller, which also wins on hybrid hosts - else 2 when
 * <cgroup2_root> (default /sys/fs/cgroup) is a cgroup2 file system, else GMP_ERR_UNAVAILABLE.  What a worker
 * consults once at start to decide whether to set GMP_EXT_CGROUP_V2. */
GMP_API int gmp_detect_cgroup_version(const gmp_mount_env* env, int* version);

/* One rule of a cgroup device filter; the values are the kernel's (BPF_DEVCG_DEV_*, BPF_DEVCG_ACC_*). */
enum { GMP_DEV_BLOCK = 1, GMP_DEV_CHAR = 2 };
enum { GMP_DEV_ACC_MKNOD = 1, GMP_DEV_ACC_READ = 2, GMP_DEV_ACC_WRITE = 4, GMP_DEV_ACC_ALL = 7 };
typedef struct gmp_dev_rule {
    int32_t  type;    /* GMP_DEV_CHAR or GMP_DEV_BLOCK */
    int32_t  major;   /* -1 = any */
    int32_t  minor;   /* -1 = any */
    uint32_t access;  /* GMP_DEV_ACC_* bits this rule covers; a request matches when all of ITS bits are covered */
    int32_t  allow;   /* 1 = allow, 0 = deny */
} gmp_dev_rule;

/* struct bpf_insn, spelled out so that callers need no kernel header: regs = dst_reg | src_reg << 4. */
typedef struct gmp_bpf_insn { uint8_t code; uint8_t regs; int16_t off; int32_t imm; } gmp_bpf_insn;

/*
 * Rules -> program, in the block form runc emits (opencontainers/runc libcontainer/cgroups/devices/devicefilter.go,
 * v1.0.0-rc93 and later; third-party, not in the reference tree): a 6-instruction prologue loads type, access,
 * major and minor into r2..r5; then one block per rule, LAST rule first (so the last matching rule of the list
 * wins), each block comparing only what the rule restricts and ending in "r0 = allow; exit"; finally
 * "r0 = default_allow; exit".  GMP_ERR_INVALID for a rule with an unknown type or no access bits,
 * GMP_ERR_BUFFER_TOO_SMALL with *n_insns = the size needed.
 */
GMP_API int gmp_devfilter_build(const gmp_dev_rule* rules, int n_rules, int default_allow,
                                gmp_bpf_insn* insns, int cap, int* n_insns);
/* Program -> rules (in list order, i.e. the reverse of block order).  GMP_ERR_UNSUPPORTED for anything that is
 * not exactly the block form above (for example a program written by crun). */
GMP_API int gmp_devfilter_decode(const gmp_bpf_insn* insns, int n_insns, gmp_dev_rule* rules, int cap,
                                 int* n_rules, int* default_allow);
/* Runs a device program on one request the way the kernel would (ALU, ALU64, JMP, JMP32, LDX from the 12-byte
 * context; no calls, no maps; at most 65536 steps).  *allowed = r0 != 0.  GMP_ERR_UNSUPPORTED for any other
 * instruction, GMP_ERR_INVALID for a jump or load out of range. */
GMP_API int gmp_devfilter_eval(const gmp_bpf_insn* insns, int n_insns, int type, uint32_t access,
                               uint32_t major, uint32_t minor, int* allowed);

/*
 * What gmp_cgroup2_update_devices would replace one attached program with, without touching the kernel.
 *  - a program in the block form above is edited rule list to rule list (see gmp_cgroup2_update_devices);
 *  - any other program (crun, systemd, hand-written) is kept verbatim behind a PREFIX: two marker instructions
 *    ("w0 = 0x474D5046; w0 = <n>"), the prologue, and one block per rule of ours (scratch register r6, so that r1
 *    still holds the context for the program behind).  A request that matches a prefix rule is decided there, every
 *    other request falls through into the original.  Revoking the last prefix rule gives the original back byte for
 *    byte.  Whether a device must be DENIED on revoke (it is reachable without our rule) is found by running the
 *    original with gmp_devfilter_eval;
 *  - a program that calls helpers or references maps cannot be read back in loadable form: GMP_ERR_UNSUPPORTED.
 */
GMP_API int gmp_devfilter_plan(const gmp_bpf_insn* insns, int n_insns, const gmp_dev_rule* rules, int n_rules, int grant,
                               gmp_bpf_insn* out, int cap, int* n_out);

/* The device programs attached to one cgroup-v2 directory (BPF_PROG_QUERY, not the effective set of its
 * ancestors).  *attach_flags receives the kernel's flags for that attachment (BPF_F_ALLOW_MULTI = 2 for runc). */
GMP_API int gmp_cgroup2_device_programs(const char* cgroup_dir, uint32_t* prog_ids, int cap, int* n,
                                        uint32_t* attach_flags);
/* The instructions of a loaded program, by id (BPF_OBJ_GET_INFO_BY_FD, translated form: identical to what was
 * loaded for programs of this type, which use no maps and no calls). */
GMP_API int gmp_cgroup2_read_program(uint32_t prog_id, gmp_bpf_insn* insns, int cap, int* n_insns);
/* Attach a program built from `rules` to a cgroup that has none (what a runtime does at container start; used by
 * the tests and the harness to play the runtime's part). */
GMP_API int gmp_cgroup2_attach_rules(const char* cgroup_dir, const gmp_dev_rule* rules, int n_rules, int default_allow);
/*
 * The v2 counterpart of writing devices.allow / devices.deny: every device program attached to `cgroup_dir` is
 * decoded, edited and atomically replaced (BPF_F_REPLACE).  grant != 0: append "allow" rules (earlier rules they
 * shadow are dropped).  grant == 0: remove exactly the allow rules a previous grant appended; where there is none
 * (the runtime itself had granted the device) append a deny rule instead.  A cgroup with no device program is
 * unrestricted at that level: nothing to do, *programs = 0.  Programs in other layouts are handled as gmp_devfilter_plan
 * describes.  Before a grant the programs ANCESTOR cgroups attached are run on the request: if one of them refuses the
 * device, no grant at this level can take effect and the call returns GMP_ERR_NODE_DENIED (what the v1 controller does
 * itself: writing a child's devices.allow fails with EPERM when the parent lacks the permission).  Any error leaves
 * every program of the cgroup untouched.
 */
GMP_API int gmp_cgroup2_update_devices(const char* cgroup_dir, const gmp_dev_rule* rules, int n_rules, int grant,
                                       int* programs);

#ifdef __cplusplus
}
#endif
#endif /* GPUMOUNTER_PROBE_H_ */
