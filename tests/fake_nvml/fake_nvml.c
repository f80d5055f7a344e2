/*
 * A fake libnvidia-ml.so.1 for GPU-less tests (test infrastructure, not product code).
 *
 * Implements exactly the NVML entry points gpumounter_b200/csrc/gmp_nvml.cpp resolves, answering from a small
 * text file named by $GMP_FAKE_NVML_CONFIG, re-read at every nvmlInit_v2():
 *
 *   gpu <uuid> <minor> <pci bus id>            one line per device, NVML index = line order
 *   proc <uuid> <pid> <bytes> compute|graphics a process running on that device
 *   p2p <i> <j> <caps index> <status>          nvmlDeviceGetP2PStatus(i, j, caps) (default: 0 = OK)
 *   nvlink <uuid> <active links>               links 0..n-1 report ENABLED
 *   fail <function name> <nvmlReturn_t>        make one entry point fail (fault injection)
 *   driver <version> / nvml <version>
 */
#include <nvml.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define MAX_GPUS 16
#define MAX_PROCS 256
#define MAX_RULES 512

struct fake_gpu { char uuid[96]; unsigned minor; char bus[32]; unsigned nvlinks; };
struct fake_proc { int gpu; unsigned pid; unsigned long long mem; int graphics; };
struct fake_p2p { int i, j, caps, status; };
struct fake_fail { char fn[64]; int code; };

static struct fake_gpu g_gpus[MAX_GPUS];
static int g_ngpus;
static struct fake_proc g_procs[MAX_PROCS];
static int g_nprocs;
static struct fake_p2p g_p2p[MAX_RULES];
static int g_np2p;
static struct fake_fail g_fail[64];
static int g_nfail;
static char g_driver[80] = "580.00.fake";
static char g_nvml[80] = "12.580.fake";
static int g_inited;

static int gpu_index(const char* uuid) {
    for (int i = 0; i < g_ngpus; ++i)
        if (strcmp(g_gpus[i].uuid, uuid) == 0) return i;
    return -1;
}

static nvmlReturn_t should_fail(const char* fn) {
    for (int i = 0; i < g_nfail; ++i)
        if (strcmp(g_fail[i].fn, fn) == 0) return (nvmlReturn_t)g_fail[i].code;
    return NVML_SUCCESS;
}

#define MAYBE_FAIL(fn)                          \
    do {                                        \
        nvmlReturn_t f__ = should_fail(fn);     \
        if (f__ != NVML_SUCCESS) return f__;    \
    } while (0)

static void load_config(void) {
    g_ngpus = g_nprocs = g_np2p = g_nfail = 0;
    const char* path = getenv("GMP_FAKE_NVML_CONFIG");
    if (!path) return;
    FILE* f = fopen(path, "r");
    if (!f) return;
    char line[512];
    while (fgets(line, sizeof(line), f)) {
        char kind[32];
        if (sscanf(line, "%31s", kind) != 1 || kind[0] == '#') continue;
        if (strcmp(kind, "gpu") == 0 && g_ngpus < MAX_GPUS) {
            struct fake_gpu* g = &g_gpus[g_ngpus];
            memset(g, 0, sizeof(*g));
            if (sscanf(line, "%*s %95s %u %31s", g->uuid, &g->minor, g->bus) >= 2) ++g_ngpus;
        } else if (strcmp(kind, "proc") == 0 && g_nprocs < MAX_PROCS) {
            char uuid[96], type[32];
            struct fake_proc* p = &g_procs[g_nprocs];
            if (sscanf(line, "%*s %95s %u %llu %31s", uuid, &p->pid, &p->mem, type) == 4) {
                p->gpu = gpu_index(uuid);
                p->graphics = strcmp(type, "graphics") == 0;
                if (p->gpu >= 0) ++g_nprocs;
            }
        } else if (strcmp(kind, "p2p") == 0 && g_np2p < MAX_RULES) {
            struct fake_p2p* r = &g_p2p[g_np2p];
            if (sscanf(line, "%*s %d %d %d %d", &r->i, &r->j, &r->caps, &r->status) == 4) ++g_np2p;
        } else if (strcmp(kind, "nvlink") == 0) {
            char uuid[96];
            unsigned links;
            if (sscanf(line, "%*s %95s %u", uuid, &links) == 2 && gpu_index(uuid) >= 0) g_gpus[gpu_index(uuid)].nvlinks = links;
        } else if (strcmp(kind, "fail") == 0 && g_nfail < 64) {
            if (sscanf(line, "%*s %63s %d", g_fail[g_nfail].fn, &g_fail[g_nfail].code) == 2) ++g_nfail;
        } else if (strcmp(kind, "driver") == 0) {
            sscanf(line, "%*s %79s", g_driver);
        } else if (strcmp(kind, "nvml") == 0) {
            sscanf(line, "%*s %79s", g_nvml);
        }
    }
    fclose(f);
}

/* device handles are 1-based indices disguised as pointers */
static nvmlDevice_t handle_of(int idx) { return (nvmlDevice_t)(size_t)(idx + 1); }
static int index_of(nvmlDevice_t d) {
    int i = (int)(size_t)d - 1;
    return (i >= 0 && i < g_ngpus) ? i : -1;
}

/* Reference-counted like the real library: only the 0 -> 1 init is a cold one (driver attach, tens of ms on a real
 * box) and only the 1 -> 0 shutdown detaches.  fake_nvml_cold_inits() lets a test count the cold ones. */
static unsigned g_cold_inits;
unsigned fake_nvml_cold_inits(void) { return g_cold_inits; }

nvmlReturn_t nvmlInit_v2(void) {
    load_config();
    MAYBE_FAIL("nvmlInit_v2");
    if (g_inited++ == 0) ++g_cold_inits;
    return NVML_SUCCESS;
}

nvmlReturn_t nvmlShutdown(void) {
    if (!g_inited) return NVML_ERROR_UNINITIALIZED;
    --g_inited;
    return NVML_SUCCESS;
}

const char* nvmlErrorString(nvmlReturn_t r) {
    switch (r) {
        case NVML_SUCCESS: return "Success";
        case NVML_ERROR_UNINITIALIZED: return "Uninitialized";
        case NVML_ERROR_INVALID_ARGUMENT: return "Invalid Argument";
        case NVML_ERROR_NOT_SUPPORTED: return "Not Supported";
        case NVML_ERROR_NO_PERMISSION: return "Insufficient Permissions";
        case NVML_ERROR_NOT_FOUND: return "Not Found";
        case NVML_ERROR_INSUFFICIENT_SIZE: return "Insufficient Size";
        case NVML_ERROR_DRIVER_NOT_LOADED: return "Driver Not Loaded";
        case NVML_ERROR_GPU_IS_LOST: return "GPU is lost";
        default: return "Unknown Error";
    }
}

nvmlReturn_t nvmlDeviceGetCount_v2(unsigned int* n) {
    if (!g_inited) return NVML_ERROR_UNINITIALIZED;
    MAYBE_FAIL("nvmlDeviceGetCount_v2");
    *n = (unsigned)g_ngpus;
    return NVML_SUCCESS;
}

nvmlReturn_t nvmlDeviceGetHandleByIndex_v2(unsigned int i, nvmlDevice_t* d) {
    if (!g_inited) return NVML_ERROR_UNINITIALIZED;
    MAYBE_FAIL("nvmlDeviceGetHandleByIndex_v2");
    if ((int)i >= g_ngpus) return NVML_ERROR_INVALID_ARGUMENT;
    *d = handle_of((int)i);
    return NVML_SUCCESS;
}

nvmlReturn_t nvmlDeviceGetHandleByUUID(const char* uuid, nvmlDevice_t* d) {
    if (!g_inited) return NVML_ERROR_UNINITIALIZED;
    MAYBE_FAIL("nvmlDeviceGetHandleByUUID");
    int i = gpu_index(uuid);
    if (i < 0) return NVML_ERROR_NOT_FOUND;
    *d = handle_of(i);
    return NVML_SUCCESS;
}

nvmlReturn_t nvmlDeviceGetMinorNumber(nvmlDevice_t d, unsigned int* minor) {
    MAYBE_FAIL("nvmlDeviceGetMinorNumber");
    int i = index_of(d);
    if (i < 0) return NVML_ERROR_INVALID_ARGUMENT;
    *minor = g_gpus[i].minor;
    return NVML_SUCCESS;
}

nvmlReturn_t nvmlDeviceGetUUID(nvmlDevice_t d, char* uuid, unsigned int len) {
    MAYBE_FAIL("nvmlDeviceGetUUID");
    int i = index_of(d);
    if (i < 0) return NVML_ERROR_INVALID_ARGUMENT;
    if (strlen(g_gpus[i].uuid) + 1 > len) return NVML_ERROR_INSUFFICIENT_SIZE;
    strcpy(uuid, g_gpus[i].uuid);
    return NVML_SUCCESS;
}

nvmlReturn_t nvmlDeviceGetPciInfo_v3(nvmlDevice_t d, nvmlPciInfo_t* pci) {
    MAYBE_FAIL("nvmlDeviceGetPciInfo_v3");
    int i = index_of(d);
    if (i < 0) return NVML_ERROR_INVALID_ARGUMENT;
    memset(pci, 0, sizeof(*pci));
    snprintf(pci->busId, sizeof(pci->busId), "%s", g_gpus[i].bus);
    return NVML_SUCCESS;
}

nvmlReturn_t nvmlDeviceGetP2PStatus(nvmlDevice_t a, nvmlDevice_t b, nvmlGpuP2PCapsIndex_t caps, nvmlGpuP2PStatus_t* st) {
    MAYBE_FAIL("nvmlDeviceGetP2PStatus");
    int i = index_of(a), j = index_of(b);
    if (i < 0 || j < 0) return NVML_ERROR_INVALID_ARGUMENT;
    *st = NVML_P2P_STATUS_OK;
    for (int r = 0; r < g_np2p; ++r)
        if (g_p2p[r].i == i && g_p2p[r].j == j && g_p2p[r].caps == (int)caps) *st = (nvmlGpuP2PStatus_t)g_p2p[r].status;
    return NVML_SUCCESS;
}

nvmlReturn_t nvmlDeviceGetNvLinkState(nvmlDevice_t d, unsigned int link, nvmlEnableState_t* st) {
    MAYBE_FAIL("nvmlDeviceGetNvLinkState");
    int i = index_of(d);
    if (i < 0 || link >= NVML_NVLINK_MAX_LINKS) return NVML_ERROR_INVALID_ARGUMENT;
    *st = link < g_gpus[i].nvlinks ? NVML_FEATURE_ENABLED : NVML_FEATURE_DISABLED;
    return NVML_SUCCESS;
}

static nvmlReturn_t list_procs(nvmlDevice_t d, unsigned int* count, nvmlProcessInfo_t* out, int graphics) {
    int i = index_of(d);
    if (i < 0) return NVML_ERROR_INVALID_ARGUMENT;
    unsigned n = 0;
    for (int p = 0; p < g_nprocs; ++p) {
        if (g_procs[p].gpu != i || g_procs[p].graphics != graphics) continue;
        if (n < *count) {
            memset(&out[n], 0, sizeof(out[n]));
            out[n].pid = g_procs[p].pid;
            out[n].usedGpuMemory = g_procs[p].mem;
        }
        ++n;
    }
    nvmlReturn_t r = n > *count ? NVML_ERROR_INSUFFICIENT_SIZE : NVML_SUCCESS;
    *count = n;
    return r;
}

nvmlReturn_t nvmlDeviceGetComputeRunningProcesses_v3(nvmlDevice_t d, unsigned int* count, nvmlProcessInfo_t* out) {
    MAYBE_FAIL("nvmlDeviceGetComputeRunningProcesses_v3");
    return list_procs(d, count, out, 0);
}

nvmlReturn_t nvmlDeviceGetGraphicsRunningProcesses_v3(nvmlDevice_t d, unsigned int* count, nvmlProcessInfo_t* out) {
    MAYBE_FAIL("nvmlDeviceGetGraphicsRunningProcesses_v3");
    return list_procs(d, count, out, 1);
}

nvmlReturn_t nvmlSystemGetDriverVersion(char* v, unsigned int len) {
    MAYBE_FAIL("nvmlSystemGetDriverVersion");
    snprintf(v, len, "%s", g_driver);
    return NVML_SUCCESS;
}

nvmlReturn_t nvmlSystemGetNVMLVersion(char* v, unsigned int len) {
    MAYBE_FAIL("nvmlSystemGetNVMLVersion");
    snprintf(v, len, "%s", g_nvml);
    return NVML_SUCCESS;
}

/* ---- health: fixed answers, overridable per device with "health <uuid> <sm> <mem> <temp> <power_mw> <reasons> <ecc_unc>" */
struct fake_health { int gpu; unsigned sm, mem, temp, power; unsigned long long reasons, ecc_unc; };
static struct fake_health g_health[MAX_GPUS];
static int g_nhealth;

static const struct fake_health* health_of(int i) {
    static const struct fake_health dflt = {-1, 1965, 3996, 41, 250000, 0, 0};
    const char* path = getenv("GMP_FAKE_NVML_CONFIG");
    g_nhealth = 0;
    if (path) {
        FILE* f = fopen(path, "r");
        char line[512];
        while (f && fgets(line, sizeof(line), f)) {
            char uuid[96];
            struct fake_health h;
            if (strncmp(line, "health ", 7) == 0 && g_nhealth < MAX_GPUS &&
                sscanf(line, "%*s %95s %u %u %u %u %llu %llu", uuid, &h.sm, &h.mem, &h.temp, &h.power, &h.reasons, &h.ecc_unc) == 7) {
                h.gpu = gpu_index(uuid);
                g_health[g_nhealth++] = h;
            }
        }
        if (f) fclose(f);
    }
    for (int k = 0; k < g_nhealth; ++k)
        if (g_health[k].gpu == i) return &g_health[k];
    return &dflt;
}

/* "remap <uuid> <correctable rows> <uncorrectable rows> <pending 0|1> <failure 0|1>"; default: nothing remapped */
nvmlReturn_t nvmlDeviceGetRemappedRows(nvmlDevice_t d, unsigned int* corr, unsigned int* unc, unsigned int* pending,
                                       unsigned int* failure) {
    MAYBE_FAIL("nvmlDeviceGetRemappedRows");
    int i = index_of(d);
    if (i < 0) return NVML_ERROR_INVALID_ARGUMENT;
    *corr = *unc = *pending = *failure = 0;
    const char* path = getenv("GMP_FAKE_NVML_CONFIG");
    FILE* f = path ? fopen(path, "r") : NULL;
    char line[512];
    while (f && fgets(line, sizeof(line), f)) {
        char uuid[96];
        unsigned a, b, c, e;
        if (strncmp(line, "remap ", 6) == 0 && sscanf(line, "%*s %95s %u %u %u %u", uuid, &a, &b, &c, &e) == 5 && gpu_index(uuid) == i) {
            *corr = a;
            *unc = b;
            *pending = c;
            *failure = e;
        }
    }
    if (f) fclose(f);
    return NVML_SUCCESS;
}

/* "nvlinkerr <uuid> <crc flit> <crc data> <replay> <recovery>": totals, answered as unsigned long long; any other
 * field id is NOT_SUPPORTED in its own nvmlReturn, as the real call does per field */
nvmlReturn_t nvmlDeviceGetFieldValues(nvmlDevice_t d, int count, nvmlFieldValue_t* values) {
    MAYBE_FAIL("nvmlDeviceGetFieldValues");
    int i = index_of(d);
    if (i < 0 || count < 0 || values == NULL) return NVML_ERROR_INVALID_ARGUMENT;
    unsigned long long v[4] = {0, 0, 0, 0};
    const char* path = getenv("GMP_FAKE_NVML_CONFIG");
    FILE* f = path ? fopen(path, "r") : NULL;
    char line[512];
    while (f && fgets(line, sizeof(line), f)) {
        char uuid[96];
        unsigned long long a, b, c, e;
        if (strncmp(line, "nvlinkerr ", 10) == 0 && sscanf(line, "%*s %95s %llu %llu %llu %llu", uuid, &a, &b, &c, &e) == 5 &&
            gpu_index(uuid) == i) {
            v[0] = a; v[1] = b; v[2] = c; v[3] = e;
        }
    }
    if (f) fclose(f);
    for (int k = 0; k < count; ++k) {
        int slot = values[k].fieldId == NVML_FI_DEV_NVLINK_CRC_FLIT_ERROR_COUNT_TOTAL ? 0
                 : values[k].fieldId == NVML_FI_DEV_NVLINK_CRC_DATA_ERROR_COUNT_TOTAL ? 1
                 : values[k].fieldId == NVML_FI_DEV_NVLINK_REPLAY_ERROR_COUNT_TOTAL ? 2
                 : values[k].fieldId == NVML_FI_DEV_NVLINK_RECOVERY_ERROR_COUNT_TOTAL ? 3 : -1;
        if (slot < 0) {
            values[k].nvmlReturn = NVML_ERROR_NOT_SUPPORTED;
            continue;
        }
        values[k].nvmlReturn = NVML_SUCCESS;
        values[k].valueType = NVML_VALUE_TYPE_UNSIGNED_LONG_LONG;
        values[k].value.ullVal = v[slot];
    }
    return NVML_SUCCESS;
}

nvmlReturn_t nvmlDeviceGetClockInfo(nvmlDevice_t d, nvmlClockType_t t, unsigned int* c) {
    MAYBE_FAIL("nvmlDeviceGetClockInfo");
    int i = index_of(d);
    if (i < 0) return NVML_ERROR_INVALID_ARGUMENT;
    *c = t == NVML_CLOCK_MEM ? health_of(i)->mem : health_of(i)->sm;
    return NVML_SUCCESS;
}

nvmlReturn_t nvmlDeviceGetMaxClockInfo(nvmlDevice_t d, nvmlClockType_t t, unsigned int* c) {
    MAYBE_FAIL("nvmlDeviceGetMaxClockInfo");
    if (index_of(d) < 0) return NVML_ERROR_INVALID_ARGUMENT;
    *c = t == NVML_CLOCK_MEM ? 3996 : 1965;
    return NVML_SUCCESS;
}

nvmlReturn_t nvmlDeviceGetTemperature(nvmlDevice_t d, nvmlTemperatureSensors_t s, unsigned int* t) {
    MAYBE_FAIL("nvmlDeviceGetTemperature");
    (void)s;
    int i = index_of(d);
    if (i < 0) return NVML_ERROR_INVALID_ARGUMENT;
    *t = health_of(i)->temp;
    return NVML_SUCCESS;
}

nvmlReturn_t nvmlDeviceGetPowerUsage(nvmlDevice_t d, unsigned int* mw) {
    MAYBE_FAIL("nvmlDeviceGetPowerUsage");
    int i = index_of(d);
    if (i < 0) return NVML_ERROR_INVALID_ARGUMENT;
    *mw = health_of(i)->power;
    return NVML_SUCCESS;
}

nvmlReturn_t nvmlDeviceGetCurrentClocksThrottleReasons(nvmlDevice_t d, unsigned long long* r) {
    MAYBE_FAIL("nvmlDeviceGetCurrentClocksThrottleReasons");
    int i = index_of(d);
    if (i < 0) return NVML_ERROR_INVALID_ARGUMENT;
    *r = health_of(i)->reasons;
    return NVML_SUCCESS;
}

nvmlReturn_t nvmlDeviceGetTotalEccErrors(nvmlDevice_t d, nvmlMemoryErrorType_t et, nvmlEccCounterType_t ct, unsigned long long* n) {
    MAYBE_FAIL("nvmlDeviceGetTotalEccErrors");
    (void)ct;
    int i = index_of(d);
    if (i < 0) return NVML_ERROR_INVALID_ARGUMENT;
    *n = et == NVML_MEMORY_ERROR_TYPE_UNCORRECTED ? health_of(i)->ecc_unc : 0;
    return NVML_SUCCESS;
}
