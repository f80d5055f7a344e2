// NVML enumeration, per-GPU process list and P2P status through a dlopen'ed
// libnvidia-ml.so.1.  B200-era replacement for the reference's cgo binding
// (ref: pkg/util/gpu/collector/nvml/{bindings.go,nvml.go,nvml_dl.go}) and its
// callers GetGPUInfo (ref: pkg/util/gpu/collector/collector.go:40-79) and
// GetRunningProcess (ref: pkg/device/nvidia.go:58-87).

#include "gmp_nvml.h"

#include <dlfcn.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <exception>
#include <map>
#include <new>
#include <mutex>
#include <string>
#include <vector>

namespace gmp {
void set_error(const char* fmt, ...) __attribute__((format(printf, 1, 2)));
void clear_error();

// What a function-try-block of an extern "C" entry point returns: the exception in flight is named and swallowed,
// nothing crosses the C ABI.  Lives here because this translation unit is part of every build of the host code.
int caught_exception() noexcept {
    try {
        throw;
    } catch (const std::bad_alloc&) {
        set_error("out of memory");
        return GMP_ERR_NOMEM;
    } catch (const std::exception& e) {
        set_error("unexpected exception: %s", e.what());
        return GMP_ERR_SYS;
    } catch (...) {
        set_error("unexpected exception");
        return GMP_ERR_SYS;
    }
}

namespace {

// The library image stays loaded for the life of the process, one handle per name.  The reference
// dlopens and dlcloses around every query (ref: nvml_dl.go:29-53); with the 580 driver's libnvidia-ml
// each fresh image leaves two eventfds and two pipes behind at its first nvmlInit_v2 that neither
// nvmlShutdown nor dlclose gives back (tools/nvml_fd_leak.py, profiles/r01_nvml_fd_leak.json), so a
// worker that mounts and unmounts for weeks would run out of descriptors.  Loaded once, the cost is
// paid once; nvmlInit_v2 / nvmlShutdown still bracket every session, so the device list is fresh.
std::mutex g_images_mu;
std::map<std::string, void*>& images() {
    static std::map<std::string, void*>* m = new std::map<std::string, void*>();  // never destroyed: no exit-time dlclose
    return *m;
}

void* load_image(const char* name) {
    std::lock_guard<std::mutex> lock(g_images_mu);
    auto it = images().find(name);
    if (it != images().end()) return it->second;
    void* h = dlopen(name, RTLD_LAZY | RTLD_LOCAL);
    if (h != nullptr) images().emplace(name, h);
    return h;
}

}  // namespace

int NvmlSession::open() {
    if (handle_ != nullptr) return GMP_OK;
    const char* name = getenv("GPUMOUNTER_NVML_LIB");
    if (name == nullptr || name[0] == '\0') name = "libnvidia-ml.so.1";  // ref: nvml_dl.go:30
    void* h = load_image(name);
    if (h == nullptr) {
        set_error("could not load NVML library (%s)", dlerror());  // ref: bindings.go:53-55
        return GMP_ERR_NVML_NOT_FOUND;
    }
    struct Sym {
        void** slot;
        const char* name;
        bool required;
    };
    const Sym syms[] = {
        {(void**)&Init, "nvmlInit_v2", true},
        {(void**)&Shutdown, "nvmlShutdown", true},
        {(void**)&ErrorString, "nvmlErrorString", true},
        {(void**)&DeviceGetCount, "nvmlDeviceGetCount_v2", true},
        {(void**)&DeviceGetHandleByIndex, "nvmlDeviceGetHandleByIndex_v2", true},
        {(void**)&DeviceGetHandleByUUID, "nvmlDeviceGetHandleByUUID", true},
        {(void**)&DeviceGetMinorNumber, "nvmlDeviceGetMinorNumber", true},
        {(void**)&DeviceGetUUID, "nvmlDeviceGetUUID", true},
        {(void**)&DeviceGetPciInfo, "nvmlDeviceGetPciInfo_v3", false},
        {(void**)&DeviceGetIndex, "nvmlDeviceGetIndex", false},
        {(void**)&DeviceGetP2PStatus, "nvmlDeviceGetP2PStatus", false},
        {(void**)&DeviceGetNvLinkState, "nvmlDeviceGetNvLinkState", false},
        {(void**)&DeviceGetComputeRunningProcesses, "nvmlDeviceGetComputeRunningProcesses_v3", false},
        {(void**)&DeviceGetGraphicsRunningProcesses, "nvmlDeviceGetGraphicsRunningProcesses_v3", false},
        {(void**)&SystemGetDriverVersion, "nvmlSystemGetDriverVersion", false},
        {(void**)&SystemGetNVMLVersion, "nvmlSystemGetNVMLVersion", false},
        {(void**)&DeviceGetClockInfo, "nvmlDeviceGetClockInfo", false},
        {(void**)&DeviceGetMaxClockInfo, "nvmlDeviceGetMaxClockInfo", false},
        {(void**)&DeviceGetTemperature, "nvmlDeviceGetTemperature", false},
        {(void**)&DeviceGetPowerUsage, "nvmlDeviceGetPowerUsage", false},
        {(void**)&DeviceGetClocksReasons, "nvmlDeviceGetCurrentClocksThrottleReasons", false},
        {(void**)&DeviceGetTotalEccErrors, "nvmlDeviceGetTotalEccErrors", false},
        {(void**)&DeviceGetRemappedRows, "nvmlDeviceGetRemappedRows", false},
        {(void**)&DeviceGetFieldValues, "nvmlDeviceGetFieldValues", false},
    };
    for (const Sym& s : syms) {
        *s.slot = dlsym(h, s.name);
        if (*s.slot == nullptr && s.required) {
            set_error("could not load NVML library (%s lacks %s)", name, s.name);
            return GMP_ERR_NVML_NOT_FOUND;
        }
    }
    nvmlReturn_t r = Init();
    if (r != NVML_SUCCESS) {
        handle_ = h;
        int rc = fail(r, "nvmlInit_v2");
        handle_ = nullptr;
        return rc;
    }
    handle_ = h;
    return GMP_OK;
}

void NvmlSession::close() {
    if (handle_ == nullptr) return;
    if (Shutdown) (void)Shutdown();
    handle_ = nullptr;  // the image stays loaded, see load_image()
}

int NvmlSession::fail(nvmlReturn_t r, const char* what) const {
    const char* s = ErrorString ? ErrorString(r) : "unknown";
    set_error("nvml: %s (%s)", s, what);  // ref: bindings.go:47-48
    return GMP_ERR_NVML;
}

}  // namespace gmp

#include "gmp_nvml_hold.h"

namespace gmp {

NvmlHold::NvmlHold() : session_(nullptr) {
    NvmlSession* s = new (std::nothrow) NvmlSession();
    if (s != nullptr && s->open() != GMP_OK) {
        delete s;
        s = nullptr;
    }
    session_ = s;
}

NvmlHold::~NvmlHold() { delete static_cast<NvmlSession*>(session_); }

namespace {
std::mutex g_holds_mu;
std::vector<NvmlSession*>& holds() {
    static std::vector<NvmlSession*>* v = new std::vector<NvmlSession*>();
    return *v;
}
}  // namespace

}  // namespace gmp

using namespace gmp;

#define GMP_NVML(sess, call, what)                       \
    do {                                                 \
        nvmlReturn_t r__ = (call);                       \
        if (r__ != NVML_SUCCESS) return (sess).fail(r__, what); \
    } while (0)

extern "C" int gmp_nvml_hold(int acquire) try {
    clear_error();
    std::lock_guard<std::mutex> lock(g_holds_mu);
    if (acquire) {
        NvmlSession* s = new (std::nothrow) NvmlSession();
        if (s == nullptr) {
            set_error("out of memory");
            return GMP_ERR_NOMEM;
        }
        const int rc = s->open();
        if (rc != GMP_OK) {
            delete s;
            return rc;
        }
        holds().push_back(s);
        return GMP_OK;
    }
    if (holds().empty()) {
        set_error("gmp_nvml_hold(0) without a matching gmp_nvml_hold(1)");
        return GMP_ERR_INVALID;
    }
    delete holds().back();
    holds().pop_back();
    return GMP_OK;
} catch (...) {
    return ::gmp::caught_exception();
}

extern "C" int gmp_enumerate_nvml(gmp_device* out, int cap, int* n) try {
    clear_error();
    if (out == nullptr || n == nullptr || cap < 0) {
        set_error("NULL argument");
        return GMP_ERR_INVALID;
    }
    *n = 0;
    NvmlSession nv;
    int rc = nv.open();
    if (rc != GMP_OK) return rc;
    unsigned int count = 0;
    GMP_NVML(nv, nv.DeviceGetCount(&count), "nvmlDeviceGetCount_v2");
    if ((int)count > cap) {
        *n = (int)count;
        set_error("%u NVML devices, capacity %d", count, cap);
        return GMP_ERR_BUFFER_TOO_SMALL;
    }
    // index order, minor then UUID, exactly as ref: collector.go:56-77
    for (unsigned int i = 0; i < count; ++i) {
        nvmlDevice_t dev;
        GMP_NVML(nv, nv.DeviceGetHandleByIndex(i, &dev), "nvmlDeviceGetHandleByIndex_v2");
        gmp_device& d = out[i];
        memset(&d, 0, sizeof(d));
        unsigned int minor = 0;
        GMP_NVML(nv, nv.DeviceGetMinorNumber(dev, &minor), "nvmlDeviceGetMinorNumber");
        GMP_NVML(nv, nv.DeviceGetUUID(dev, d.uuid, GMP_UUID_LEN), "nvmlDeviceGetUUID");
        d.minor = minor;
        d.nvml_index = (int32_t)i;
        d.cuda_ordinal = -1;
        if (nv.DeviceGetPciInfo) {
            nvmlPciInfo_t pci;
            memset(&pci, 0, sizeof(pci));
            if (nv.DeviceGetPciInfo(dev, &pci) == NVML_SUCCESS) snprintf(d.pci_bus_id, GMP_PCI_LEN, "%s", pci.busId);
        }
    }
    *n = (int)count;
    return GMP_OK;
} catch (...) {
    return ::gmp::caught_exception();
}

extern "C" int gmp_nvml_versions(char* driver, size_t driver_cap, char* nvml, size_t nvml_cap) try {
    clear_error();
    NvmlSession nv;
    int rc = nv.open();
    if (rc != GMP_OK) return rc;
    if (driver != nullptr && driver_cap > 0) {
        driver[0] = '\0';
        if (nv.SystemGetDriverVersion)
            GMP_NVML(nv, nv.SystemGetDriverVersion(driver, (unsigned)driver_cap), "nvmlSystemGetDriverVersion");
    }
    if (nvml != nullptr && nvml_cap > 0) {
        nvml[0] = '\0';
        if (nv.SystemGetNVMLVersion)
            GMP_NVML(nv, nv.SystemGetNVMLVersion(nvml, (unsigned)nvml_cap), "nvmlSystemGetNVMLVersion");
    }
    return GMP_OK;
} catch (...) {
    return ::gmp::caught_exception();
}

extern "C" int gmp_gpu_running_processes(const char* uuid, gmp_process_info* out, int cap, int* n) try {
    clear_error();
    if (uuid == nullptr || n == nullptr || cap < 0 || (cap > 0 && out == nullptr)) {
        set_error("NULL argument");
        return GMP_ERR_INVALID;
    }
    *n = 0;
    NvmlSession nv;  // init -> by-UUID -> graphics + compute -> shutdown, ref: nvidia.go:58-87
    int rc = nv.open();
    if (rc != GMP_OK) return rc;
    nvmlDevice_t dev;
    GMP_NVML(nv, nv.DeviceGetHandleByUUID(uuid, &dev), "nvmlDeviceGetHandleByUUID");
    if (!nv.DeviceGetGraphicsRunningProcesses || !nv.DeviceGetComputeRunningProcesses) {
        set_error("nvml: running-process queries not exported by this NVML");
        return GMP_ERR_NVML;
    }
    // The reference sizes both queries at 1024 entries (ref: nvidia.go:74,80).
    std::vector<nvmlProcessInfo_t> procs(1024);
    int total = 0;
    for (int pass = 0; pass < 2; ++pass) {  // graphics first, then compute (ref: nvidia.go:74-86)
        unsigned int count = (unsigned)procs.size();
        nvmlReturn_t r = pass == 0 ? nv.DeviceGetGraphicsRunningProcesses(dev, &count, procs.data())
                                   : nv.DeviceGetComputeRunningProcesses(dev, &count, procs.data());
        if (r != NVML_SUCCESS)
            return nv.fail(r, pass == 0 ? "nvmlDeviceGetGraphicsRunningProcesses" : "nvmlDeviceGetComputeRunningProcesses");
        for (unsigned int i = 0; i < count; ++i) {
            if (total < cap) {
                out[total].pid = procs[i].pid;
                out[total].used_gpu_memory = procs[i].usedGpuMemory;
            }
            ++total;
        }
    }
    *n = total;
    if (total > cap) {
        set_error("%d GPU processes, capacity %d", total, cap);
        return GMP_ERR_BUFFER_TOO_SMALL;
    }
    return GMP_OK;
} catch (...) {
    return ::gmp::caught_exception();
}

extern "C" int gmp_nvml_p2p_status(const char* const* uuids, int n, int caps_index, int32_t* status) try {
    clear_error();
    if (uuids == nullptr || status == nullptr || n <= 0 || n > GMP_MAX_DEVICES) {
        set_error("bad argument (n=%d)", n);
        return GMP_ERR_INVALID;
    }
    if (caps_index < 0 || caps_index > (int)NVML_P2P_CAPS_INDEX_PCI) {
        set_error("unknown P2P caps index %d", caps_index);
        return GMP_ERR_INVALID;
    }
    NvmlSession nv;
    int rc = nv.open();
    if (rc != GMP_OK) return rc;
    if (!nv.DeviceGetP2PStatus) {
        set_error("nvml: nvmlDeviceGetP2PStatus not exported by this NVML");
        return GMP_ERR_NVML;
    }
    nvmlDevice_t dev[GMP_MAX_DEVICES];
    for (int i = 0; i < n; ++i) GMP_NVML(nv, nv.DeviceGetHandleByUUID(uuids[i], &dev[i]), "nvmlDeviceGetHandleByUUID");
    for (int i = 0; i < n; ++i) {
        for (int j = 0; j < n; ++j) {
            if (i == j) {
                status[i * n + j] = NVML_P2P_STATUS_OK;
                continue;
            }
            nvmlGpuP2PStatus_t st = NVML_P2P_STATUS_UNKNOWN;
            GMP_NVML(nv, nv.DeviceGetP2PStatus(dev[i], dev[j], (nvmlGpuP2PCapsIndex_t)caps_index, &st),
                     "nvmlDeviceGetP2PStatus");
            status[i * n + j] = (int32_t)st;
        }
    }
    return GMP_OK;
} catch (...) {
    return ::gmp::caught_exception();
}

extern "C" int gmp_nvml_health(const char* uuid, gmp_gpu_health* out) try {
    clear_error();
    if (uuid == nullptr || out == nullptr) {
        set_error("NULL argument");
        return GMP_ERR_INVALID;
    }
    memset(out, 0xFF, sizeof(*out));  // every field "unknown" until NVML answers it
    NvmlSession nv;
    int rc = nv.open();
    if (rc != GMP_OK) return rc;
    nvmlDevice_t dev;
    GMP_NVML(nv, nv.DeviceGetHandleByUUID(uuid, &dev), "nvmlDeviceGetHandleByUUID");
    unsigned int v = 0;
    if (nv.DeviceGetClockInfo && nv.DeviceGetClockInfo(dev, NVML_CLOCK_SM, &v) == NVML_SUCCESS) out->sm_mhz = v;
    if (nv.DeviceGetClockInfo && nv.DeviceGetClockInfo(dev, NVML_CLOCK_MEM, &v) == NVML_SUCCESS) out->mem_mhz = v;
    if (nv.DeviceGetMaxClockInfo && nv.DeviceGetMaxClockInfo(dev, NVML_CLOCK_SM, &v) == NVML_SUCCESS) out->sm_max_mhz = v;
    if (nv.DeviceGetMaxClockInfo && nv.DeviceGetMaxClockInfo(dev, NVML_CLOCK_MEM, &v) == NVML_SUCCESS) out->mem_max_mhz = v;
    if (nv.DeviceGetTemperature && nv.DeviceGetTemperature(dev, NVML_TEMPERATURE_GPU, &v) == NVML_SUCCESS) out->temperature_c = v;
    if (nv.DeviceGetPowerUsage && nv.DeviceGetPowerUsage(dev, &v) == NVML_SUCCESS) out->power_mw = v;
    unsigned long long w = 0;
    if (nv.DeviceGetClocksReasons && nv.DeviceGetClocksReasons(dev, &w) == NVML_SUCCESS) out->throttle_reasons = w;
    if (nv.DeviceGetTotalEccErrors) {
        if (nv.DeviceGetTotalEccErrors(dev, NVML_MEMORY_ERROR_TYPE_UNCORRECTED, NVML_VOLATILE_ECC, &w) == NVML_SUCCESS)
            out->ecc_uncorrected_volatile = w;
        if (nv.DeviceGetTotalEccErrors(dev, NVML_MEMORY_ERROR_TYPE_CORRECTED, NVML_VOLATILE_ECC, &w) == NVML_SUCCESS)
            out->ecc_corrected_volatile = w;
    }
    unsigned int corr = 0, unc = 0, pending = 0, failure = 0;
    if (nv.DeviceGetRemappedRows && nv.DeviceGetRemappedRows(dev, &corr, &unc, &pending, &failure) == NVML_SUCCESS) {
        out->remapped_rows_correctable = corr;
        out->remapped_rows_uncorrectable = unc;
        out->remap_pending = pending;
        out->remap_failure = failure;
    }
    if (nv.DeviceGetFieldValues) {
        nvmlFieldValue_t fv[4];
        memset(fv, 0, sizeof(fv));
        fv[0].fieldId = NVML_FI_DEV_NVLINK_CRC_FLIT_ERROR_COUNT_TOTAL;
        fv[1].fieldId = NVML_FI_DEV_NVLINK_CRC_DATA_ERROR_COUNT_TOTAL;
        fv[2].fieldId = NVML_FI_DEV_NVLINK_REPLAY_ERROR_COUNT_TOTAL;
        fv[3].fieldId = NVML_FI_DEV_NVLINK_RECOVERY_ERROR_COUNT_TOTAL;
        if (nv.DeviceGetFieldValues(dev, 4, fv) == NVML_SUCCESS) {
            uint64_t* slots[4] = {&out->nvlink_crc_flit_errors, &out->nvlink_crc_data_errors, &out->nvlink_replay_errors,
                                  &out->nvlink_recovery_errors};
            for (int k = 0; k < 4; ++k) {
                if (fv[k].nvmlReturn != NVML_SUCCESS) continue;
                switch (fv[k].valueType) {
                    case NVML_VALUE_TYPE_UNSIGNED_INT: *slots[k] = fv[k].value.uiVal; break;
                    case NVML_VALUE_TYPE_UNSIGNED_LONG: *slots[k] = fv[k].value.ulVal; break;
                    case NVML_VALUE_TYPE_UNSIGNED_LONG_LONG: *slots[k] = fv[k].value.ullVal; break;
                    default: break;  // a type these counters are not documented to have: left unknown
                }
            }
        }
    }
    return GMP_OK;
} catch (...) {
    return ::gmp::caught_exception();
}

extern "C" int gmp_nvml_nvlink_active(const char* uuid, int* active_links, int* max_links) try {
    clear_error();
    if (uuid == nullptr || active_links == nullptr) {
        set_error("NULL argument");
        return GMP_ERR_INVALID;
    }
    *active_links = 0;
    if (max_links) *max_links = NVML_NVLINK_MAX_LINKS;
    NvmlSession nv;
    int rc = nv.open();
    if (rc != GMP_OK) return rc;
    nvmlDevice_t dev;
    GMP_NVML(nv, nv.DeviceGetHandleByUUID(uuid, &dev), "nvmlDeviceGetHandleByUUID");
    if (!nv.DeviceGetNvLinkState) return GMP_OK;
    for (unsigned int l = 0; l < NVML_NVLINK_MAX_LINKS; ++l) {
        nvmlEnableState_t st = NVML_FEATURE_DISABLED;
        if (nv.DeviceGetNvLinkState(dev, l, &st) == NVML_SUCCESS && st == NVML_FEATURE_ENABLED) ++*active_links;
    }
    return GMP_OK;
} catch (...) {
    return ::gmp::caught_exception();
}

// NVML's record of the devices named by UUID, one by one (minor, index, bus id filled in; `found[i]` says which).
// For a caller INSIDE a pod that holds only some of the node's GPUs: there the index walk of gmp_enumerate_nvml, which
// follows the reference to the letter (ref: collector.go:56-77 returns at the first failing handle), stops at the first
// GPU whose node the pod's device cgroup does not let NVML open, while a by-UUID lookup of a mounted GPU still works.
namespace gmp {
int nvml_lookup_by_uuid(gmp_device* devs, int n, int* n_found) {
    *n_found = 0;
    NvmlSession nv;
    int rc = nv.open();
    if (rc != GMP_OK) return rc;
    int first_error = GMP_OK;
    for (int i = 0; i < n; ++i) {
        nvmlDevice_t dev;
        unsigned int minor = 0;
        nvmlReturn_t r = nv.DeviceGetHandleByUUID(devs[i].uuid, &dev);
        if (r == NVML_SUCCESS) r = nv.DeviceGetMinorNumber(dev, &minor);
        if (r != NVML_SUCCESS) {
            if (first_error == GMP_OK) first_error = nv.fail(r, "nvmlDeviceGetHandleByUUID");
            devs[i].nvml_index = -1;
            devs[i].minor = UINT32_MAX;
            continue;
        }
        devs[i].minor = minor;
        devs[i].nvml_index = -1;
        unsigned int index = 0;
        if (nv.DeviceGetIndex && nv.DeviceGetIndex(dev, &index) == NVML_SUCCESS) devs[i].nvml_index = (int32_t)index;
        if (nv.DeviceGetPciInfo) {
            nvmlPciInfo_t pci;
            memset(&pci, 0, sizeof(pci));
            if (nv.DeviceGetPciInfo(dev, &pci) == NVML_SUCCESS) snprintf(devs[i].pci_bus_id, GMP_PCI_LEN, "%s", pci.busId);
        }
        ++*n_found;
    }
    return *n_found == n ? GMP_OK : first_error;
}
}  // namespace gmp

// All four capability matrices and the active-link counts in ONE NVML session (internal; the P2P probe
// would otherwise pay NVML's ~100 ms load+init five times over).
namespace gmp {
int nvml_p2p_everything(const char* const* uuids, int n, int32_t* caps_out[4], int32_t* nvlink_active) {
    NvmlSession nv;
    int rc = nv.open();
    if (rc != GMP_OK) return rc;
    if (!nv.DeviceGetP2PStatus) {
        set_error("nvml: nvmlDeviceGetP2PStatus not exported by this NVML");
        return GMP_ERR_NVML;
    }
    nvmlDevice_t dev[GMP_MAX_DEVICES];
    for (int i = 0; i < n; ++i) GMP_NVML(nv, nv.DeviceGetHandleByUUID(uuids[i], &dev[i]), "nvmlDeviceGetHandleByUUID");
    for (int caps = 0; caps < 4; ++caps)
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < n; ++j) {
                nvmlGpuP2PStatus_t st = NVML_P2P_STATUS_OK;
                if (i != j)
                    GMP_NVML(nv, nv.DeviceGetP2PStatus(dev[i], dev[j], (nvmlGpuP2PCapsIndex_t)caps, &st), "nvmlDeviceGetP2PStatus");
                caps_out[caps][i * n + j] = (int32_t)st;
            }
    for (int i = 0; i < n; ++i) {
        nvlink_active[i] = 0;
        if (!nv.DeviceGetNvLinkState) continue;
        for (unsigned int l = 0; l < NVML_NVLINK_MAX_LINKS; ++l) {
            nvmlEnableState_t st = NVML_FEATURE_DISABLED;
            if (nv.DeviceGetNvLinkState(dev[i], l, &st) == NVML_SUCCESS && st == NVML_FEATURE_ENABLED) ++nvlink_active[i];
        }
    }
    return GMP_OK;
}
}  // namespace gmp
