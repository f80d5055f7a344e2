// dlopen-based NVML session used by the enumerator, the process query and the
// P2P matrix check.  Built against the CUDA toolkit's nvml.h (API v12: 18 NVLink
// links, nvmlProcessInfo_t v3) for types only; no NVML symbol is linked, so the
// library loads on boxes without a driver
// (same idea as ref: pkg/util/gpu/collector/nvml/nvml_dl.go:29-36, bindings.go:20).
#pragma once

#include <nvml.h>

#include "gpumounter_probe.h"

namespace gmp {

class NvmlSession {
  public:
    NvmlSession() = default;
    ~NvmlSession() { close(); }
    NvmlSession(const NvmlSession&) = delete;
    NvmlSession& operator=(const NvmlSession&) = delete;

    // dlopen (once per process and library name; the image is kept) + nvmlInit_v2.  GMP_OK, GMP_ERR_NVML_NOT_FOUND or GMP_ERR_NVML.
    int open();
    // nvmlShutdown.  Unlike ref: nvml_dl.go:39-53 there is no dlclose: every fresh image of the driver's
    // library leaks descriptors at its first init (gmp_nvml.cpp, load_image).
    void close();
    bool is_open() const { return handle_ != nullptr; }

    // Records "nvml: <nvmlErrorString>" (ref: bindings.go:43-49) and returns GMP_ERR_NVML.
    int fail(nvmlReturn_t r, const char* what) const;

    nvmlReturn_t (*Init)(void) = nullptr;
    nvmlReturn_t (*Shutdown)(void) = nullptr;
    const char* (*ErrorString)(nvmlReturn_t) = nullptr;
    nvmlReturn_t (*DeviceGetCount)(unsigned int*) = nullptr;
    nvmlReturn_t (*DeviceGetHandleByIndex)(unsigned int, nvmlDevice_t*) = nullptr;
    nvmlReturn_t (*DeviceGetHandleByUUID)(const char*, nvmlDevice_t*) = nullptr;
    nvmlReturn_t (*DeviceGetMinorNumber)(nvmlDevice_t, unsigned int*) = nullptr;
    nvmlReturn_t (*DeviceGetUUID)(nvmlDevice_t, char*, unsigned int) = nullptr;
    nvmlReturn_t (*DeviceGetPciInfo)(nvmlDevice_t, nvmlPciInfo_t*) = nullptr;
    nvmlReturn_t (*DeviceGetIndex)(nvmlDevice_t, unsigned int*) = nullptr;
    nvmlReturn_t (*DeviceGetP2PStatus)(nvmlDevice_t, nvmlDevice_t, nvmlGpuP2PCapsIndex_t, nvmlGpuP2PStatus_t*) = nullptr;
    nvmlReturn_t (*DeviceGetNvLinkState)(nvmlDevice_t, unsigned int, nvmlEnableState_t*) = nullptr;
    nvmlReturn_t (*DeviceGetComputeRunningProcesses)(nvmlDevice_t, unsigned int*, nvmlProcessInfo_t*) = nullptr;
    nvmlReturn_t (*DeviceGetGraphicsRunningProcesses)(nvmlDevice_t, unsigned int*, nvmlProcessInfo_t*) = nullptr;
    nvmlReturn_t (*SystemGetDriverVersion)(char*, unsigned int) = nullptr;
    nvmlReturn_t (*SystemGetNVMLVersion)(char*, unsigned int) = nullptr;
    nvmlReturn_t (*DeviceGetClockInfo)(nvmlDevice_t, nvmlClockType_t, unsigned int*) = nullptr;
    nvmlReturn_t (*DeviceGetMaxClockInfo)(nvmlDevice_t, nvmlClockType_t, unsigned int*) = nullptr;
    nvmlReturn_t (*DeviceGetTemperature)(nvmlDevice_t, nvmlTemperatureSensors_t, unsigned int*) = nullptr;
    nvmlReturn_t (*DeviceGetPowerUsage)(nvmlDevice_t, unsigned int*) = nullptr;
    nvmlReturn_t (*DeviceGetClocksReasons)(nvmlDevice_t, unsigned long long*) = nullptr;
    nvmlReturn_t (*DeviceGetTotalEccErrors)(nvmlDevice_t, nvmlMemoryErrorType_t, nvmlEccCounterType_t, unsigned long long*) = nullptr;
    nvmlReturn_t (*DeviceGetRemappedRows)(nvmlDevice_t, unsigned int*, unsigned int*, unsigned int*, unsigned int*) = nullptr;
    nvmlReturn_t (*DeviceGetFieldValues)(nvmlDevice_t, int, nvmlFieldValue_t*) = nullptr;

  private:
    void* handle_ = nullptr;
};

}  // namespace gmp
