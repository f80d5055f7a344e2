// No dependencies on purpose: included by the host-only translation units (gmp_mount.cpp) as well.
#pragma once

namespace gmp {

// Holds one nvmlInit_v2 reference for its lifetime, so that the per-GPU queries a batched call makes (each of which
// brackets itself with init/shutdown, as ref: nvidia.go:58-63 does) pay the driver attach once, not once per GPU.
// Best effort: if NVML cannot be opened the hold is empty and the queries report the error themselves.
class NvmlHold {
  public:
    NvmlHold();
    ~NvmlHold();
    NvmlHold(const NvmlHold&) = delete;
    NvmlHold& operator=(const NvmlHold&) = delete;

  private:
    void* session_;
};

}  // namespace gmp
