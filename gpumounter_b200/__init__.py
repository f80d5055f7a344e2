"""gpumounter_b200 — B200-native post-mount GPU probe and node-local mount path for GPUMounter.

The product is ``libgpumounter_probe.so`` (C ABI in ``include/gpumounter_probe.h``,
C++/CUDA sources in ``gpumounter_b200/csrc``).  This package is the thin host-side
mirror of the reference's Go packages over that ABI via ctypes, so that tests read
like the reference's own call sites:

    gpumounter_b200.probe      the new post-mount probe (no reference counterpart)
    gpumounter_b200.nvml       ref: pkg/util/gpu/collector/nvml
    gpumounter_b200.device     ref: pkg/device
    gpumounter_b200.cgroup     ref: pkg/util/cgroup
    gpumounter_b200.namespace  ref: pkg/util/namespace
    gpumounter_b200.util       ref: pkg/util/util.go
    gpumounter_b200.allocator  ref: pkg/util/gpu/allocator (selection logic only)
    gpumounter_b200.collector  ref: pkg/util/gpu/collector
    gpumounter_b200.api        ref: pkg/api/gpu-mount (wire contract, hand-built descriptor)
    gpumounter_b200.server     ref: pkg/server/gpu-mount + cmd/GPUMounter-worker (handlers over a Cluster interface)

There is no CPU fallback anywhere: importing works without a GPU, but every compute
entry point raises ``ProbeError(GMP_ERR_UNAVAILABLE)`` on a box with no CUDA driver, and
loading fails loudly when the shared library has not been built.
"""

from ._lib import ProbeError, lib, lib_path  # noqa: F401

__all__ = ["ProbeError", "lib", "lib_path"]
