#!/usr/bin/env python
"""What the one-shot probe call spends outside its kernels: cudaMalloc / cudaFree of the probe buffers, cudaHostAlloc /
cudaFreeHost of the DMA staging buffer, and the stream-ordered allocator (cudaMallocAsync with the pool kept) as the
alternative, per size.  Guides the work on the `e2e` call.  GPU box only; prints one JSON object.

    gpurun -- 'python tools/alloc_costs.py > gpurun_out/alloc_costs.json'
"""

from __future__ import annotations

import json
import statistics
import sys
import time

try:
    from cuda.bindings import driver as drv, runtime as rt
except ImportError:  # older cuda-python
    from cuda import cuda as drv, cudart as rt


def ok(ret):
    err, *rest = ret
    if int(err) != 0:
        raise RuntimeError(f"{err}")
    return rest[0] if len(rest) == 1 else (rest or None)


def main():
    err, n = rt.cudaGetDeviceCount()
    if int(err) != 0 or n == 0:
        print(json.dumps({"unavailable": f"no CUDA device ({err})"}))
        return 0
    ok(rt.cudaSetDevice(0))
    ok(rt.cudaFree(0))
    stream = ok(rt.cudaStreamCreate())
    out = {"device_malloc_free_ms": {}, "pinned_alloc_free_ms": {}, "malloc_async_pool_kept_ms": {}}
    for gib in (1, 2, 4, 8, 16):
        size = gib << 30
        allocs, frees = [], []
        for _ in range(5):
            t0 = time.perf_counter()
            p = ok(rt.cudaMalloc(size))
            t1 = time.perf_counter()
            ok(rt.cudaFree(p))
            t2 = time.perf_counter()
            allocs.append((t1 - t0) * 1e3)
            frees.append((t2 - t1) * 1e3)
        out["device_malloc_free_ms"][f"{gib}GiB"] = {"malloc_p50": statistics.median(allocs), "free_p50": statistics.median(frees)}
    for mib in (1, 4, 16, 64):
        size = mib << 20
        allocs, frees = [], []
        for _ in range(5):
            t0 = time.perf_counter()
            p = ok(rt.cudaHostAlloc(size, 0))
            t1 = time.perf_counter()
            ok(rt.cudaFreeHost(p))
            t2 = time.perf_counter()
            allocs.append((t1 - t0) * 1e3)
            frees.append((t2 - t1) * 1e3)
        out["pinned_alloc_free_ms"][f"{mib}MiB"] = {"alloc_p50": statistics.median(allocs), "free_p50": statistics.median(frees)}
    # stream-ordered allocator with the pool's release threshold raised: memory stays with the process between calls
    try:
        pool_section(out, stream)
    except Exception as e:  # the binding of this one call has changed between cuda-python releases
        out["malloc_async_pool_kept_ms"] = {"unavailable": f"{type(e).__name__}: {e}"}
    print(json.dumps(out))
    return 0


def pool_section(out, stream):
    pool = ok(rt.cudaDeviceGetDefaultMemPool(0))
    ok(rt.cudaMemPoolSetAttribute(pool, rt.cudaMemPoolAttr.cudaMemPoolAttrReleaseThreshold, drv.cuuint64_t((1 << 64) - 1)))
    for gib in (1, 8):
        size = gib << 30
        allocs = []
        for _ in range(5):
            t0 = time.perf_counter()
            p = ok(rt.cudaMallocAsync(size, stream))
            ok(rt.cudaStreamSynchronize(stream))
            t1 = time.perf_counter()
            ok(rt.cudaFreeAsync(p, stream))
            ok(rt.cudaStreamSynchronize(stream))
            allocs.append((t1 - t0) * 1e3)
        out["malloc_async_pool_kept_ms"][f"{gib}GiB"] = {"first": allocs[0], "later_p50": statistics.median(allocs[1:])}


if __name__ == "__main__":
    sys.exit(main())
