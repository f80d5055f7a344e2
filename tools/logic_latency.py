#!/usr/bin/env python
"""Metric 0 of SURVEY.md 8(d): microseconds per call of the pure path/parsing/selection logic, library (C++ behind the C
ABI) next to the oracle's plain-C restatement, both through ctypes with the cost of an empty ctypes call subtracted.
The reference has no such measurement (and no tests for these functions); the point of the number is only that this
logic is nowhere near the mount latency (hundreds of microseconds) - it is string handling.  CPU only.  One JSON object."""

from __future__ import annotations

import ctypes
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from gpumounter_b200 import cgroup2, lib  # noqa: E402
from gpumounter_b200._model import ContainerResources, Pod, gmp_gpu, gmp_ns_config  # noqa: E402
from oracle import gmo_ref_ctypes as ref  # noqa: E402  (comparison arm of a measurement tool, like bench.py's cpu_baseline)

UID = "0a1b2c3d-1111-2222-3333-444455556666"
CID = "c0ffee" * 10 + "beef"
N = 20000


def per_call_us(fn, n=N):
    fn()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    return 1e6 * (time.perf_counter() - t0) / n


def main():
    L, O = lib(), ref._lib()
    empty = min(per_call_us(lambda: L.gmp_abi_version()) for _ in range(3))
    buf = ctypes.create_string_buffer(8192)
    res = dict(cpu_request="500m", memory_request="1Gi", cpu_limit="2", memory_limit="1Gi")  # a Burstable pod
    cp = Pod("gpu-pod", "default", uid=UID, container_id="docker://" + CID, containers=[ContainerResources(**res)]).to_c()
    op = ref.Pod(uid=UID, container_id="docker://" + CID, containers=[ref.Resources(**res)]).c()
    q = ctypes.c_int()
    g = gmp_gpu()
    L.gmp_device_new(3, b"GPU-f61ffc1a-9e61-1c0e-2211-4f8f252fe7bc", ctypes.byref(g))
    og = ref.gmo_gpu()
    O.gmo_device_new(3, b"GPU-f61ffc1a-9e61-1c0e-2211-4f8f252fe7bc", ctypes.byref(og))
    cfg = gmp_ns_config()
    cfg.mount, cfg.target = 1, 4242
    argc = ctypes.c_int()
    rules = (cgroup2.gmp_dev_rule * 12)(*[cgroup2.Rule(2, 1, i, 7).to_c() for i in range(11)], cgroup2.gpu_rule(3).to_c())
    insns, n_insns = (cgroup2.gmp_bpf_insn * 256)(), ctypes.c_int()
    L.gmp_devfilter_build(rules, 11, 0, insns, 256, ctypes.byref(n_insns))
    base_n = n_insns.value
    out_insns, n_out = (cgroup2.gmp_bpf_insn * 256)(), ctypes.c_int()
    one = (cgroup2.gmp_dev_rule * 1)(cgroup2.gpu_rule(5).to_c())

    rows = {
        "GetPodQOS": (lambda: L.gmp_get_pod_qos(ctypes.byref(cp), ctypes.byref(q)), None),
        "GetCgroupName(systemd)": (lambda: L.gmp_get_cgroup_name(b"systemd", ctypes.byref(cp), CID.encode(), buf, 8192),
                                   lambda: O.gmo_get_cgroup_name(b"systemd", ctypes.byref(op), CID.encode(), buf, 8192)),
        "GetCgroupName(cgroupfs)": (lambda: L.gmp_get_cgroup_name(b"cgroupfs", ctypes.byref(cp), CID.encode(), buf, 8192),
                                    lambda: O.gmo_get_cgroup_name(b"cgroupfs", ctypes.byref(op), CID.encode(), buf, 8192)),
        "NvidiaGPU.String": (lambda: L.gmp_device_string(ctypes.byref(g), buf, 8192), lambda: O.gmo_device_string(ctypes.byref(og), buf, 8192)),
        "namespace.buildCommand": (lambda: L.gmp_ns_build_command(ctypes.byref(cfg), buf, 8192, ctypes.byref(argc)), None),
        "devfilter build (12 rules)": (lambda: L.gmp_devfilter_build(rules, 12, 0, insns, 256, ctypes.byref(n_insns)), None),
        "devfilter plan (grant 1 GPU)": (lambda: L.gmp_devfilter_plan(insns, base_n, one, 1, 1, out_insns, 256, ctypes.byref(n_out)), None),
    }
    out = {"empty_ctypes_call_us": round(empty, 3), "calls_per_row": N, "host_cores_used": 1, "rows": {}}
    for name, (product, oracle) in rows.items():
        row = {"library_us": round(max(0.0, per_call_us(product) - empty), 3)}
        if oracle is not None:
            row["oracle_us"] = round(max(0.0, per_call_us(oracle) - empty), 3)
        out["rows"][name] = row
    print(json.dumps(out))
    return 0


if __name__ == "__main__":
    sys.exit(main())
