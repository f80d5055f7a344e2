#!/usr/bin/env python
"""Mount latency, node-local: the reference's fork/exec path (restated in oracle/gmo_reference.c) next to the
native path, on a cluster-less target; and, on a GPU box, mount-to-usable with the in-pod probe.

    reference : per GPU `sh -c "echo 'c 195:N rw' > .../devices.allow"` + `nsenter --target P --mount sh -c "mknod ..."`,
                serial over the G GPUs as at ref: pkg/server/gpu-mount/server.go:80-95; single-threaded.
    native    : gmp_mount_gpus — write(2) per GPU, one setns child with G mknod(2) calls.
    to-usable : t0 = entry of the native mount -> device nodes made -> gpumounter-probe fexecve'd inside the target
                (device cgroup + mount namespace) -> CLOCK_MONOTONIC instant at which the first kernel result verified
                on EVERY mounted GPU (the probe reports it).  Cold process each time: cuInit, context, cudaMalloc.

Slave-pod scheduling (seconds, ref: allocator.go:40-99) needs a cluster and is not part of these numbers.
Prints one JSON object.
"""

from __future__ import annotations

import argparse
import json
import os
import stat
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from gpumounter_b200 import device, probe, util  # noqa: E402
from gpumounter_b200._model import MountEnv, Pod  # noqa: E402
from tools import standin_pod as harness  # noqa: E402  (the stand-in pod: test + bench infrastructure)


def pct(xs, q):
    xs = sorted(xs)
    if not xs:
        return None
    k = min(len(xs) - 1, max(0, int(round(q * (len(xs) - 1)))))
    return xs[k]


def summary(xs_s):
    return {"p50_ms": 1e3 * statistics.median(xs_s), "p99_ms": 1e3 * pct(xs_s, 0.99), "min_ms": 1e3 * min(xs_s),
            "mean_ms": 1e3 * statistics.fmean(xs_s), "n": len(xs_s)}


def product_pod():
    return Pod("gpu-pod", uid=harness.UID, container_id="docker://" + harness.CID)


def env_of(t):
    if getattr(t, "cgroup2", False):  # extension (header section 7): the unified hierarchy, eBPF device program
        from gpumounter_b200._model import EXT_CGROUP_V2
        return MountEnv(cgroup2_root=t.devices_root, dev_prefix=t.dev_prefix, cgroup_driver="cgroupfs", extensions=EXT_CGROUP_V2)
    return MountEnv(devices_root=t.devices_root, dev_prefix=t.dev_prefix, cgroup_driver="cgroupfs")


def fake_gpus(n):
    return [(f"GPU-{i:08x}-0000-0000-0000-{i:012x}", i) for i in range(n)]


def open_fds():
    return len(os.listdir("/proc/self/fd"))


def reference_running_processes(uuid):
    """GetRunningProcess as the reference does it (ref: pkg/device/nvidia.go:58-87 over nvml_dl.go:29-53): dlopen, nvmlInit_v2,
    handle by UUID, graphics then compute processes (1024 entries each), nvmlShutdown, dlclose - once per GPU.  Only used on a
    box with a real driver, so that the reference arm of the unmount timing pays the same NVML busy check the native arm does."""
    import ctypes
    import _ctypes

    class ProcessInfo(ctypes.Structure):  # nvmlProcessInfo_t (v3): pid, usedGpuMemory, gpuInstanceId, computeInstanceId
        _fields_ = [("pid", ctypes.c_uint), ("mem", ctypes.c_ulonglong), ("gi", ctypes.c_uint), ("ci", ctypes.c_uint)]

    h = _ctypes.dlopen("libnvidia-ml.so.1", os.RTLD_LAZY | os.RTLD_LOCAL)
    try:
        lib = ctypes.CDLL("libnvidia-ml.so.1", handle=h)
        assert lib.nvmlInit_v2() == 0
        try:
            dev = ctypes.c_void_p()
            assert lib.nvmlDeviceGetHandleByUUID(uuid.encode(), ctypes.byref(dev)) == 0
            pids = []
            for fn in (lib.nvmlDeviceGetGraphicsRunningProcesses_v3, lib.nvmlDeviceGetComputeRunningProcesses_v3):
                buf, n = (ProcessInfo * 1024)(), ctypes.c_uint(1024)
                assert fn(dev, ctypes.byref(n), buf) == 0
                pids += [buf[i].pid for i in range(n.value)]
            return pids
        finally:
            lib.nvmlShutdown()
    finally:
        _ctypes.dlclose(h)


def bench_reference(t, gpus, iters, nvml_busy_check=False):
    """The oracle's fork/exec restatement: the timed CPU baseline (1 thread)."""
    from oracle import gmo_ref_ctypes as ref  # the checker doubles as the timed reference arm, nothing else

    pod = ref.Pod(uid=harness.UID, container_id="docker://" + harness.CID)
    kw = dict(devices_root=t.devices_root, dev_prefix=t.dev_prefix, cgroup_driver="cgroupfs")
    recs = [ref.device_new(m, u) for u, m in gpus]
    mount_s, unmount_s = [], []
    for _ in range(iters):
        t0 = time.perf_counter()
        for r in recs:  # serial, as server.go:80-95
            ref.mount_gpu(pod, r, **kw)
        t1 = time.perf_counter()
        for (u, _), r in zip(gpus, recs):  # server.go:155-168
            ref.unmount_gpu(pod, r, False, reference_running_processes(u) if nvml_busy_check else [], **kw)
        t2 = time.perf_counter()
        mount_s.append(t1 - t0)
        unmount_s.append(t2 - t1)
    return summary(mount_s), summary(unmount_s)


def bench_native(t, gpus, iters, batched=True):
    pod, env = product_pod(), env_of(t)
    recs = [device.New(m, u) for u, m in gpus]
    mount_s, unmount_s = [], []
    for _ in range(iters):
        t0 = time.perf_counter()
        if batched:
            util.MountGPUs(pod, recs, env)
        else:
            for r in recs:
                util.MountGPU(pod, r, env)
        t1 = time.perf_counter()
        if batched:
            util.UnmountGPUs(pod, recs, False, env)
        else:
            for r in recs:
                util.UnmountGPU(pod, r, False, env)
        t2 = time.perf_counter()
        mount_s.append(t1 - t0)
        unmount_s.append(t2 - t1)
    return summary(mount_s), summary(unmount_s)


def churn(t, rounds):
    """1 -> 8 -> 1 GPUs, `rounds` times, native path; checks that nothing leaks."""
    pod, env = product_pod(), env_of(t)
    recs = [device.New(m, u) for u, m in fake_gpus(8)]
    fds0 = open_fds()
    step_s = []
    t_all = time.perf_counter()
    for _ in range(rounds):
        for k in list(range(1, 9)) + list(range(7, 0, -1)):
            t0 = time.perf_counter()
            util.MountGPUs(pod, recs[:k], env)
            assert len(t.nodes()) == k
            util.UnmountGPUs(pod, recs[:k], False, env)
            step_s.append(time.perf_counter() - t0)
        assert t.nodes() == []
    total = time.perf_counter() - t_all
    return {"rounds": rounds, "steps": len(step_s), "total_s": total, "step": summary(step_s),
            "leaked_fds": open_fds() - fds0, "leaked_nodes": len(t.nodes()), "host_sees": t.host_sees(),
            "zombies": sum(1 for p in os.listdir("/proc") if p.isdigit() and _is_zombie_child(p))}


def _is_zombie_child(pid):
    try:
        with open(f"/proc/{pid}/stat") as f:
            parts = f.read().rsplit(")", 1)[1].split()
        return parts[0] == "Z" and int(parts[1]) == os.getpid()
    except Exception:
        return False


def env_in(t):
    """MountEnv of a target: a pod-like one uses the reference's own /dev/nvidia<N> names inside its mount namespace."""
    if getattr(t, "pod_like", False):
        return MountEnv(devices_root=t.devices_root, cgroup_driver="cgroupfs")
    return env_of(t)


def mknod_like(src, dst):
    st = os.stat(src)
    os.mknod(dst, stat.S_IFCHR | 0o666, st.st_rdev)
    os.chmod(dst, 0o666)


TARGET_POD_LIKE = "private /dev + real devices cgroup"
TARGET_SCRATCH = "scratch dev directory + scratch cgroup files (host /dev shared; CUDA restricted by CUDA_VISIBLE_DEVICES)"


def pod_like_available():
    return harness.can_run_pod_like() is None


class GpuTarget:
    """The stand-in pod for the GPU measurements.  pod_like (the default wherever the box has a writable cgroup-v1
    devices hierarchy): tmpfs ON /dev of the target's mount namespace without any /dev/nvidia<N>, a real devices cgroup
    with a container's whitelist, and NO CUDA_VISIBLE_DEVICES - CUDA in the pod can only open what the mount path made.
    The scratch shape (round 1) remains as a fallback and says so in the output."""

    def __init__(self, pod_like=None):
        self.pod_like = pod_like_available() if pod_like is None else pod_like
        self.t = harness.Target(pod_like=True) if self.pod_like else harness.Target()
        self.label = TARGET_POD_LIKE if self.pod_like else TARGET_SCRATCH

    def __enter__(self):
        self.t.start()
        if not self.pod_like:
            for name in ("nvidiactl", "nvidia-uvm"):
                mknod_like(f"/dev/{name}", f"/proc/{self.t.pid}/root{self.t.devdir}/{name}")
        return self

    def __exit__(self, *exc):
        if self.pod_like:  # the main process must leave the group before the group can be removed
            try:
                with open(os.path.join(harness.V1_DEVICES_ROOT, "cgroup.procs"), "w") as f:
                    f.write(f"{self.t.pid}\n")
            except OSError:
                pass
        self.t.stop()

    def env(self):
        return env_in(self.t)

    def probe_args(self):
        return [] if self.pod_like else ["--dev-dir", self.t.devdir]

    def restrict_cuda(self, gpus):
        """pod-like: NOTHING may restrict CUDA through the environment the in-pod probe inherits (a box's launcher often
        exports CUDA_VISIBLE_DEVICES); scratch: the environment is the only restriction there is."""
        self.saved_env = os.environ.pop("CUDA_VISIBLE_DEVICES", None)
        if not self.pod_like:
            os.environ["CUDA_VISIBLE_DEVICES"] = ",".join(u for u, _ in gpus)

    def unrestrict_cuda(self):
        os.environ.pop("CUDA_VISIBLE_DEVICES", None)
        if getattr(self, "saved_env", None) is not None:
            os.environ["CUDA_VISIBLE_DEVICES"] = self.saved_env


def mounted_set(n_gpus):
    nv = {d.uuid: d for d in probe.enumerate_nvml()}
    cu = probe.enumerate_cuda()[:n_gpus]
    return [(d.uuid, nv[d.uuid].minor) for d in cu]


def to_usable(iters, hbm_bytes, n_gpus, pod_like=None, p2p_bytes=256 << 20, fanout=False):
    """Real GPUs only: native mount into the target, then the in-pod probe; t1 is reported by the probe.
    fanout: gmp_probe_in_pod_fanout - one probe process per GPU, each confined to its own GPU (see the header)."""
    gpus = mounted_set(n_gpus)
    lat, lat_first, phases, last, label = [], [], [], None, None
    for _ in range(iters):
        with GpuTarget(pod_like) as g:
            label = g.label
            recs = [device.New(m, u) for u, m in gpus]
            pod, env = product_pod(), g.env()
            g.restrict_cuda(gpus)
            args = g.probe_args() + ["--bytes", str(hbm_bytes), "--iters", "1", "--dma-bytes", "16M",
                                     "--p2p-bytes", str(p2p_bytes), "--p2p-iters", "1"]
            t0 = time.monotonic_ns()
            util.MountGPUs(pod, recs, env)
            t_mounted = time.monotonic_ns()
            if fanout:
                rc, rep = util.ProbeInPodFanout(pod, recs, args, env, raise_on_error=False)
            else:
                rc, rep = util.ProbeInPod(pod, recs, args, env, raise_on_error=False)
            t_done = time.monotonic_ns()
            g.unrestrict_cuda()
            assert rc == 0 and rep.get("status") == 0, (rc, rep, probe.lib().gmp_last_error())
            last = rep
            lat.append((rep["mono_usable_ns"] - t0) / 1e9)
            first = min(h["mono_first_verified_ns"] for h in rep["hbm"])
            lat_first.append((first - t0) / 1e9)
            phases.append({"mount_ms": (t_mounted - t0) / 1e6, "exec_to_probe_start_ms": (rep["mono_start_ns"] - t_mounted) / 1e6,
                           "probe_start_to_usable_ms": rep["ms_to_usable"], "cuda_init_ms": rep["ms_cuda_init"],
                           "nvml_ms_overlapped": rep["ms_nvml"], "context_ms": max(h["ms_context"] for h in rep["hbm"]),
                           "alloc_ms": max(h["ms_alloc"] for h in rep["hbm"]),
                           "first_kernel_ms": max(h["ms_first_kernel"] for h in rep["hbm"]),
                           "first_gpu_usable_ms": (first - t0) / 1e6,
                           "all_gpus_usable_ms": (rep["mono_usable_ns"] - t0) / 1e6,
                           "full_probe_ms": (t_done - t_mounted) / 1e6})
    med = {k: statistics.median(p[k] for p in phases) for k in phases[0]}
    out = {"gpus": len(gpus), "hbm_bytes_per_gpu": hbm_bytes, "iters": iters, "target": label,
           "probe": "one process per GPU (fan-out)" if fanout else "one process, one host thread per GPU",
           "p50_s": statistics.median(lat), "p99_s": pct(lat, 0.99), "min_s": min(lat), "max_s": max(lat),
           "first_gpu_usable_p50_s": statistics.median(lat_first), "cuda_init_ms": med["cuda_init_ms"],
           "median_phases_ms": med,
           "device_set_equals_nvml": last["device_set_status"] == 0 and last["nvml_status"] == 0,
           "context_timeline_last_ms": [{"uuid": h["uuid"], "begin": (h.get("mono_context_begin_ns", 0) - last["mono_start_ns"]) / 1e6,
                                         "end": (h.get("mono_context_end_ns", 0) - last["mono_start_ns"]) / 1e6}
                                        for h in last["hbm"]],
           "hbm_gbs_per_gpu_last": [{"write": h["write_gbs"], "verify": h["read_gbs"], "copy": h["copy_gbs"]} for h in last["hbm"]]}
    if "p2p" in last:
        out["p2p"] = p2p_summary(last["p2p"])
    return out


def p2p_summary(q):
    n = len(q["can_access"])
    off = [(i, j) for i in range(n) for j in range(n) if i != j]

    def stats(m):
        v = [m[i][j] for i, j in off]
        return {"min": min(v), "mean": statistics.fmean(v), "max": max(v)}

    return {"matrix_match_nvml": q["matrix_match"] == 1, "pairs": len(off),
            "all_pairs_accessible": all(q["can_access"][i][j] == 1 for i, j in off),
            "all_pairs_nvlink": all(q["nvml_nvlink"][i][j] == 0 for i, j in off),
            "store_gbs": stats(q["write_gbs"]), "load_gbs": stats(q["read_gbs"]), "memcpy_peer_gbs": stats(q["memcpy_peer_gbs"]),
            "bidir_gbs_per_gpu_min": min(q["bidir_gbs"]), "alltoall_gbs": q["alltoall_gbs"],
            "alltoall_gbs_per_gpu": q["alltoall_gbs"] / n,
            "nvlink_active": q["nvlink_active"], "bytes_per_pair": q["bytes"], "mismatch_lanes": q["mismatch_lanes"]}


def to_usable_resident(iters, hbm_bytes, n_gpus, keep_contexts, skip_p2p=False, release_at_end=True, pod_like=None):
    """Two-stage hook with a resident worker (this process stands in for it: on the host, sees every GPU, cuInit paid once at
    start-up): native mount -> gmp_check_nodes_in_pod (a helper joins the pod's device cgroup + mount namespace and opens the
    nodes there; nothing executed, no CUDA) -> gmp_probe_all(HOST_VIEW) in-process on the mounted GPUs.  keep_contexts: the worker retains the
    primary contexts between calls (KEEP_DEVICE) or releases them (cudaDeviceReset) after each probe.
    What this does NOT time is the tenant's own CUDA start-up in the pod - that is to_usable()."""
    gpus = mounted_set(n_gpus)  # cuInit happens here at the latest
    expected = [probe.Device(u, m, "", -1, -1) for u, m in gpus]
    flags = probe.PROBE_F_HOST_VIEW | probe.PROBE_F_SKIP_NODES | (probe.PROBE_F_KEEP_DEVICE if keep_contexts else 0)
    if skip_p2p:  # a caller that shares the process with other CUDA users (bench.py) must not toggle peer access
        flags |= probe.PROBE_F_SKIP_P2P
    lat, phases, label = [], [], None
    for it in range(iters + 1):  # iteration 0 warms the worker (first contexts, module load); not counted
        with GpuTarget(pod_like) as g:
            label = g.label
            recs = [device.New(m, u) for u, m in gpus]
            pod, env = product_pod(), g.env()
            t0 = time.monotonic_ns()
            util.MountGPUs(pod, recs, env)
            t_mounted = time.monotonic_ns()
            rc, failed = util.CheckNodesInPod(pod, recs, env, raise_on_error=False)   # helper in the pod's cgroup + mount ns, no exec
            t_nodes = time.monotonic_ns()
            assert rc == 0, (rc, failed, probe.lib().gmp_last_error())
            r = probe.probe_all(expected, hbm_bytes=hbm_bytes, hbm_iters=1, dma_bytes=16 << 20, p2p_bytes=256 << 20, p2p_iters=1,
                                flags=flags, raise_on_error=False)
            t_done = time.monotonic_ns()
            assert r.status == 0, r.detail
            rep = json.loads(probe.report_json(r))
            util.UnmountGPUs(pod, recs, False, env)
            if it == 0:
                continue
            lat.append((rep["mono_usable_ns"] - t0) / 1e9)
            phases.append({"mount_ms": (t_mounted - t0) / 1e6, "nodes_only_in_pod_ms": (t_nodes - t_mounted) / 1e6,
                           "probe_start_to_usable_ms": rep["ms_to_usable"], "context_ms": max(h["ms_context"] for h in rep["hbm"]),
                           "alloc_ms": max(h["ms_alloc"] for h in rep["hbm"]),
                           "first_kernel_ms": max(h["ms_first_kernel"] for h in rep["hbm"]),
                           "full_probe_ms": (t_done - t_nodes) / 1e6})
    if keep_contexts and release_at_end:
        # hand the contexts back: one last probe without KEEP_DEVICE resets every device it touched
        probe.probe_all(expected, hbm_bytes=1 << 20, hbm_iters=1, dma_bytes=0,
                        flags=probe.PROBE_F_HOST_VIEW | probe.PROBE_F_SKIP_NODES | probe.PROBE_F_SKIP_P2P, raise_on_error=False)
    return {"gpus": len(gpus), "hbm_bytes_per_gpu": hbm_bytes, "iters": iters, "contexts_retained": keep_contexts,
            "target": label, "kernels_run_from": "the worker's own process on the host (not inside the pod)",
            "p50_s": statistics.median(lat), "p99_s": pct(lat, 0.99), "min_s": min(lat), "max_s": max(lat),
            "median_phases_ms": {k: statistics.median(p[k] for p in phases) for k in phases[0]}}


def device_free_bytes():
    """cudaMemGetInfo of every visible GPU, from a child process so that this one creates no context of its own."""
    import subprocess

    code = ("from cuda.bindings import runtime as rt\nimport json\nout=[]\n"
            "n=rt.cudaGetDeviceCount()[1]\n"
            "for i in range(n):\n    rt.cudaSetDevice(i); e,f,t=rt.cudaMemGetInfo(); out.append(int(f)); rt.cudaDeviceReset()\n"
            "print(json.dumps(out))\n")
    try:
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
        return json.loads(r.stdout.strip().splitlines()[-1])
    except Exception:
        return None


def churn_real(rounds, hbm_bytes=256 << 20, nodes_only=False, pod_like=None):
    """Config 5 on real GPUs: 1 -> N -> 1 mounted GPUs, `rounds` times, in ONE long-lived stand-in pod.  Every step: native
    mount, in-pod probe (HBM only; `nodes_only`: the device nodes opened from inside the pod, no CUDA start-up), native
    unmount, in-pod check that the nodes are gone.  Leak counters afterwards: this process's descriptors, zombies, NVML's
    process list of every GPU, the pod's devices.list against its state before the first mount, free device memory."""
    nv = {d.uuid: d for d in probe.enumerate_nvml()}
    cu = probe.enumerate_cuda()
    gpus = [(d.uuid, nv[d.uuid].minor) for d in cu]
    n = len(gpus)
    recs_all = [device.New(m, u) for u, m in gpus]
    baseline = [len(r.GetRunningProcess()) for r in recs_all]
    free0 = device_free_bytes()
    fds0 = open_fds()
    usable_s, gone_s, steps = [], [], 0
    # a resident worker keeps NVML attached (gmp_nvml_hold): the busy check of every unmount is then a query, not a
    # cold nvmlInit_v2 (~140 ms on an 8-GPU host)
    with GpuTarget(pod_like) as g, util.NvmlAttached():
        t = g.t
        pod, env = product_pod(), g.env()
        devices_list0 = t.read("devices.list")
        t_all = time.monotonic_ns()
        for _ in range(rounds):
            for k in list(range(1, n + 1)) + list(range(n - 1, 0, -1)):
                recs = recs_all[:k]
                g.restrict_cuda(gpus[:k])
                t0 = time.monotonic_ns()
                util.MountGPUs(pod, recs, env)
                if nodes_only:
                    rc, failed = util.CheckNodesInPod(pod, recs, env, raise_on_error=False)
                    assert rc == 0, (rc, failed, probe.lib().gmp_last_error())
                    usable_s.append((time.monotonic_ns() - t0) / 1e9)
                else:
                    hook = util.ProbeInPodFanout if (k >= 2 and g.pod_like) else util.ProbeInPod  # the product's hook
                    rc, rep = hook(pod, recs, g.probe_args() + ["--bytes", str(hbm_bytes), "--iters", "1",
                                                                "--dma-bytes", "1M", "--skip-p2p"], env, raise_on_error=False)
                    assert rc == 0 and rep.get("status") == 0, (rc, rep, probe.lib().gmp_last_error())
                    usable_s.append((rep["mono_usable_ns"] - t0) / 1e9)
                g.unrestrict_cuda()
                t1 = time.monotonic_ns()
                util.UnmountGPUs(pod, recs, False, env)
                rc, failed = util.VerifyGoneInPod(pod, recs, env, raise_on_error=False)
                gone_s.append((time.monotonic_ns() - t1) / 1e9)
                assert rc == 0, (rc, failed, probe.lib().gmp_last_error())
                assert t.gpu_nodes() == []
                steps += 1
        total_s = (time.monotonic_ns() - t_all) / 1e9
        devices_list1 = t.read("devices.list")
        label = g.label
    leaked_procs = [len(r.GetRunningProcess()) - b for r, b in zip(recs_all, baseline)]
    free1 = device_free_bytes()
    return {"gpus": n, "rounds": rounds, "steps": steps, "hbm_bytes_per_gpu": hbm_bytes, "target": label, "total_s": total_s,
            "per_step_check": "device nodes opened from inside the pod (no CUDA)" if nodes_only
            else "cold in-pod HBM probe (one process per GPU from 2 GPUs on, then the whole-pod device-set check)",
            "mount_to_usable": {"p50_s": statistics.median(usable_s), "p99_s": pct(usable_s, 0.99), "max_s": max(usable_s)},
            "unmount_to_verified_gone": {"p50_ms": 1e3 * statistics.median(gone_s), "p99_ms": 1e3 * pct(gone_s, 0.99)},
            "leaked_fds": open_fds() - fds0, "leaked_gpu_processes": leaked_procs,
            "devices_list_back_to_baseline": devices_list0 == devices_list1,
            "device_free_bytes_delta": None if not (free0 and free1) else [b - a for a, b in zip(free0, free1)],
            "zombies": sum(1 for p in os.listdir("/proc") if p.isdigit() and _is_zombie_child(p))}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=100)
    ap.add_argument("--gpus", default="1,2,4,8")
    ap.add_argument("--churn-rounds", type=int, default=100)
    ap.add_argument("--usable-iters", type=int, default=0, help="> 0 only on a box with real GPUs")
    ap.add_argument("--usable-bytes", type=int, default=1 << 30)
    ap.add_argument("--resident-iters", type=int, default=0,
                    help="> 0 only on a box with real GPUs: two-stage hook, kernels run from this (resident) process")
    ap.add_argument("--scratch-target", action="store_true",
                    help="round-1 stand-in (scratch dev directory, scratch cgroup files, CUDA_VISIBLE_DEVICES) instead of the "
                         "pod-like one (tmpfs on /dev, real cgroup-v1 devices group)")
    ap.add_argument("--fanout", action="store_true", help="also measure mount-to-usable with one probe process per GPU")
    ap.add_argument("--usable-sizes", default="1,2,4,8")
    ap.add_argument("--churn-real-rounds", type=int, default=0, help="> 0 only on a box with real GPUs: cold in-pod probe per step")
    ap.add_argument("--churn-nodes-rounds", type=int, default=0,
                    help="> 0 only on a box with real GPUs: nodes opened from inside the pod per step, no CUDA start-up")
    ap.add_argument("--skip-mount-bench", action="store_true", help="leave out the reference-vs-native mount/unmount timing")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    # same context-size setting gpumounter-probe gives itself (probe_main.cpp); must precede this process's cuInit
    os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "1")
    os.environ.setdefault("CUDA_DEVICE_MAX_COPY_CONNECTIONS", "1")
    why = harness.can_run()
    if why:
        print(json.dumps({"unavailable": why}))
        return 0
    os.environ.setdefault("GPUMOUNTER_NVML_LIB", os.environ.get("GPUMOUNTER_NVML_LIB", ""))
    have_gpu = os.path.exists("/dev/nvidiactl")
    if not have_gpu:
        # the unmount path asks NVML for the GPU's processes; without a driver use the test double
        fake = os.path.join(ROOT, "tests", "fake_nvml", "libfake_nvml.so")
        cfg = "/tmp/gmp_latency_fake_nvml.cfg"
        with open(cfg, "w") as f:
            f.write("".join(f"gpu {u} {m} 0000:00:00.0\n" for u, m in fake_gpus(8)))
        os.environ["GPUMOUNTER_NVML_LIB"] = fake
        os.environ["GMP_FAKE_NVML_CONFIG"] = cfg
    else:
        os.environ.pop("GPUMOUNTER_NVML_LIB", None)
    pod_like = None if not args.scratch_target else False
    try:
        from tests import harness as v2_harness  # the cgroup-v2 stand-in needs the oracle's device-program builder

        v2_ok = v2_harness.can_run_cgroup2() is None
    except Exception:
        v2_harness, v2_ok = None, False
    out = {"host_cores": os.cpu_count(), "threads_used": 1, "iters": args.iters, "mount": {}, "unmount": {},
           "unmount_nvml_busy_check": "real NVML, both arms" if have_gpu else "fake NVML in the native arm, none in the reference arm"}
    # On a GPU box the unmount path's NVML query names a UUID: use real UUIDs where they exist, made-up ones beyond
    real = []
    if have_gpu:
        nv = probe.enumerate_nvml()
        real = [(d.uuid, d.minor + 100) for d in nv]  # minors shifted: never touch a real node name in the private /dev
    for g in ([] if args.skip_mount_bench else [int(x) for x in args.gpus.split(",")]):
        gpus = (real + [x for x in fake_gpus(8)])[:g] if not have_gpu else None
        if have_gpu:
            # NVML only answers for real UUIDs; repeat them with distinct minors to model G GPUs' worth of work
            gpus = [(real[i % len(real)][0], 100 + i) for i in range(g)]
        with harness.Target() as t:
            ref_m, ref_u = bench_reference(t, gpus, args.iters, nvml_busy_check=have_gpu)
        with harness.Target() as t:
            nat_m, nat_u = bench_native(t, gpus, args.iters, batched=True)
        with harness.Target() as t:
            ser_m, ser_u = bench_native(t, gpus, args.iters, batched=False)
        out["mount"][str(g)] = {"reference_fork_exec": ref_m, "native_batched": nat_m, "native_serial": ser_m,
                                "speedup_p50": ref_m["p50_ms"] / nat_m["p50_ms"]}
        if v2_ok:
            # no reference arm: on the unified hierarchy the reference's path resolves a directory that does not exist
            with v2_harness.Target(cgroup2=True) as t:
                v2_m, v2_u = bench_native(t, gpus, args.iters, batched=True)
            out["mount"][str(g)]["native_batched_cgroup2"] = v2_m
            out.setdefault("unmount_cgroup2", {})[str(g)] = v2_u
        out["unmount"][str(g)] = {"reference_fork_exec": ref_u, "native_batched": nat_u, "native_serial": ser_u,
                                  "speedup_p50": ref_u["p50_ms"] / nat_u["p50_ms"]}
    if args.churn_rounds and not have_gpu:
        with harness.Target() as t:
            out["churn_1_8_1"] = churn(t, args.churn_rounds)
        if v2_ok:
            with v2_harness.Target(cgroup2=True) as t:
                out["churn_1_8_1_cgroup2"] = churn(t, args.churn_rounds)
    elif args.churn_rounds:
        out["churn_1_8_1"] = {"skipped": "churn uses made-up UUIDs, which a real NVML refuses; run on a driver-less box"}
    if args.usable_iters > 0 and have_gpu:
        n_vis = len(probe.enumerate_cuda())
        sizes = [int(x) for x in args.usable_sizes.split(",") if int(x) <= n_vis]
        out["mount_to_usable"] = {str(n): to_usable(args.usable_iters, args.usable_bytes, n, pod_like) for n in sizes}
        if args.fanout:
            out["mount_to_usable_fanout"] = {str(n): to_usable(args.usable_iters, args.usable_bytes, n, pod_like, fanout=True)
                                             for n in sizes}
    if args.resident_iters > 0 and have_gpu:
        n_vis = len(probe.enumerate_cuda())
        out["mount_to_usable_resident_worker"] = {
            mode: {str(n): to_usable_resident(args.resident_iters, args.usable_bytes, n, keep, pod_like=pod_like)
                   for n in (1, 2, 4, 8) if n <= n_vis}
            for mode, keep in (("contexts_released", False), ("contexts_retained", True))}
    if args.churn_real_rounds > 0 and have_gpu:
        out["churn_real"] = churn_real(args.churn_real_rounds, pod_like=pod_like)
    if args.churn_nodes_rounds > 0 and have_gpu:
        out["churn_real_nodes_only"] = churn_real(args.churn_nodes_rounds, nodes_only=True, pod_like=pod_like)
    text = json.dumps(out)
    print(text)
    if args.out:
        with open(args.out, "w") as f:
            f.write(text + "\n")
    return 0


if __name__ == "__main__":
    sys.exit(main())
