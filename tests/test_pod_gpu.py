"""The reference's own acceptance test, cluster-less, on a real GPU box (GPU only).

ref: docs/guide/QuickStart.md:43-46,63-69,87-91 - inside the pod `nvidia-smi -L` says "No devices found.", then lists the
mounted GPUs after AddGPU, then the remaining ones after RemoveGPU.  The stand-in pod here (tools/standin_pod.py,
`pod_like=True`) has a tmpfs ON /dev of its own mount namespace, populated like a container started with
NVIDIA_VISIBLE_DEVICES=none (control nodes, no /dev/nvidia<N>), and lives in a real cgroup-v1 `devices` group with a
container's whitelist.  Nothing restricts CUDA through the environment: a process in the pod reaches a GPU only through
the node (ref: pkg/util/namespace/namespace.go:167-177) and the devices.allow line (ref: pkg/util/cgroup/cgroup.go:143-155)
that the mount path under test makes.
"""

import os
import shutil
import subprocess
import sys
import time

import pytest

from gpumounter_b200 import ProbeError, _lib, device, probe, util
from gpumounter_b200._model import MountEnv, Pod
from tests import harness

pytestmark = pytest.mark.gpu


@pytest.fixture
def pod_target(monkeypatch):
    why = harness.can_run_pod_like()
    if why:
        pytest.skip(why)
    # these tests prove that the device cgroup and the pod's /dev gate CUDA, not the environment: whatever the box's
    # launcher exported must not reach the processes started in the pod (they inherit this process's environment)
    monkeypatch.delenv("CUDA_VISIBLE_DEVICES", raising=False)
    monkeypatch.delenv("NVIDIA_VISIBLE_DEVICES", raising=False)
    t = harness.Target(pod_like=True).start()
    try:
        yield t
    finally:
        try:  # the main process must leave the group before the group can be removed
            with open(os.path.join(harness.standin_pod.V1_DEVICES_ROOT, "cgroup.procs"), "w") as f:
                f.write(f"{t.pid}\n")
        except OSError:
            pass
        t.stop()


def the_pod():
    return Pod("gpu-pod", uid=harness.UID, container_id="docker://" + harness.CID)


def env_of(t):
    return MountEnv(devices_root=t.devices_root, cgroup_driver="cgroupfs")  # dev_prefix: the reference's "/dev/nvidia"


def node_gpus():
    """(uuid, minor) of every GPU of the box, NVML order (ref: collector.go:56-77)."""
    return [(d.uuid, d.minor) for d in probe.enumerate_nvml()]


def smi_in_pod(t):
    """`nvidia-smi -L` run as a process of the pod; None when the box has no nvidia-smi."""
    if shutil.which("nvidia-smi") is None:
        return None
    p = t.spawn_in_pod(["nvidia-smi", "-L"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    out, _ = p.communicate(timeout=120)
    return out


def uuids_listed(smi_output):
    return {line.split("UUID: ")[1].rstrip(")\n ") for line in smi_output.splitlines() if line.startswith("GPU ") and "UUID: " in line}


def probe_in_pod(t, recs, extra):
    return util.ProbeInPod(the_pod(), recs, extra, env_of(t), raise_on_error=False)


def assert_pod_has_no_gpu(t, recs):
    # the mounted-GPU check, asked about GPUs that are not mounted: the node is not there
    rc, rep = probe_in_pod(t, recs, ["--skip-hbm", "--skip-p2p"])
    assert rc == _lib.GMP_ERR_NODE_GPU, (rc, rep)
    assert rep["nodes_status"] == _lib.GMP_ERR_NODE_GPU and "/dev/nvidia" in rep["detail"]
    assert t.gpu_nodes() == []
    # and CUDA itself, started in the pod with nothing named: not one device (the control nodes alone get it nowhere)
    rc, rep = probe_in_pod(t, [], ["--skip-nodes", "--skip-hbm", "--skip-p2p"])
    assert rc == _lib.GMP_ERR_UNAVAILABLE, (rc, rep)
    assert rep["n_gpus"] == 0
    smi = smi_in_pod(t)
    if smi is not None:
        assert uuids_listed(smi) == set(), smi  # ref: QuickStart.md:43-46 "No devices found."


def test_mount_makes_exactly_the_mounted_gpus_usable_in_the_pod(pod_target):
    t = pod_target
    gpus = node_gpus()
    # all but one GPU where the box has several, so that "exactly the mounted set" differs from "everything on the node"
    k = len(gpus) - 1 if len(gpus) > 1 else 1
    mounted, left_out = gpus[:k], gpus[k:]
    recs = [device.New(m, u) for u, m in mounted]
    assert t.whitelist_refused == [] or all("195:" not in line for line in t.whitelist_refused), t.whitelist_refused

    # (a) before the mount: "No devices found."
    assert_pod_has_no_gpu(t, recs)

    # (b) after the mount: CUDA in the pod sees exactly the mounted UUIDs; every one of them runs and verifies kernels
    assert util.MountGPUs(the_pod(), recs, env_of(t)) == k
    assert t.gpu_nodes() == sorted(f"nvidia{m}" for _, m in mounted)
    rc, rep = probe_in_pod(t, recs, ["--bytes", "256M", "--iters", "1", "--dma-bytes", "4M", "--p2p-bytes", "64M", "--p2p-iters", "1"])
    assert rc == 0 and rep["status"] == 0, (rc, rep.get("detail"), rep)
    assert rep["device_set_status"] == 0 and rep["n_gpus"] == k
    assert {d["uuid"] for d in rep["devices"]} == {u for u, _ in mounted}
    assert {d["uuid"]: d["minor"] for d in rep["devices"]} == dict(mounted)
    assert rep["nvml_status"] == 0, "NVML in the pod must answer for the mounted GPUs (by UUID when its index walk is cut short)"
    assert all(h["status"] == 0 and h["mismatch_lanes"] == 0 and h["write_gbs"] > 0 for h in rep["hbm"])
    assert rep["mono_usable_ns"] > rep["mono_start_ns"]
    # nothing named, nothing skipped: CUDA's own enumeration in the pod is the mounted set, no more
    rc, rep = probe_in_pod(t, [], ["--skip-nodes", "--skip-hbm", "--skip-p2p", "--skip-nvml"])
    assert rc == 0 and {d["uuid"] for d in rep["devices"]} == {u for u, _ in mounted}, (rc, rep)
    smi = smi_in_pod(t)
    if smi is not None:
        assert uuids_listed(smi) == {u for u, _ in mounted}, smi  # ref: QuickStart.md:63-69
    # a GPU of the node that was NOT mounted stays out of reach
    if left_out:
        other = [device.New(m, u) for u, m in left_out[:1]]
        rc, rep = probe_in_pod(t, other, ["--skip-hbm", "--skip-p2p"])
        assert rc == _lib.GMP_ERR_NODE_GPU, (rc, rep)
        assert t.try_open(left_out[0][1]) != 0

    # (c) after the unmount: back to (a), verified from inside
    assert util.UnmountGPUs(the_pod(), recs, False, env_of(t)) == k
    rc, failed = util.VerifyGoneInPod(the_pod(), recs, env_of(t), raise_on_error=False)
    assert rc == 0, (rc, failed)
    assert_pod_has_no_gpu(t, recs)


HOLDER = (
    "import sys, time\n"
    "from cuda.bindings import runtime as rt\n"
    "assert int(rt.cudaSetDevice(0)[0]) == 0\n"
    "err, p = rt.cudaMalloc(64 << 20)\n"
    "assert int(err) == 0, err\n"
    "print('holding', flush=True)\n"
    "time.sleep(600)\n"
)


def test_unmount_refuses_a_busy_gpu_on_real_nvml(pod_target):
    """ref: pkg/util/util.go:100-109 + pkg/device/nvidia.go:58-87: a process of the pod that holds the GPU makes UnmountGPU
    answer GPUBusy (here: the real NVML's process list intersected with the real cgroup.procs); force kills it."""
    t = pod_target
    uuid, minor = node_gpus()[0]
    rec = device.New(minor, uuid)
    util.MountGPUs(the_pod(), [rec], env_of(t))
    holder = t.spawn_in_pod([sys.executable, "-c", HOLDER], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    line = holder.stdout.readline()
    assert line.strip() == "holding", line + holder.stderr.read()
    # NVML names the holder, and it is a member of the pod's cgroup: that intersection is GetPodGPUProcesses
    pids = {p.Pid for p in rec.GetRunningProcess()}
    if pids and holder.pid not in pids:
        pytest.skip(f"NVML on this box reports PIDs of another PID namespace ({pids}; the holder is {holder.pid}): "
                    "the busy check needs hostPID, as the worker has it (ref: deploy/gpu-mounter-workers.yaml:16)")
    assert holder.pid in pids, (holder.pid, pids)
    assert util.GetPodGPUProcesses(the_pod(), rec, env_of(t)) == [str(holder.pid)]
    with pytest.raises(ProbeError) as e:
        util.UnmountGPUs(the_pod(), [rec], False, env_of(t))
    assert e.value.code == _lib.GMP_ERR_BUSY
    assert t.gpu_nodes() == [f"nvidia{minor}"], "a refused unmount must not have removed anything"
    assert t.try_open(minor) == 0
    # force: permission revoked, node removed, holder signalled (ref: util.go:137-142)
    util.UnmountGPUs(the_pod(), [rec], True, env_of(t))
    assert holder.wait(timeout=30) != 0
    assert t.gpu_nodes() == []
    for _ in range(100):  # the driver releases the context a moment after the process is gone
        if holder.pid not in {p.Pid for p in rec.GetRunningProcess()}:
            break
        time.sleep(0.1)
    assert holder.pid not in {p.Pid for p in rec.GetRunningProcess()}


def test_churn_in_one_pod_leaves_nothing_behind(pod_target):
    """BASELINE configs[4] in small: 1 -> n -> 1 GPUs, 3 rounds, cold in-pod probe at every step; devices.list, the pod's
    /dev, this process's descriptors and NVML's process lists end where they began."""
    t = pod_target
    gpus = node_gpus()[:4]
    recs_all = [device.New(m, u) for u, m in gpus]
    list0 = t.read("devices.list")
    fds0 = len(os.listdir("/proc/self/fd"))
    procs0 = [len(r.GetRunningProcess()) for r in recs_all]
    fds_after_first = None
    for rnd in range(3):
        for k in list(range(1, len(gpus) + 1)) + list(range(len(gpus) - 1, 0, -1)):
            recs = recs_all[:k]
            util.MountGPUs(the_pod(), recs, env_of(t))
            rc, rep = probe_in_pod(t, recs, ["--bytes", "64M", "--iters", "1", "--dma-bytes", "1M", "--skip-p2p"])
            assert rc == 0 and rep["n_gpus"] == k, (rc, rep)
            util.UnmountGPUs(the_pod(), recs, False, env_of(t))
            rc, failed = util.VerifyGoneInPod(the_pod(), recs, env_of(t), raise_on_error=False)
            assert rc == 0 and t.gpu_nodes() == []
        if fds_after_first is None:
            fds_after_first = len(os.listdir("/proc/self/fd"))
    assert t.read("devices.list") == list0
    assert len(os.listdir("/proc/self/fd")) == fds_after_first <= fds0 + 4  # the NVML image's one-off descriptors
    assert [len(r.GetRunningProcess()) for r in recs_all] == procs0


def test_fanout_probe_one_process_per_gpu(pod_target):
    """gmp_probe_in_pod_fanout on real GPUs: every per-GPU process sees exactly ONE CUDA device (its child group hides the
    others), all of them verify, and the whole-pod process that follows sees exactly the mounted set."""
    t = pod_target
    gpus = node_gpus()
    k = len(gpus) - 1 if len(gpus) > 1 else 1
    mounted = gpus[:k]
    recs = [device.New(m, u) for u, m in mounted]
    util.MountGPUs(the_pod(), recs, env_of(t))
    rc, rep = util.ProbeInPodFanout(the_pod(), recs, ["--bytes", "256M", "--iters", "1", "--dma-bytes", "4M", "--p2p-bytes", "64M",
                                                      "--p2p-iters", "1"], env_of(t), raise_on_error=False)
    assert rc == 0 and rep["status"] == 0, (rc, rep.get("detail"), rep)
    assert [r["n_gpus"] for r in rep["reports"]] == [1] * k
    assert [r["devices"][0]["uuid"] for r in rep["reports"]] == [u for u, _ in mounted]
    assert all(h["status"] == 0 and h["mismatch_lanes"] == 0 and h["write_gbs"] > 0 for h in rep["hbm"]) and len(rep["hbm"]) == k
    assert rep["mono_usable_ns"] == max(h["mono_first_verified_ns"] for h in rep["hbm"])
    w = rep["whole_pod"]
    assert w["status"] == 0 and w["device_set_status"] == 0 and {d["uuid"] for d in w["devices"]} == {u for u, _ in mounted}
    if k >= 2:
        assert w["p2p"]["matrix_match"] == 1 and w["p2p"]["mismatch_lanes"] == 0
    assert [d for d in os.listdir(t.cgroup_dir) if d.startswith("gmp-probe-")] == []
    util.UnmountGPUs(the_pod(), recs, False, env_of(t))
    assert t.read("devices.list").count("195:") == 1  # the control node alone, as before the mount


def test_incremental_mount_expects_what_the_pod_held_before(pod_target):
    """A second AddGPU into a pod that already has a GPU: CUDA in the pod sees both; the probe of the NEW one names the old
    one with --also-mounted (visible, not probed).  Without it the set check rightly fails; so does naming a GPU the pod
    does not hold."""
    t = pod_target
    gpus = node_gpus()
    if len(gpus) < 2:
        pytest.skip("needs at least 2 GPUs")
    (u0, m0), (u1, m1) = gpus[0], gpus[1]
    first, second = device.New(m0, u0), device.New(m1, u1)
    util.MountGPUs(the_pod(), [first], env_of(t))
    util.MountGPUs(the_pod(), [second], env_of(t))
    small = ["--bytes", "64M", "--iters", "1", "--dma-bytes", "1M", "--skip-p2p"]
    rc, rep = probe_in_pod(t, [second], small + ["--also-mounted", f"{u0}:{m0}"])
    assert rc == 0 and [h["uuid"] for h in rep["hbm"]] == [u1] and rep["device_set_status"] == 0, (rc, rep)
    rc, rep = probe_in_pod(t, [second], small)
    assert rc == _lib.GMP_ERR_DEVICE_SET and "2" in rep["detail"], (rc, rep)
    if len(gpus) >= 3:
        rc, rep = probe_in_pod(t, [second], small + ["--also-mounted", f"{gpus[2][0]}:{gpus[2][1]}", "--skip-nodes"])
        assert rc == _lib.GMP_ERR_DEVICE_SET and gpus[2][0] in rep["detail"], (rc, rep)
    # the fan-out form: the per-GPU child is narrowed to the new GPU, the whole-pod step checks the full set
    rc, rep = util.ProbeInPodFanout(the_pod(), [second], small + ["--also-mounted", f"{u0}:{m0}"], env_of(t), raise_on_error=False)
    assert rc == 0 and rep["reports"][0]["n_gpus"] == 1 and rep["whole_pod"]["device_set_status"] == 0, (rc, rep)
    util.UnmountGPUs(the_pod(), [first, second], False, env_of(t))
