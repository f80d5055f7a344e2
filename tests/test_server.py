"""The worker's two RPC handlers (ref: pkg/server/gpu-mount/server.go) driven with a fake cluster and a cluster-less
target, in process and through a real gRPC channel on the reference's method paths.  CPU only."""

import grpc
import pytest

from gpumounter_b200 import ProbeError

from gpumounter_b200 import api, device, server
from gpumounter_b200._model import MountEnv, Pod
from tests import harness
from tests.conftest import FAKE_UUIDS

reason = harness.can_run()
pytestmark = pytest.mark.skipif(reason is not None, reason=reason or "")


class FakeCluster(server.Cluster):
    """Eight GPUs on one node; slave pods take free GPUs, stay Pending for `pending_polls` status reads, are
    Unschedulable when the node has too few free GPUs."""

    def __init__(self, pods, uuids, pending_polls=2):
        self.pods = {(p.namespace, p.name): p for p in pods}
        self.free = list(uuids)
        self.alloc = {}          # slave pod name -> uuids
        self.native = {}         # (ns, pod) -> uuids the pod owns natively
        self.polls = {}
        self.pending_polls = pending_polls
        self.deleted = []        # (name, grace)
        self.unschedulable = set()
        self.fail_get = False
        self.linger = {}         # name -> status reads before a deleted pod disappears

    def get_pod(self, namespace, name):
        if self.fail_get:
            raise server.ClusterError("apiserver down")
        try:
            return self.pods[(namespace, name)]
        except KeyError:
            raise server.NotFound(name)

    def create_slave_pod(self, name, owner, gpu_num):
        assert name.startswith(owner.name + "-slave-pod-") and len(name) == len(owner.name) + len("-slave-pod-") + 6
        if gpu_num <= len(self.free):
            self.alloc[name] = [self.free.pop(0) for _ in range(gpu_num)]
        else:
            self.unschedulable.add(name)
        self.polls[name] = 0

    def pod_status(self, namespace, name):
        assert namespace == "gpu-pool"
        if name in self.linger:
            self.linger[name] -= 1
            if self.linger[name] <= 0:
                del self.linger[name]
            return server.PodStatus("Running")
        if name in self.unschedulable:
            return server.PodStatus("Pending", "Unschedulable")
        if name not in self.alloc:
            raise server.NotFound(name)
        self.polls[name] += 1
        if self.polls[name] <= self.pending_polls:
            return server.PodStatus("Pending", None if self.polls[name] == 1 else "ContainersNotReady")
        return server.PodStatus("Running")

    def delete_pod(self, namespace, name, grace_period_seconds=None):
        self.deleted.append((name, grace_period_seconds))
        self.unschedulable.discard(name)
        if name in self.alloc:
            self.free.extend(self.alloc.pop(name))
            if grace_period_seconds is None:
                self.linger[name] = 2  # default grace: the pod is still there for a while

    def list_pod_resources(self):
        out = [(name, "gpu-pool", [("gpu-container", [("nvidia.com/gpu", list(u))])]) for name, u in self.alloc.items()]
        out += [(pod, ns, [("main", [("nvidia.com/gpu", list(u))])]) for (ns, pod), u in self.native.items()]
        return out


def target_pod():
    return Pod("gpu-pod", "default", uid=harness.UID, node_name="node-1", container_id="docker://" + harness.CID)


@pytest.fixture
def world(fake_nvml):
    cfg = fake_nvml()
    gpus = [device.New(m, u) for u, m, _ in cfg]
    with harness.Target() as t:
        cluster = FakeCluster([target_pod()], [u for u, _, _ in cfg])
        env = MountEnv(devices_root=t.devices_root, dev_prefix=t.dev_prefix, cgroup_driver="cgroupfs")
        probes = []

        def hook(pod, mounted):
            probes.append([g.UUID for g in mounted])
            return 0, {"n_gpus": len(mounted)}

        impl = server.GPUMountImpl(cluster, gpus=gpus, env=env, probe_hook=hook)
        yield impl, cluster, t, fake_nvml, probes


def add(pod="gpu-pod", ns="default", n=1, entire=False):
    return api.AddGPURequest(pod_name=pod, namespace=ns, gpu_num=n, is_entire_mount=entire)


def test_single_mount_then_policy(world):
    impl, cluster, t, _, probes = world
    assert impl.AddGPU(add(n=2)).add_gpu_result == api.AddGPUResult_Success
    assert len(cluster.alloc) == 2 and all(len(u) == 1 for u in cluster.alloc.values())     # one slave pod per GPU
    assert t.nodes() == ["nvidia2", "nvidia3"]                                               # minors of NVML index 0,1
    assert probes == [[FAKE_UUIDS[0], FAKE_UUIDS[1]]] and impl.last_probe_report["status"] == 0
    assert impl.GetMountType(target_pod()) == "single-mount"
    with pytest.raises(server.RpcError, match="^FailedCreated$"):                            # server.go:57-59
        impl.AddGPU(add(n=2, entire=True))
    assert impl.AddGPU(add(n=1)).add_gpu_result == api.AddGPUResult_Success                  # more single mounts are fine
    assert t.nodes() == ["nvidia1", "nvidia2", "nvidia3"]


def test_entire_mount_uses_one_slave_pod_and_blocks_further_mounts(world):
    impl, cluster, t, _, _ = world
    assert impl.AddGPU(add(n=4, entire=True)).add_gpu_result == api.AddGPUResult_Success
    assert [len(u) for u in cluster.alloc.values()] == [4]
    assert len(t.nodes()) == 4 and impl.GetMountType(target_pod()) == "entire-mount"
    for entire in (False, True):
        with pytest.raises(server.RpcError, match="^FailedCreated$"):
            impl.AddGPU(add(n=1, entire=entire))


def test_pod_not_found_insufficient_and_internal_errors(world):
    impl, cluster, t, _, _ = world
    assert impl.AddGPU(add(pod="ghost")).add_gpu_result == api.AddGPUResult_PodNotFound
    assert impl.RemoveGPU(api.RemoveGPURequest(pod_name="ghost", namespace="default")).remove_gpu_result == \
        api.RemoveGPUResult_PodNotFound
    assert impl.AddGPU(add(n=9)).add_gpu_result == api.AddGPUResult_InsufficientGPU          # 8 free, 9 asked: ninth pod unschedulable
    assert sorted(g for _, g in cluster.deleted) == [0] * 9 and not cluster.alloc            # all nine recycled with grace 0
    assert t.nodes() == [] and len(cluster.free) == 8
    cluster.fail_get = True
    with pytest.raises(server.RpcError, match="^Service Internal Error $"):
        impl.AddGPU(add())
    cluster.fail_get = False
    with pytest.raises(server.RpcError, match="^Service Internal Error $"):                  # Go would divide by zero here
        impl.AddGPU(add(n=0, entire=True))
    assert impl.AddGPU(add(n=0)).add_gpu_result == api.AddGPUResult_Success                  # zero single mounts: a no-op success


def test_mount_failure_rolls_back_slave_pods_but_not_nodes(world):
    impl, cluster, t, _, _ = world
    assert impl.AddGPU(add(n=1)).add_gpu_result == api.AddGPUResult_Success
    # make the next GPU's node exist already so that its mknod fails
    import os
    os.mknod(f"/proc/{t.pid}/root{t.devdir}/nvidia2", 0o020666, os.makedev(195, 2))
    with pytest.raises(server.RpcError, match="^Service Internal Error $"):
        impl.AddGPU(add(n=1))
    assert len(cluster.alloc) == 1 and cluster.deleted[-1][1] == 0                           # server.go:86-91
    assert "nvidia3" in t.nodes()                                                            # earlier mount untouched


def test_partial_mount_failure_revokes_what_was_granted_before_the_gpus_are_freed(world):
    """The batched mount writes ALL n devices.allow lines before the first mknod.  When it fails part-way the rollback
    must not hand GPUs back to the scheduler that this pod can still open: lines revoked, nodes made so far removed.
    revoke_on_mount_failure=False is the reference's rollback to the letter (slave pods only, server.go:86-91)."""
    import os

    impl, cluster, t, _, _ = world
    # three GPUs asked; the SECOND one's node exists already, so mknod fails after the first node was made
    os.mknod(f"/proc/{t.pid}/root{t.devdir}/nvidia2", 0o020666, os.makedev(195, 2))
    with pytest.raises(server.RpcError, match="^Service Internal Error $"):
        impl.AddGPU(add(n=3))
    assert not cluster.alloc and t.nodes() == []                       # nothing left mounted, nothing left allocated
    assert (t.read("devices.deny") or "").strip() != ""                # the grants were taken back
    impl.revoke_on_mount_failure = False
    for minor in range(8):                                             # whichever GPUs come next, their first mknod fails
        os.mknod(f"/proc/{t.pid}/root{t.devdir}/nvidia{minor}", 0o020666, os.makedev(195, minor))
    deny_before = t.read("devices.deny")
    with pytest.raises(server.RpcError, match="^Service Internal Error $"):
        impl.AddGPU(add(n=3))
    assert not cluster.alloc and t.read("devices.deny") == deny_before  # the reference's rollback: slave pods only


def test_probe_verdict_policy(world):
    """A failing or raising probe hook never escapes as a crashed RPC; with fail_on_probe_error it undoes the mount."""
    impl, cluster, t, _, _ = world

    def failing_hook(pod, gpus):
        return -7, {"detail": "mismatch"}

    def raising_hook(pod, gpus):
        raise ProbeError(-27, "in-pod probe did not finish", "ProbeInPod")

    impl.probe_hook = failing_hook
    assert impl.AddGPU(add(n=1)).add_gpu_result == api.AddGPUResult_Success      # default policy: reported, not enforced
    assert impl.last_probe_report["status"] == -7 and len(t.nodes()) == 1
    impl.probe_hook = raising_hook
    assert impl.AddGPU(add(n=1)).add_gpu_result == api.AddGPUResult_Success
    assert impl.last_probe_report["status"] == -27 and "did not finish" in impl.last_probe_report["detail"]
    assert len(t.nodes()) == 2 and len(cluster.alloc) == 2
    impl.fail_on_probe_error = True
    with pytest.raises(server.RpcError, match="^Service Internal Error $"):
        impl.AddGPU(add(n=2))
    assert len(t.nodes()) == 2 and len(cluster.alloc) == 2                       # the two new GPUs are gone again, the old stay


def test_the_products_hook_reports_the_pods_view_and_never_raises(world):
    """server.in_pod_probe_hook on a box without a driver: stage 1 (nodes opened from inside the pod) already fails, with the
    node named; the verdict travels in last_probe_report, and with fail_on_probe_error the mount is undone."""
    impl, cluster, t, _, _ = world
    impl.probe_hook = server.in_pod_probe_hook(impl.env)
    assert impl.AddGPU(add(n=1)).add_gpu_result == api.AddGPUResult_Success
    rep = impl.last_probe_report
    assert rep["status"] == -18 and rep["stage"] == "nodes" and "nvidiactl" in rep["detail"]      # GMP_ERR_NODE_CTL
    impl.fail_on_probe_error = True
    with pytest.raises(server.RpcError, match="^Service Internal Error $"):
        impl.AddGPU(add(n=1))
    assert len(t.nodes()) == 1 and len(cluster.alloc) == 1


def test_hook_is_told_what_the_pod_held_before(world):
    """A second AddGPU mounts one more GPU into a pod that already has two: the in-pod probe must expect CUDA to see all
    three and probe only the new one."""
    impl, cluster, t, _, _ = world
    seen = []

    def hook(pod, gpus, already):
        seen.append(([g.UUID for g in gpus], sorted(g.UUID for g in already)))
        return 0, {}

    impl.probe_hook = hook
    impl.AddGPU(add(n=2))
    impl.AddGPU(add(n=1))
    assert seen == [([FAKE_UUIDS[0], FAKE_UUIDS[1]], []), ([FAKE_UUIDS[2]], sorted([FAKE_UUIDS[0], FAKE_UUIDS[1]]))]


def test_status_refresh_and_query_are_one_critical_section(world):
    """Concurrent handlers (gRPC serves them from a thread pool) never see the GPU list half reset."""
    import threading

    impl, cluster, t, _, _ = world
    assert impl.AddGPU(add(n=2)).add_gpu_result == api.AddGPUResult_Success
    wrong, stop = [], threading.Event()

    def reader():
        while not stop.is_set():
            if impl.GetMountType(target_pod()) != "single-mount" or len(impl.GetPodGPUResources("gpu-pod", "default")) != 2:
                wrong.append(1)

    threads = [threading.Thread(target=reader) for _ in range(4)]
    for th in threads:
        th.start()
    for _ in range(300):
        impl.GetRemoveGPU(target_pod(), [])
    stop.set()
    for th in threads:
        th.join()
    assert not wrong


def remove(uuids, force=False):
    return api.RemoveGPURequest(pod_name="gpu-pod", namespace="default", uuids=uuids, force=force)


def test_remove_paths(world):
    impl, cluster, t, fake_nvml, _ = world
    assert impl.AddGPU(add(n=2)).add_gpu_result == api.AddGPUResult_Success
    assert impl.RemoveGPU(remove(["GPU-unknown"])).remove_gpu_result == api.RemoveGPUResult_GPUNotFound
    assert impl.RemoveGPU(remove([])).remove_gpu_result == api.RemoveGPUResult_GPUNotFound
    member = t.spawn_member()
    fake_nvml(procs=[(FAKE_UUIDS[0], member.pid, 1 << 20, "compute")])
    assert impl.RemoveGPU(remove([FAKE_UUIDS[0]])).remove_gpu_result == api.RemoveGPUResult_GPUBusy
    assert t.nodes() == ["nvidia2", "nvidia3"] and len(cluster.alloc) == 2
    assert impl.RemoveGPU(remove([FAKE_UUIDS[1]])).remove_gpu_result == api.RemoveGPUResult_Success   # the idle one
    assert t.nodes() == ["nvidia3"] and len(cluster.alloc) == 1
    assert cluster.deleted[-1][1] is None                                                     # default grace (allocator.go:136)
    assert impl.RemoveGPU(remove([FAKE_UUIDS[0]], force=True)).remove_gpu_result == api.RemoveGPUResult_Success
    assert member.wait(timeout=5) == -15 and t.nodes() == [] and not cluster.alloc
    assert impl.GetMountType(target_pod()) == "no-mount"


def test_entire_mount_removal_needs_the_right_count(world):
    impl, cluster, t, _, _ = world
    impl.AddGPU(add(n=3, entire=True))
    assert impl.RemoveGPU(remove([FAKE_UUIDS[0]])).remove_gpu_result == api.RemoveGPUResult_GPUNotFound  # allocator.go:121-123
    assert impl.RemoveGPU(remove(["x", "y", "z"])).remove_gpu_result == api.RemoveGPUResult_Success      # values irrelevant
    assert t.nodes() == [] and not cluster.alloc


def test_over_the_wire_on_the_reference_method_paths(world):
    impl, cluster, t, _, _ = world
    srv = server.serve(impl, "127.0.0.1:0")
    try:
        ch = grpc.insecure_channel(f"127.0.0.1:{srv.bound_port}")   # ref: grpc.Dial(workerIP:1200, WithInsecure)
        add_rpc = ch.unary_unary(api.ADD_GPU_METHOD, request_serializer=lambda m: m.SerializeToString(),
                                 response_deserializer=api.AddGPUResponse.FromString)
        rm_rpc = ch.unary_unary(api.REMOVE_GPU_METHOD, request_serializer=lambda m: m.SerializeToString(),
                                response_deserializer=api.RemoveGPUResponse.FromString)
        assert add_rpc(add(n=2), timeout=30).add_gpu_result == api.AddGPUResult_Success
        assert add_rpc(add(pod="ghost"), timeout=30).add_gpu_result == api.AddGPUResult_PodNotFound
        with pytest.raises(grpc.RpcError) as e:
            add_rpc(add(n=1, entire=True), timeout=30)
        assert e.value.code() == grpc.StatusCode.UNKNOWN and e.value.details() == "FailedCreated"
        assert rm_rpc(remove([FAKE_UUIDS[0], FAKE_UUIDS[1]]), timeout=30).remove_gpu_result == api.RemoveGPUResult_Success
        assert t.nodes() == []
        # raw bytes on the wire are what a Go client would send
        raw = ch.unary_unary(api.ADD_GPU_METHOD)
        assert raw(b"\x0a\x05ghost\x12\x07default\x18\x01", timeout=30) == b"\x08\x02"
        ch.close()
    finally:
        srv.stop(0)


@pytest.mark.skipif(harness.can_run_cgroup2() is not None, reason="no writable cgroup2 mount / bpf(2) here")
def test_add_and_remove_on_a_containerd_pod_on_the_unified_hierarchy(fake_nvml):
    """The same handlers with the two mount-path extensions on (header section 7): a containerd container id and a real
    cgroup-v2 directory whose device program the kernel enforces."""
    import errno

    from gpumounter_b200 import cgroup2
    from gpumounter_b200._model import EXT_CGROUP_V2, EXT_CRI_RUNTIMES

    cfg = fake_nvml()
    gpus = [device.New(m, u) for u, m, _ in cfg]
    pod = Pod("gpu-pod", "default", uid=harness.UID, node_name="node-1", container_id="containerd://" + harness.CID)
    with harness.Target(cgroup2=True) as t:
        cluster = FakeCluster([pod], [u for u, _, _ in cfg])
        env = MountEnv(cgroup2_root=t.devices_root, dev_prefix=t.dev_prefix, cgroup_driver="cgroupfs",
                       extensions=EXT_CGROUP_V2 | EXT_CRI_RUNTIMES)
        impl = server.GPUMountImpl(cluster, gpus=gpus, env=env, probe_hook=lambda p, m: (0, {"n_gpus": len(m)}))
        baseline = [cgroup2.read_program(i) for i in cgroup2.device_programs(t.cgroup_dir)[0]]
        assert impl.AddGPU(add(n=2)).add_gpu_result == api.AddGPUResult_Success
        assert t.nodes() == ["nvidia2", "nvidia3"]
        assert t.try_open(2) in (errno.ENXIO, errno.ENODEV) and t.try_open(3) in (errno.ENXIO, errno.ENODEV)
        member = t.spawn_member()
        fake_nvml(procs=[(FAKE_UUIDS[0], member.pid, 1 << 20, "compute")])
        assert impl.RemoveGPU(remove([FAKE_UUIDS[0]])).remove_gpu_result == api.RemoveGPUResult_GPUBusy
        assert impl.RemoveGPU(remove([FAKE_UUIDS[0], FAKE_UUIDS[1]], force=True)).remove_gpu_result == api.RemoveGPUResult_Success
        assert member.wait(timeout=5) == -15 and t.nodes() == [] and not cluster.alloc
        assert [cgroup2.read_program(i) for i in cgroup2.device_programs(t.cgroup_dir)[0]] == baseline


def test_a_cluster_with_a_watch_is_woken_by_events_instead_of_polling(fake_nvml):
    """Extension: where the reference spins on Get (allocator.go:246,295), a cluster that can watch
    blocks the handler until a pod event arrives.  The slave pod here becomes Running 50 ms after creation; the handler
    reads its status exactly twice (Pending, then Running after the one event) and never sleeps on its own."""
    import threading
    import time

    cfg = fake_nvml()
    gpus = [device.New(m, u) for u, m, _ in cfg]

    class WatchingCluster(FakeCluster):
        def __init__(self, *a):
            super().__init__(*a, pending_polls=0)
            self.event = threading.Event()
            self.running = set()
            self.status_reads = 0
            self.waits = 0

        def create_slave_pod(self, name, owner, gpu_num):
            super().create_slave_pod(name, owner, gpu_num)

            def start():
                time.sleep(0.05)
                self.running.add(name)
                self.event.set()

            threading.Thread(target=start, daemon=True).start()

        def pod_status(self, namespace, name):
            self.status_reads += 1
            if name in self.alloc and name not in self.running:
                return server.PodStatus("Pending", None)
            return super().pod_status(namespace, name)

        def wait_for_pod_event(self, namespace, names, timeout_s):
            assert namespace == "gpu-pool" and len(names) == 1
            self.waits += 1
            self.event.wait(timeout_s)
            self.event.clear()
            return True

    with harness.Target() as t:
        cluster = WatchingCluster([target_pod()], [u for u, _, _ in cfg])
        env = MountEnv(devices_root=t.devices_root, dev_prefix=t.dev_prefix, cgroup_driver="cgroupfs")
        impl = server.GPUMountImpl(cluster, gpus=gpus, env=env, poll_interval_s=30.0)   # a sleep would blow the timing
        t0 = time.monotonic()
        assert impl.AddGPU(add(n=1)).add_gpu_result == api.AddGPUResult_Success
        assert time.monotonic() - t0 < 5.0
        assert cluster.waits == 1 and cluster.status_reads == 2
        assert len(t.nodes()) == 1


def test_remove_gpu_attaches_nvml_once(world):
    """RemoveGPU asks NVML about every GPU twice (server.go:139-153, then util.go:100-109).  One attach covers it all."""
    import ctypes
    import os

    impl, cluster, t, fake_nvml, _ = world
    assert impl.AddGPU(add(n=4)).add_gpu_result == api.AddGPUResult_Success
    cold = ctypes.CDLL(os.environ["GPUMOUNTER_NVML_LIB"]).fake_nvml_cold_inits
    cold.restype = ctypes.c_uint
    before = cold()
    assert impl.RemoveGPU(remove(FAKE_UUIDS[:4])).remove_gpu_result == api.RemoveGPUResult_Success
    assert cold() - before == 1 and t.nodes() == []
