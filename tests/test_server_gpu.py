"""BASELINE configs[1]-[3] through the reference's own wire API on real GPUs (GPU only): AddGPU / RemoveGPU over a real
gRPC channel on /gpu_mount.AddGPUService/AddGPU and /gpu_mount.RemoveGPUService/RemoveGPU (ref: pkg/api/gpu-mount/api.pb.go:390,
462) -> the worker mirror (ref: pkg/server/gpu-mount/server.go:34-179) -> native mount into a pod-like stand-in -> the
in-pod probe hook -> additive response fields.  Only the cluster is a double (slave pods, kubelet listing): there is no
cluster on any box; the GPUs, NVML, the device cgroup, the pod's /dev and the probe are real."""

import json
import os

import grpc
import pytest

from gpumounter_b200 import api, device, probe, server
from gpumounter_b200._model import MountEnv, Pod
from tests import harness
from tests.test_server import FakeCluster

pytestmark = pytest.mark.gpu


def test_add_probe_remove_over_the_wire_on_real_gpus(monkeypatch):
    why = harness.can_run_pod_like()
    if why:
        pytest.skip(why)
    monkeypatch.delenv("CUDA_VISIBLE_DEVICES", raising=False)
    nv = probe.enumerate_nvml()
    gpus = [device.New(d.minor, d.uuid) for d in nv]
    n = min(2, len(gpus))
    t = harness.Target(pod_like=True).start()
    srv = None
    try:
        env = MountEnv(devices_root=t.devices_root, cgroup_driver="cgroupfs", probe_timeout_ms=120_000)
        pod = Pod("gpu-pod", "default", uid=harness.UID, node_name="node-1", container_id="docker://" + harness.CID)
        cluster = FakeCluster([pod], [g.UUID for g in gpus], pending_polls=1)
        impl = server.GPUMountImpl(cluster, gpus=gpus, env=env, extended_response=True, fail_on_probe_error=True,
                                   probe_hook=server.in_pod_probe_hook(env, ("--bytes", "256M", "--iters", "1", "--p2p-bytes", "64M",
                                                                             "--p2p-iters", "1")))
        srv = server.serve(impl, "127.0.0.1:0")
        ch = grpc.insecure_channel(f"127.0.0.1:{srv.bound_port}")
        add_rpc = ch.unary_unary(api.ADD_GPU_METHOD, request_serializer=lambda m: m.SerializeToString(),
                                 response_deserializer=api.ExtendedAddGPUResponse.FromString)
        old_add_rpc = ch.unary_unary(api.ADD_GPU_METHOD, request_serializer=lambda m: m.SerializeToString(),
                                     response_deserializer=api.AddGPUResponse.FromString)
        rm_rpc = ch.unary_unary(api.REMOVE_GPU_METHOD, request_serializer=lambda m: m.SerializeToString(),
                                response_deserializer=api.RemoveGPUResponse.FromString)
        assert t.gpu_nodes() == []
        r = add_rpc(api.AddGPURequest(pod_name="gpu-pod", namespace="default", gpu_num=n, is_entire_mount=False), timeout=180)
        assert r.add_gpu_result == api.AddGPUResult_Success and r.probe_status == 0
        mounted = list(r.mounted_uuids)
        assert len(mounted) == n and set(mounted) <= {g.UUID for g in gpus}
        rep = json.loads(r.probe_report)
        assert rep["status"] == 0 and {h["uuid"] for h in rep["hbm"]} == set(mounted) and rep["mono_usable_ns"] > 0
        assert all(h["mismatch_lanes"] == 0 for h in rep["hbm"])
        assert len(t.gpu_nodes()) == n
        # a client built from the reference's api.proto reads the same bytes as the plain response
        if len(gpus) > n:
            plain = old_add_rpc(api.AddGPURequest(pod_name="gpu-pod", namespace="default", gpu_num=1, is_entire_mount=False), timeout=180)
            assert plain.add_gpu_result == api.AddGPUResult_Success and len(t.gpu_nodes()) == n + 1
            mounted = [g.UUID for g in impl.GetPodGPUResources("gpu-pod", "default")]
        # RemoveGPU: everything mounted goes, the pod is empty again, verified from inside
        rr = rm_rpc(api.RemoveGPURequest(pod_name="gpu-pod", namespace="default", uuids=mounted, force=False), timeout=180)
        assert rr.remove_gpu_result == api.RemoveGPUResult_Success
        assert t.gpu_nodes() == [] and not cluster.alloc
        assert t.read("devices.list").count("195:") == 1
    finally:
        if srv is not None:
            srv.stop(0)
        with open(os.path.join(harness.standin_pod.V1_DEVICES_ROOT, "cgroup.procs"), "w") as f:
            f.write(f"{t.pid}\n")
        t.stop()
