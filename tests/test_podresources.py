"""The kubelet PodResources client (ref: collector.go:165-194) against a stand-in kubelet on a unix socket, feeding
GPUCollector.UpdateGPUStatus end to end.  CPU only."""

from concurrent import futures

import grpc
import pytest

from gpumounter_b200 import collector, device, podresources
from tests.conftest import FAKE_UUIDS


def serve_fake_kubelet(sock, response=None, fail=False, services=("v1alpha1",)):
    def handler(raw, context):
        if fail:
            context.abort(grpc.StatusCode.UNAVAILABLE, "kubelet restarting")
        return response

    server = grpc.server(futures.ThreadPoolExecutor(max_workers=2))
    for package in services:
        server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(f"{package}.PodResourcesLister", {
            "List": grpc.unary_unary_rpc_method_handler(handler, request_deserializer=lambda b: b,
                                                        response_serializer=lambda m: m.SerializeToString())}),))
    server.add_insecure_port("unix://" + sock)
    server.start()
    return server


def sample_response():
    r = podresources.ListPodResourcesResponse()
    p = r.pod_resources.add(name="gpu-pod", namespace="default")
    c = p.containers.add(name="main")
    c.devices.add(resource_name="nvidia.com/gpu", device_ids=[FAKE_UUIDS[0]])
    c.devices.add(resource_name="example.com/fpga", device_ids=["f0"])
    s = r.pod_resources.add(name="gpu-pod-slave-pod-0a1b2c", namespace="gpu-pool")
    s.containers.add(name="gpu-container").devices.add(resource_name="nvidia.com/gpu", device_ids=[FAKE_UUIDS[1], FAKE_UUIDS[2]])
    r.pod_resources.add(name="no-devices", namespace="kube-system").containers.add(name="c")
    return r


def test_wire_bytes_are_the_published_field_numbers():
    r = podresources.ListPodResourcesResponse()
    r.pod_resources.add(name="p", namespace="n").containers.add(name="c").devices.add(resource_name="r", device_ids=["d"])
    # field 1 (pod_resources) { 1:"p" 2:"n" 3:{ 1:"c" 2:{ 1:"r" 2:"d" } } }
    assert r.SerializeToString() == bytes.fromhex("0a13" "0a0170" "12016e" "1a0b" "0a0163" "1206" "0a0172" "120164")
    assert podresources.LIST_METHOD == "/v1alpha1.PodResourcesLister/List"


def test_list_pods_feeds_update_gpu_status(tmp_path):
    sock = str(tmp_path / "kubelet.sock")
    srv = serve_fake_kubelet(sock, sample_response())
    try:
        listing = podresources.ListPods(sock, timeout=5)
    finally:
        srv.stop(0)
    assert listing[0] == ("gpu-pod", "default", [("main", [("nvidia.com/gpu", [FAKE_UUIDS[0]]), ("example.com/fpga", ["f0"])])])
    c = collector.GPUCollector([device.New(i, FAKE_UUIDS[i]) for i in range(4)])
    c.UpdateGPUStatus(listing)
    assert [(g.PodName, g.Namespace) for g in c.GPUList] == [("gpu-pod", "default"), ("gpu-pod-slave-pod-0a1b2c", "gpu-pool"),
                                                              ("gpu-pod-slave-pod-0a1b2c", "gpu-pool"), ("", "")]
    assert [g.MinorNumber for g in c.GetPodGPUResources("gpu-pod", "default")] == [0, 1, 2]


def test_errors_are_reported_like_the_reference(tmp_path):
    with pytest.raises(FileNotFoundError, match="Can not found"):
        podresources.ListPods(str(tmp_path / "absent.sock"))
    sock = str(tmp_path / "kubelet.sock")
    srv = serve_fake_kubelet(sock, fail=True)
    try:
        with pytest.raises(ConnectionError, match="failure getting pod resources"):
            podresources.ListPods(sock, timeout=5)
    finally:
        srv.stop(0)
    dead = tmp_path / "dead.sock"
    dead.write_text("")  # exists, nobody listens
    with pytest.raises(ConnectionError, match="failure connecting to"):
        podresources.ListPods(str(dead), timeout=0.5)


def v1_response():
    r = podresources.V1ListPodResourcesResponse()
    c = r.pod_resources.add(name="gpu-pod", namespace="default").containers.add(name="main", cpu_ids=[0, 1, 64, 65])
    d = c.devices.add(resource_name="nvidia.com/gpu", device_ids=[FAKE_UUIDS[0]])
    d.topology.nodes.add(ID=1)
    c.devices.add(resource_name="nvidia.com/gpu", device_ids=[FAKE_UUIDS[3]])
    return r


def test_v1_api_is_an_opt_in_extension(tmp_path):
    """Kubernetes 1.20+ serves v1 (and from 1.27 on only v1); the reference asks for v1alpha1.  Same listing either way,
    plus the NUMA node v1 reports for a device."""
    resp = v1_response()
    assert podresources.numa_nodes(resp) == {FAKE_UUIDS[0]: [1]}
    # a v1alpha1 client reading v1 bytes sees the same pods/containers/devices: topology and cpu_ids are unknown fields
    old = podresources.ListPodResourcesResponse.FromString(resp.SerializeToString())
    assert podresources.to_listing(old) == podresources.to_listing(resp)
    sock = str(tmp_path / "kubelet.sock")
    srv = serve_fake_kubelet(sock, resp, services=("v1",))  # a modern kubelet: no v1alpha1 any more
    try:
        want = [("gpu-pod", "default", [("main", [("nvidia.com/gpu", [FAKE_UUIDS[0]]), ("nvidia.com/gpu", [FAKE_UUIDS[3]])])])]
        assert podresources.ListPods(sock, timeout=5, api="v1") == want
        assert podresources.ListPods(sock, timeout=5, api="auto") == want
        with pytest.raises(ConnectionError, match="UNIMPLEMENTED"):   # the reference's call against that kubelet
            podresources.ListPods(sock, timeout=5)
    finally:
        srv.stop(0)
    sock2 = str(tmp_path / "old.sock")
    srv = serve_fake_kubelet(sock2, sample_response(), services=("v1alpha1",))  # a 1.18 kubelet
    try:
        assert podresources.ListPods(sock2, timeout=5, api="auto") == podresources.ListPods(sock2, timeout=5)
        with pytest.raises(ConnectionError, match="UNIMPLEMENTED"):
            podresources.ListPods(sock2, timeout=5, api="v1")
    finally:
        srv.stop(0)
    with pytest.raises(ValueError):
        podresources.ListPods(sock2, api="v2")
