"""ref: pkg/util/gpu/collector/collector.go:165-194 — the kubelet PodResources client (connectToServer + ListPods).

The reference imports k8s.io/kubernetes v1.18.6 pkg/kubelet/apis/podresources/v1alpha1 (go.mod:14), which is not vendored;
the five messages and the one method are restated here from that API's published definition (parity
unpinned — there is no kubelet in this environment; tests run against a stand-in server on a unix socket):

    service PodResourcesLister { rpc List(ListPodResourcesRequest) returns (ListPodResourcesResponse) }   package v1alpha1
    ListPodResourcesResponse { repeated PodResources pod_resources = 1 }
    PodResources            { string name = 1; string namespace = 2; repeated ContainerResources containers = 3 }
    ContainerResources      { string name = 1; repeated ContainerDevices devices = 2 }
    ContainerDevices        { string resource_name = 1; repeated string device_ids = 2 }

Extension (no reference counterpart): the same service in package `v1` (Kubernetes >= 1.20, the only
one served from 1.27 on), which adds `TopologyInfo topology = 3 { repeated NUMANode nodes = 1 { int64 ID = 1 } }` to
ContainerDevices and `repeated int64 cpu_ids = 3` to ContainerResources; `ListPods(api="v1" | "auto")`.
"""

from __future__ import annotations

import os
from typing import List, Tuple

from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

SocketDir = "/var/lib/kubelet/pod-resources"   # ref: pkg/util/gpu/types.go:6
SocketPath = SocketDir + "/kubelet.sock"       # ref: types.go:7
ConnectionTimeout = 10.0                        # seconds, ref: types.go:9
LIST_METHOD = "/v1alpha1.PodResourcesLister/List"

_F = descriptor_pb2.FieldDescriptorProto


def _file(package: str) -> descriptor_pb2.FileDescriptorProto:
    fd = descriptor_pb2.FileDescriptorProto(name=f"podresources_{package}.proto", package=package, syntax="proto3")

    def msg(name, fields):
        m = fd.message_type.add(name=name)
        for fname, num, ftype, label, tname in fields:
            f = m.field.add(name=fname, number=num, type=ftype, label=label)
            if tname:
                f.type_name = f".{package}." + tname

    opt, rep = _F.LABEL_OPTIONAL, _F.LABEL_REPEATED
    msg("ListPodResourcesRequest", [])
    msg("ListPodResourcesResponse", [("pod_resources", 1, _F.TYPE_MESSAGE, rep, "PodResources")])
    msg("PodResources", [("name", 1, _F.TYPE_STRING, opt, None), ("namespace", 2, _F.TYPE_STRING, opt, None),
                         ("containers", 3, _F.TYPE_MESSAGE, rep, "ContainerResources")])
    container = [("name", 1, _F.TYPE_STRING, opt, None), ("devices", 2, _F.TYPE_MESSAGE, rep, "ContainerDevices")]
    devices = [("resource_name", 1, _F.TYPE_STRING, opt, None), ("device_ids", 2, _F.TYPE_STRING, rep, None)]
    if package == "v1":
        # what v1 adds that matters here: the NUMA node of each device; cpu_ids (3), memory (4) and
        # dynamic_resources (5) of ContainerResources are left undeclared and skipped as unknown fields
        container.append(("cpu_ids", 3, _F.TYPE_INT64, rep, None))
        devices.append(("topology", 3, _F.TYPE_MESSAGE, opt, "TopologyInfo"))
        msg("TopologyInfo", [("nodes", 1, _F.TYPE_MESSAGE, rep, "NUMANode")])
        msg("NUMANode", [("ID", 1, _F.TYPE_INT64, opt, None)])
    msg("ContainerResources", container)
    msg("ContainerDevices", devices)
    s = fd.service.add(name="PodResourcesLister")
    s.method.add(name="List", input_type=f".{package}.ListPodResourcesRequest", output_type=f".{package}.ListPodResourcesResponse")
    return fd


_pool = descriptor_pool.DescriptorPool()
_pool.Add(_file("v1alpha1"))
_pool.Add(_file("v1"))
ListPodResourcesRequest = message_factory.GetMessageClass(_pool.FindMessageTypeByName("v1alpha1.ListPodResourcesRequest"))
ListPodResourcesResponse = message_factory.GetMessageClass(_pool.FindMessageTypeByName("v1alpha1.ListPodResourcesResponse"))
# Extension, no reference counterpart: the v1 API that replaced v1alpha1 (GA in Kubernetes 1.20,
# package "v1", same socket).  Same field numbers for everything UpdateGPUStatus reads.
V1_LIST_METHOD = "/v1.PodResourcesLister/List"
V1ListPodResourcesRequest = message_factory.GetMessageClass(_pool.FindMessageTypeByName("v1.ListPodResourcesRequest"))
V1ListPodResourcesResponse = message_factory.GetMessageClass(_pool.FindMessageTypeByName("v1.ListPodResourcesResponse"))

Listing = List[Tuple[str, str, List[Tuple[str, List[Tuple[str, List[str]]]]]]]


def to_listing(resp) -> Listing:
    """The (pod, namespace, [(container, [(resource_name, [device_ids])])]) form GPUCollector.UpdateGPUStatus takes."""
    return [(p.name, p.namespace, [(c.name, [(d.resource_name, list(d.device_ids)) for d in c.devices]) for c in p.containers])
            for p in resp.pod_resources]


def numa_nodes(resp) -> dict:
    """v1 only: device id -> NUMA node ids, for the devices that report a topology."""
    out = {}
    for p in resp.pod_resources:
        for c in p.containers:
            for d in c.devices:
                if d.HasField("topology"):
                    for dev in d.device_ids:
                        out[dev] = [n.ID for n in d.topology.nodes]
    return out


def ListPods(socket_path: str = SocketPath, timeout: float = ConnectionTimeout, api: str = "v1alpha1") -> Listing:
    """ref: collector.go:92-111 (stat the socket), :165-180 (blocking dial over unix, 10 s), :182-194 (List, 10 s).

    api: "v1alpha1" (the reference's, default), "v1" (extension), or "auto" (v1, falling back to v1alpha1 when the
    kubelet answers UNIMPLEMENTED)."""
    import grpc

    if api not in ("v1alpha1", "v1", "auto"):
        raise ValueError(f"unknown PodResources API version {api!r}")
    if not os.path.exists(socket_path):
        raise FileNotFoundError(f"Can not found {socket_path}")  # collector.go:93-97
    channel = grpc.insecure_channel("unix://" + socket_path)
    try:
        try:
            grpc.channel_ready_future(channel).result(timeout=timeout)  # grpc.WithBlock()
        except grpc.FutureTimeoutError:
            raise ConnectionError(f"failure connecting to {socket_path}: context deadline exceeded")
        versions = {"v1alpha1": [(LIST_METHOD, ListPodResourcesRequest, ListPodResourcesResponse)],
                    "v1": [(V1_LIST_METHOD, V1ListPodResourcesRequest, V1ListPodResourcesResponse)]}
        versions["auto"] = versions["v1"] + versions["v1alpha1"]
        last = None
        for method, request, response in versions[api]:
            call = channel.unary_unary(method, request_serializer=lambda m: m.SerializeToString(),
                                       response_deserializer=response.FromString)
            try:
                return to_listing(call(request(), timeout=timeout))
            except grpc.RpcError as e:
                last = e
                if e.code() != grpc.StatusCode.UNIMPLEMENTED:
                    break
        raise ConnectionError(f"failure getting pod resources {last.code().name}: {last.details()}")
    finally:
        channel.close()
