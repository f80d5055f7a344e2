"""ref: pkg/api/gpu-mount/api.proto — the worker's gRPC contract, built by hand (no protoc in this image).

The descriptor below is written out field by field from api.proto:4-45 and is checked in tests against the
FileDescriptorProto that protoc-gen-go embedded in the reference's api.pb.go (tests/golden/api_filedescriptor.pb).
Field numbers, enum values (note the gap at 3 in RemoveGPUResult) and the method paths are frozen.

`EXTENDED` is the same file with ADDITIVE fields only (new numbers, proto3 optional-by-default): a probe verdict on
AddGPUResponse.  An unmodified GPUMounter master parses an extended response and simply does not see the new fields;
an extended client parses an old worker's response and sees them empty.
"""

from __future__ import annotations

from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

PACKAGE = "gpu_mount"
ADD_GPU_METHOD = "/gpu_mount.AddGPUService/AddGPU"            # ref: api.pb.go:390
REMOVE_GPU_METHOD = "/gpu_mount.RemoveGPUService/RemoveGPU"   # ref: api.pb.go:462
WORKER_PORT = 1200                                             # ref: cmd/GPUMounter-worker/main.go:24

_F = descriptor_pb2.FieldDescriptorProto


def _field(msg, name, number, ftype, label=_F.LABEL_OPTIONAL, type_name=None, json_name=None):
    f = msg.field.add()
    f.name, f.number, f.type, f.label = name, number, ftype, label
    if type_name:
        f.type_name = type_name
    parts = name.split("_")
    f.json_name = json_name or (parts[0] + "".join(p.capitalize() for p in parts[1:]))
    return f


def file_descriptor_proto(extended: bool = False) -> descriptor_pb2.FileDescriptorProto:
    fd = descriptor_pb2.FileDescriptorProto()
    fd.name = "api_extended.proto" if extended else "api.proto"
    fd.package = "gpu_mount_ext" if extended else PACKAGE
    fd.syntax = "proto3"
    pkg = fd.package

    m = fd.message_type.add()  # api.proto:4-9
    m.name = "AddGPURequest"
    _field(m, "pod_name", 1, _F.TYPE_STRING)
    _field(m, "namespace", 2, _F.TYPE_STRING)
    _field(m, "gpu_num", 3, _F.TYPE_INT32)
    _field(m, "is_entire_mount", 4, _F.TYPE_BOOL)

    m = fd.message_type.add()  # api.proto:11-19
    m.name = "AddGPUResponse"
    _field(m, "add_gpu_result", 1, _F.TYPE_ENUM, type_name=f".{pkg}.AddGPUResponse.AddGPUResult")
    e = m.enum_type.add()
    e.name = "AddGPUResult"
    for name, number in (("Success", 0), ("InsufficientGPU", 1), ("PodNotFound", 2)):
        v = e.value.add()
        v.name, v.number = name, number
    if extended:  # additive only: new field numbers >= 2
        _field(m, "probe_status", 2, _F.TYPE_SINT32)      # GMP_OK / GMP_ERR_* of the post-mount probe
        _field(m, "probe_report", 3, _F.TYPE_STRING)      # the probe's JSON line
        _field(m, "mounted_uuids", 4, _F.TYPE_STRING, label=_F.LABEL_REPEATED)

    m = fd.message_type.add()  # api.proto:25-30
    m.name = "RemoveGPURequest"
    _field(m, "pod_name", 1, _F.TYPE_STRING)
    _field(m, "namespace", 2, _F.TYPE_STRING)
    _field(m, "uuids", 3, _F.TYPE_STRING, label=_F.LABEL_REPEATED)
    _field(m, "force", 4, _F.TYPE_BOOL)

    m = fd.message_type.add()  # api.proto:32-41
    m.name = "RemoveGPUResponse"
    _field(m, "remove_gpu_result", 1, _F.TYPE_ENUM, type_name=f".{pkg}.RemoveGPUResponse.RemoveGPUResult")
    e = m.enum_type.add()
    e.name = "RemoveGPUResult"
    for name, number in (("Success", 0), ("GPUBusy", 1), ("PodNotFound", 2), ("GPUNotFound", 4)):
        v = e.value.add()
        v.name, v.number = name, number

    for svc, method, req, resp in (("AddGPUService", "AddGPU", "AddGPURequest", "AddGPUResponse"),
                                   ("RemoveGPUService", "RemoveGPU", "RemoveGPURequest", "RemoveGPUResponse")):
        s = fd.service.add()  # api.proto:21-23, 43-45
        s.name = svc
        mt = s.method.add()
        mt.name, mt.input_type, mt.output_type = method, f".{pkg}.{req}", f".{pkg}.{resp}"
        mt.options.SetInParent()  # `rpc X (A) returns (B) {};` carries an empty options block
    return fd


_pool = descriptor_pool.DescriptorPool()
_pool.Add(file_descriptor_proto(False))
_pool.Add(file_descriptor_proto(True))


def _cls(full_name):
    return message_factory.GetMessageClass(_pool.FindMessageTypeByName(full_name))


AddGPURequest = _cls("gpu_mount.AddGPURequest")
AddGPUResponse = _cls("gpu_mount.AddGPUResponse")
RemoveGPURequest = _cls("gpu_mount.RemoveGPURequest")
RemoveGPUResponse = _cls("gpu_mount.RemoveGPUResponse")
ExtendedAddGPUResponse = _cls("gpu_mount_ext.AddGPUResponse")

# enum values, ref: api.pb.go:29-33,57-62
AddGPUResult_Success, AddGPUResult_InsufficientGPU, AddGPUResult_PodNotFound = 0, 1, 2
RemoveGPUResult_Success, RemoveGPUResult_GPUBusy, RemoveGPUResult_PodNotFound, RemoveGPUResult_GPUNotFound = 0, 1, 2, 4


def add_result_http(result: int, pod_name: str, node_name: str):
    """HTTP status + body the reference master writes for a worker result (ref: cmd/GPUMounter-master/main.go:103-116)."""
    if result == AddGPUResult_Success:
        return 200, "Add GPU Success\n"
    if result == AddGPUResult_InsufficientGPU:
        return 500, f"Insufficient GPU on Node: {node_name}\n"
    if result == AddGPUResult_PodNotFound:
        return 400, f"No Pod{pod_name} on Node: {node_name}\n"
    return 200, ""


def remove_result_http(result: int, pod_name: str, node_name: str, uuids):
    """ref: cmd/GPUMounter-master/main.go:206-224"""
    joined = ", ".join(uuids)
    if result == RemoveGPUResult_PodNotFound:
        return 400, f"No Pod{pod_name} on Node: {node_name}\n"
    if result == RemoveGPUResult_GPUBusy:
        return 400, f"Pod: {pod_name} has running processes on GPU: {joined}\n"
    if result == RemoveGPUResult_GPUNotFound:
        return 400, f"Invalid UUIDs: {joined}\n"
    if result == RemoveGPUResult_Success:
        return 200, "Remove GPU Success\n"
    return 200, ""
