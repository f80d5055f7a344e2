"""The gRPC wire contract (CPU only): the hand-built descriptor against the FileDescriptorProto the reference's own
api.pb.go embeds (tests/golden/api_filedescriptor.pb, extracted by tools/extract_ref_descriptor.py), known-answer
wire bytes, and forward/backward compatibility of the additive probe fields."""

import os

from google.protobuf import descriptor_pb2

from gpumounter_b200 import api

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "api_filedescriptor.pb")


def test_descriptor_equals_the_one_embedded_in_the_reference():
    want = descriptor_pb2.FileDescriptorProto()
    want.ParseFromString(open(GOLDEN, "rb").read())
    got = api.file_descriptor_proto()
    assert got == want, "hand-built api.proto descriptor differs from protoc-gen-go's embedded one"
    assert got.SerializeToString(deterministic=True) == want.SerializeToString(deterministic=True)


def test_method_paths_and_port():
    assert api.ADD_GPU_METHOD == "/gpu_mount.AddGPUService/AddGPU"           # ref: api.pb.go:390
    assert api.REMOVE_GPU_METHOD == "/gpu_mount.RemoveGPUService/RemoveGPU"  # ref: api.pb.go:462
    assert api.WORKER_PORT == 1200                                            # ref: cmd/GPUMounter-worker/main.go:24


def test_known_answer_wire_bytes():
    # proto3 encoding worked out by hand: tag = (field << 3) | wire_type
    req = api.AddGPURequest(pod_name="p", namespace="d", gpu_num=4, is_entire_mount=True)
    assert req.SerializeToString() == bytes([0x0A, 1, ord("p"), 0x12, 1, ord("d"), 0x18, 4, 0x20, 1])
    assert api.AddGPUResponse(add_gpu_result=0).SerializeToString() == b""          # Success is the zero value
    assert api.AddGPUResponse(add_gpu_result=1).SerializeToString() == b"\x08\x01"  # InsufficientGPU
    assert api.AddGPUResponse(add_gpu_result=2).SerializeToString() == b"\x08\x02"  # PodNotFound
    rm = api.RemoveGPURequest(pod_name="p", namespace="d", uuids=["u1", "u2"], force=True)
    assert rm.SerializeToString() == b"\x0a\x01p\x12\x01d\x1a\x02u1\x1a\x02u2\x20\x01"
    assert api.RemoveGPUResponse(remove_gpu_result=4).SerializeToString() == b"\x08\x04"  # GPUNotFound, the value after the gap
    assert api.RemoveGPURequest.FromString(b"\x1a\x02u1").uuids == ["u1"]


def test_enum_values_and_the_gap_at_three():
    e = api.RemoveGPUResponse.DESCRIPTOR.enum_types_by_name["RemoveGPUResult"]
    assert {v.name: v.number for v in e.values} == {"Success": 0, "GPUBusy": 1, "PodNotFound": 2, "GPUNotFound": 4}
    assert 3 not in e.values_by_number
    a = api.AddGPUResponse.DESCRIPTOR.enum_types_by_name["AddGPUResult"]
    assert {v.name: v.number for v in a.values} == {"Success": 0, "InsufficientGPU": 1, "PodNotFound": 2}


def test_additive_fields_are_invisible_to_old_peers_and_empty_from_old_workers():
    ext = api.ExtendedAddGPUResponse(add_gpu_result=0, probe_status=-7, probe_report='{"status":-7}', mounted_uuids=["GPU-a"])
    old_view = api.AddGPUResponse.FromString(ext.SerializeToString())   # an unmodified master parsing a new worker's reply
    assert old_view.add_gpu_result == api.AddGPUResult_Success
    # and it re-serialises what it did not understand untouched (proto3 keeps unknown fields)
    assert api.ExtendedAddGPUResponse.FromString(old_view.SerializeToString()).probe_report == '{"status":-7}'
    new_view = api.ExtendedAddGPUResponse.FromString(api.AddGPUResponse(add_gpu_result=1).SerializeToString())
    assert (new_view.add_gpu_result, new_view.probe_status, new_view.probe_report, list(new_view.mounted_uuids)) == (1, 0, "", [])
    # nothing existing was renumbered
    base = {f.name: (f.number, f.type) for f in api.AddGPUResponse.DESCRIPTOR.fields}
    extd = {f.name: (f.number, f.type) for f in api.ExtendedAddGPUResponse.DESCRIPTOR.fields}
    assert all(extd[k] == v for k, v in base.items()) and min(n for k, (n, _) in extd.items() if k not in base) >= 2


def test_master_http_mapping_matches_the_oracle():
    from oracle import gmo_ref_ctypes as ref

    for result in (0, 1, 2, 3):
        assert api.add_result_http(result, "p", "n1") == ref.add_result_http(result, "p", "n1")
    for result in (0, 1, 2, 3, 4):
        assert api.remove_result_http(result, "p", "n1", ["u1", "u2"]) == ref.remove_result_http(result, "p", "n1", ["u1", "u2"])
