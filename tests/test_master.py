"""The REST front end (ref: cmd/GPUMounter-master/main.go) over a real HTTP socket and a real gRPC channel to the worker
mirror: the reference's routes, codes and bodies, plus the opt-in pass-through of the additive probe fields
(SURVEY.md 8f rank 3).  CPU only."""

import http.client
import json
import urllib.parse

import pytest

from gpumounter_b200 import api, device, master, server
from gpumounter_b200._model import MountEnv, Pod
from oracle import gmo_ref_ctypes as ref
from tests import harness
from tests.conftest import FAKE_UUIDS
from tests.test_server import FakeCluster, target_pod

pytestmark = pytest.mark.skipif(harness.can_run() is not None, reason=str(harness.can_run()))


class FakeMasterCluster(master.MasterCluster):
    def __init__(self, pods, workers):
        self.pods, self.workers, self.fail = {(p.namespace, p.name): p for p in pods}, workers, False

    def get_pod(self, namespace, name):
        if self.fail:
            raise server.ClusterError("apiserver down")
        try:
            return self.pods[(namespace, name)]
        except KeyError:
            raise server.NotFound(name)

    def worker_addresses(self):
        return self.workers


@pytest.fixture
def stack(fake_nvml):
    cfg = fake_nvml()
    gpus = [device.New(m, u) for u, m, _ in cfg]
    with harness.Target() as t:
        cluster = FakeCluster([target_pod()], [u for u, _, _ in cfg])
        env = MountEnv(devices_root=t.devices_root, dev_prefix=t.dev_prefix, cgroup_driver="cgroupfs")
        hook = lambda pod, mounted: (0, {"n_gpus": len(mounted), "verdict": "ok"})  # noqa: E731
        impl = server.GPUMountImpl(cluster, gpus=gpus, env=env, probe_hook=hook, extended_response=True)
        worker = server.serve(impl, "127.0.0.1:0")
        elsewhere = Pod("lonely", "default", uid=harness.UID, node_name="node-9", container_id="docker://x")
        mc = FakeMasterCluster([target_pod(), elsewhere], {"node-1": f"127.0.0.1:{worker.bound_port}"})
        httpd = master.serve_http(master.Master(mc), "127.0.0.1", 0)
        try:
            yield httpd.server_address[1], cluster, t, mc
        finally:
            httpd.shutdown()
            worker.stop(0)


def call(port, method, path, form=None):
    conn = http.client.HTTPConnection("127.0.0.1", port, timeout=30)
    body, headers = None, {}
    if form is not None:
        body = urllib.parse.urlencode(form, doseq=True)
        headers["Content-Type"] = "application/x-www-form-urlencoded"
    conn.request(method, path, body=body, headers=headers)
    r = conn.getresponse()
    out = (r.status, r.read().decode(), dict(r.getheaders()))
    conn.close()
    return out


ADD = "/addgpu/namespace/default/pod/gpu-pod/gpu/{n}/isEntireMount/{e}"
REMOVE = "/removegpu/namespace/default/pod/gpu-pod/force/{f}"


def test_reference_routes_codes_and_bodies(stack):
    port, cluster, t, mc = stack
    assert call(port, "GET", "/")[:2] == (200, "This is gpu mounter api!\n")
    # parameter errors, main.go:31-44
    assert call(port, "GET", ADD.format(n="two", e="false"))[:2] == (400, "Invalid param gpuNum: two\n")
    assert call(port, "GET", ADD.format(n="3000000000", e="false"))[:2] == (400, "Invalid param gpuNum: 3000000000\n")
    assert call(port, "GET", ADD.format(n="1", e="yes"))[:2] == (400, "Invalid param isEntireMount: yes(should be true or false)\n")
    # pod lookup, main.go:52-65
    assert call(port, "GET", "/addgpu/namespace/default/pod/ghost/gpu/1/isEntireMount/false")[:2] == \
        (404, "No pod: ghost in namespace: default\n")
    assert call(port, "GET", "/addgpu/namespace/default/pod/lonely/gpu/1/isEntireMount/false")[:2] == (500, "Service Internal Error\n")
    # success and the worker's results, bodies as the oracle's restatement of main.go:103-116
    status, body, headers = call(port, "GET", ADD.format(n=2, e="false"))
    assert (status, body) == ref.add_result_http(api.AddGPUResult_Success, "gpu-pod", "node-1") == (200, "Add GPU Success\n")
    assert len(t.nodes()) == 2
    status, body, headers = call(port, "GET", ADD.format(n=7, e="T"))     # ParseBool takes T; entire mount on a mounted pod
    assert (status, body) == (500, "Service Internal Error\n")            # worker: FailedCreated -> gRPC error -> 500
    assert headers["X-Content-Type-Options"] == "nosniff"                 # http.Error's headers
    cluster.free = []                                                      # nothing left on the node
    assert call(port, "GET", ADD.format(n=1, e="0"))[:2] == ref.add_result_http(api.AddGPUResult_InsufficientGPU, "gpu-pod", "node-1")
    # removal, main.go:118-224
    assert call(port, "POST", REMOVE.format(f="0"))[:2] == (400, "Invalid parameter\n")
    assert call(port, "POST", REMOVE.format(f="maybe"), {"uuids": ["x"]})[:2] == \
        (400, "Invalid parameter force: maybe(should be true or false)\n")
    assert call(port, "POST", REMOVE.format(f="0"), {"uuids": ["GPU-a", "GPU-b"]})[:2] == \
        ref.remove_result_http(api.RemoveGPUResult_GPUNotFound, "gpu-pod", "node-1", ["GPU-a", "GPU-b"])
    member = t.spawn_member()
    fake = [(FAKE_UUIDS[0], member.pid, 1 << 20, "compute")]
    import os
    with open(os.environ["GMP_FAKE_NVML_CONFIG"], "a") as f:
        f.write("".join(f"proc {u} {pid} {mem} {kind}\n" for u, pid, mem, kind in fake))
    assert call(port, "POST", REMOVE.format(f="false"), {"uuids": [FAKE_UUIDS[0]]})[:2] == \
        ref.remove_result_http(api.RemoveGPUResult_GPUBusy, "gpu-pod", "node-1", [FAKE_UUIDS[0]])
    assert call(port, "POST", REMOVE.format(f="1"), {"uuids": [FAKE_UUIDS[0], FAKE_UUIDS[1]]})[:2] == (200, "Remove GPU Success\n")
    assert member.wait(timeout=5) == -15 and t.nodes() == []
    # routing
    assert call(port, "GET", "/nope")[:2] == (404, "404 page not found\n")
    assert call(port, "POST", ADD.format(n=1, e="false"), {})[0] == 405 and call(port, "GET", REMOVE.format(f="0"))[0] == 405
    mc.fail = True
    assert call(port, "GET", ADD.format(n=1, e="false"))[:2] == (500, "apiserver down\n")


def test_probe_verdict_passes_through_only_when_asked(stack):
    port, cluster, t, _ = stack
    plain = call(port, "GET", ADD.format(n=1, e="false"))
    assert plain[:2] == (200, "Add GPU Success\n")                        # old clients: not a byte of difference
    status, body, headers = call(port, "GET", ADD.format(n=2, e="false") + "?report=json")
    assert status == 200 and headers["Content-Type"] == "application/json"
    doc = json.loads(body)
    assert doc["result"] == "Success" and doc["probe_status"] == 0 and doc["probe"] == {"n_gpus": 2, "verdict": "ok"}
    assert doc["mounted_uuids"] == [FAKE_UUIDS[1], FAKE_UUIDS[2]]
    # failures keep the reference's answer whether or not the report was asked for
    assert call(port, "GET", "/addgpu/namespace/default/pod/ghost/gpu/1/isEntireMount/false?report=json")[:2] == \
        (404, "No pod: ghost in namespace: default\n")


def test_master_against_a_worker_built_from_the_reference_proto(fake_nvml):
    """A worker that answers with the reference's plain AddGPUResponse (no fields 2-4): `?report=json` still works and
    says so with nulls."""
    cfg = fake_nvml()
    gpus = [device.New(m, u) for u, m, _ in cfg]
    with harness.Target() as t:
        cluster = FakeCluster([target_pod()], [u for u, _, _ in cfg])
        env = MountEnv(devices_root=t.devices_root, dev_prefix=t.dev_prefix, cgroup_driver="cgroupfs")
        worker = server.serve(server.GPUMountImpl(cluster, gpus=gpus, env=env), "127.0.0.1:0")
        mc = FakeMasterCluster([target_pod()], {"node-1": f"127.0.0.1:{worker.bound_port}"})
        m = master.Master(mc)
        try:
            status, body, _ = m.handle("GET", ADD.format(n=1, e="false") + "?report=json")
            assert status == 200 and json.loads(body) == {"result": "Success", "probe_status": None, "mounted_uuids": None, "probe": None}
        finally:
            worker.stop(0)


def test_go_parsers():
    assert [master.parse_bool(s) for s in ("1", "t", "T", "TRUE", "true", "True")] == [True] * 6
    assert [master.parse_bool(s) for s in ("0", "f", "F", "FALSE", "false", "False")] == [False] * 6
    assert [master.parse_bool(s) for s in ("", "yes", "tRuE", "2", " true")] == [None] * 5
    assert [master.parse_int32(s) for s in ("0", "8", "+8", "-1", "2147483647", "-2147483648")] == [0, 8, 8, -1, 2147483647, -2147483648]
    assert [master.parse_int32(s) for s in ("", "+", "2147483648", "1_0", "0x10", "1.0", " 1", "١")] == [None] * 8
