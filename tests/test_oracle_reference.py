"""oracle/gmo_reference.c against the golden vectors (CPU only).

One vector is the reference's own (comment at pkg/util/cgroup/cgroup.go:47-49); the rest are hand-derived
from the cited Go code — see tests/golden/appendix_a.json.
"""

import json
import os

import pytest

from oracle import gmo_ref_ctypes as ref

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "appendix_a.json")))
U, C = GOLD["U"], GOLD["C"]
MT = {"entire-mount": 0, "single-mount": 1, "no-mount": 2, "unknown-mount": 3}


def make_gpus(owners):
    gpus = []
    for i, g in enumerate(GOLD["gpus"]):
        d = ref.device_new(g["minor"], g["uuid"])
        if str(i) in owners:
            d.pod_name, d.namespace = owners[str(i)]
            d.state = "GPU_ALLOCATED_STATE"
        gpus.append(d)
    return gpus


def pod_with(containers, uid=U):
    return ref.Pod(uid=uid, containers=[ref.Resources(**c) for c in containers])


def test_the_reference_comment_vector():
    v = GOLD["to_systemd"][0]
    assert v["source"].startswith("reference")
    assert ref.to_systemd(v["parts"]) == v["want"]


@pytest.mark.parametrize("v", GOLD["to_systemd"])
def test_to_systemd(v):
    assert ref.to_systemd(v["parts"]) == v["want"]


@pytest.mark.parametrize("v", GOLD["expand_slice"])
def test_expand_slice(v):
    if v.get("error"):
        with pytest.raises(ref.OracleError) as e:
            ref.expand_slice(v["in"])
        assert e.value.code == ref.E_SLICE
    else:
        assert ref.expand_slice(v["in"]) == v["want"]


@pytest.mark.parametrize("v", GOLD["to_cgroupfs"])
def test_to_cgroupfs(v):
    assert ref.to_cgroupfs(v["parts"]) == v["want"]


@pytest.mark.parametrize("v", GOLD["name_check"])
def test_new_cgroup_name_panics(v):
    assert (ref.new_cgroup_name_check(v["component"]) == 0) == v["ok"]


@pytest.mark.parametrize("v", GOLD["cgroup_driver"])
def test_get_cgroup_driver(v):
    if "error" in v:
        with pytest.raises(ref.OracleError) as e:
            ref.get_cgroup_driver(v["env"])
        assert e.value.code == ref.E_DRIVER and str(e.value).endswith(v["error"])
    else:
        assert ref.get_cgroup_driver(v["env"]) == v["want"]


@pytest.mark.parametrize("v", GOLD["qos"])
def test_pod_qos(v):
    assert ref.pod_qos(pod_with(v["containers"])) == v["want"]


@pytest.mark.parametrize("a,b,want", GOLD["quantity_cmp"])
def test_quantity_cmp(a, b, want):
    assert ref.quantity_cmp(a, b) == want
    assert ref.quantity_cmp(b, a) == -want


def test_quantity_rejects_garbage():
    for bad in ("", "abc", "1Qi", "1..2", "--1", "1 Gi"):
        with pytest.raises(ref.OracleError):
            ref.quantity_cmp(bad, "1")


@pytest.mark.parametrize("v", GOLD["cgroup_name"])
def test_get_cgroup_name(v):
    pod = pod_with(v["qos_containers"], uid=v.get("uid", U))
    if "error" in v:
        with pytest.raises(ref.OracleError) as e:
            ref.get_cgroup_name(v["driver"], pod, C)
        assert e.value.code == {"driver": ref.E_DRIVER, "name": ref.E_PANIC_NAME}[v["error"]]
    else:
        assert ref.get_cgroup_name(v["driver"], pod, C) == v["want"].replace("{U}", U).replace("{C}", C)


def test_device_record():
    d = GOLD["device"]
    g = ref.device_new(d["minor"], d["uuid"])
    assert (g.device_file_path, g.state, g.pod_name, g.namespace) == (d["device_file_path"], d["state"], "", "")
    assert ref.device_string(g) == d["string"]
    g.pod_name = 'we"ird<&>\\'
    assert json.loads(ref.device_string(g))["PodName"] == g.pod_name
    assert "\\u003c\\u0026\\u003e" in ref.device_string(g)  # Go's json.Marshal HTML-escapes


def test_permission_command_and_device_group_path():
    p = GOLD["permission"]
    assert ref.permission_command(p["cgroup_path"], p["minor"], True) == p["allow"]
    assert ref.permission_command(p["cgroup_path"], p["minor"], False) == p["deny"]
    assert ref.device_group_path("/x") == "/sys/fs/cgroup/devices/x"
    assert ref.device_group_path("/x", "/tmp/root") == "/tmp/root/x"


@pytest.mark.parametrize("v", GOLD["nsenter"])
def test_nsenter_argv(v):
    if "error" in v:
        with pytest.raises(ref.OracleError) as e:
            ref.nsenter_argv(v["program"], v["args"], **v["config"])
        assert e.value.code == ref.E_TARGET and v["error"] in str(e.value)
    else:
        assert ref.nsenter_argv(v["program"], v["args"], **v["config"]) == v["want"]


def test_shell_command_strings():
    c = GOLD["commands"]
    assert ref.mknod_command(3) == c["mknod"]
    assert ref.rm_command(3) == c["rm"]
    assert ref.kill_command(c["kill_pids"]) == c["kill"]


@pytest.mark.parametrize("v", GOLD["cgroup_procs"])
def test_get_cgroup_pids(v, tmp_path):
    d = tmp_path / "kubepods" / "x"
    d.mkdir(parents=True)
    (d / "cgroup.procs").write_text(v["content"])
    assert ref.get_cgroup_pids(str(tmp_path), "/kubepods/x") == v["want"]


def test_get_cgroup_pids_missing_file(tmp_path):
    with pytest.raises(ref.OracleError) as e:
        ref.get_cgroup_pids(str(tmp_path), "/nope")
    assert e.value.code == ref.E_IO


@pytest.mark.parametrize("v", GOLD["can_mount"])
def test_can_mount(v):
    assert ref.can_mount(MT[v["type"]], v["entire"]) == v["want"]


@pytest.mark.parametrize("v", GOLD["mount_type"], ids=lambda v: v["name"])
def test_get_mount_type(v):
    t = ref.get_mount_type(make_gpus(v["owners"]), v["pod"], v["ns"])
    assert ref.mount_type_string(t) == v["want"]


@pytest.mark.parametrize("v", GOLD["remove_gpu"], ids=lambda v: v["name"])
def test_get_remove_gpu(v):
    assert ref.get_remove_gpu(make_gpus(v["owners"]), v["pod"], v["ns"], v["uuids"]) == v["want"]


def test_slave_pod_naming_and_count():
    s = GOLD["slave_pod"]
    assert ref.slave_pod_name(s["owner"], bytes(s["rand"])) == s["name"]
    for gpu_num, entire, count, per in s["counts"]:
        assert ref.slave_pod_count(gpu_num, entire) == (count, per)


@pytest.mark.parametrize("v", GOLD["pod_gpu_processes"])
def test_pod_gpu_processes(v):
    assert ref.pod_gpu_processes(v["cgroup"], v["gpu"]) == v["want"]


def test_update_gpu_status_and_pod_resources():
    gpus = make_gpus({"3": ["stale", "ns"]})
    u = [g["uuid"] for g in GOLD["gpus"]]
    listing = [
        ("a", "default", [("main", [("nvidia.com/gpu", [u[0]]), ("example.com/fpga", ["x"])])]),
        ("a-slave-pod-aaaaaa", "gpu-pool", [("gpu-container", [("nvidia.com/gpu", [u[1], u[2]])])]),
        ("b", "default", [("main", [])]),
    ]
    out = ref.update_gpu_status(gpus, listing)
    assert [(g.pod_name, g.namespace, g.state) for g in out] == [
        ("a", "default", "GPU_ALLOCATED_STATE"), ("a-slave-pod-aaaaaa", "gpu-pool", "GPU_ALLOCATED_STATE"),
        ("a-slave-pod-aaaaaa", "gpu-pool", "GPU_ALLOCATED_STATE"), ("", "", "GPU_FREE_STATE")]
    assert ref.pod_gpu_resources(out, "a", "default") == [0, 1, 2]
    assert ref.pod_gpu_resources(out, "a-slave-pod-aaaaaa", "gpu-pool") == [1, 2]
    assert ref.pod_gpu_resources(out, "b", "default") == []
    with pytest.raises(ref.OracleError):
        ref.update_gpu_status(gpus, [("z", "default", [("c", [("nvidia.com/gpu", ["GPU-unknown"])])])])


def test_http_mapping():
    for result, code, body in GOLD["http_add"]:
        assert ref.add_result_http(result, "p", "n1") == (code, body)
    for result, code, body in GOLD["http_remove"]:
        assert ref.remove_result_http(result, "p", "n1", ["u1", "u2"]) == (code, body)
