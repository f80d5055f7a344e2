"""The product's host logic (C++ behind the C ABI) against the golden vectors and, on random inputs, against the
oracle's C restatement of the reference (CPU only)."""

import json
import os

import pytest
from hypothesis import given, settings
from hypothesis import strategies as st

from gpumounter_b200 import ProbeError, _lib, allocator, cgroup, device, namespace, util
from gpumounter_b200._model import ContainerResources, MountEnv, Pod
from oracle import gmo_ref_ctypes as ref

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "appendix_a.json")))
U, C = GOLD["U"], GOLD["C"]


def pod_with(containers, uid=U):
    return Pod("gpu-pod", uid=uid, container_id="docker://" + C, containers=[ContainerResources(**c) for c in containers])


def make_allocator(owners):
    a = allocator.GPUAllocator()
    for i, g in enumerate(GOLD["gpus"]):
        d = device.New(g["minor"], g["uuid"])
        if str(i) in owners:
            d.PodName, d.Namespace = owners[str(i)]
            d.State = device.GPU_ALLOCATED_STATE
        a.GPUList.append(d)
    return a


@pytest.mark.parametrize("v", GOLD["to_systemd"])
def test_to_systemd(v):
    assert cgroup.CgroupName(v["parts"]).ToSystemd() == v["want"]


@pytest.mark.parametrize("v", GOLD["to_cgroupfs"])
def test_to_cgroupfs(v):
    assert cgroup.CgroupName(v["parts"]).ToCgroupfs() == v["want"]


@pytest.mark.parametrize("v", GOLD["expand_slice"])
def test_expand_slice_through_to_systemd(v):
    # ToSystemd(parts) == ExpandSlice(escape(parts) joined by "-" + ".slice"): drive ExpandSlice with one part per component
    name = v["in"]
    if not name.endswith(".slice") or "/" in name or "_" in name:
        pytest.skip("not expressible as a CgroupName")
    parts = name[:-len(".slice")].split("-")
    if parts in (["", ""], [""]):
        pytest.skip("'-.slice' / '.slice' are not reachable from a CgroupName (an empty name short-circuits to '/')")
    if v.get("error"):
        with pytest.raises(ProbeError) as e:
            cgroup.CgroupName(parts).ToSystemd()
        assert e.value.code == _lib.GMP_ERR_NAME
    else:
        assert cgroup.CgroupName(parts).ToSystemd() == v["want"]


@pytest.mark.parametrize("v", GOLD["name_check"])
def test_new_cgroup_name(v):
    if v["ok"]:
        assert cgroup.NewCgroupName(["kubepods"], v["component"]) == ["kubepods", v["component"]]
    else:
        with pytest.raises(ProbeError) as e:
            cgroup.NewCgroupName(["kubepods"], v["component"])
        assert e.value.code == _lib.GMP_ERR_NAME


@pytest.mark.parametrize("v", GOLD["cgroup_driver"])
def test_get_cgroup_driver(v, monkeypatch):
    if v["env"] is None:
        monkeypatch.delenv("CGROUP_DRIVER", raising=False)
    else:
        monkeypatch.setenv("CGROUP_DRIVER", v["env"])
    if "error" in v:
        with pytest.raises(ProbeError) as e:
            cgroup.GetCgroupDriver()
        assert e.value.code == _lib.GMP_ERR_CGROUP_DRIVER and e.value.detail == v["error"]
    else:
        assert cgroup.GetCgroupDriver() == v["want"]


@pytest.mark.parametrize("v", GOLD["qos"])
def test_pod_qos(v):
    assert cgroup.GetPodQOS(pod_with(v["containers"])) == v["want"]


@pytest.mark.parametrize("v", GOLD["cgroup_name"])
def test_get_cgroup_name(v):
    pod = pod_with(v["qos_containers"], uid=v.get("uid", U))
    if "error" in v:
        with pytest.raises(ProbeError) as e:
            cgroup.GetCgroupName(v["driver"], pod, C)
        assert e.value.code == {"driver": _lib.GMP_ERR_CGROUP_DRIVER, "name": _lib.GMP_ERR_NAME}[v["error"]]
    else:
        assert cgroup.GetCgroupName(v["driver"], pod, C) == v["want"].replace("{U}", U).replace("{C}", C)


def test_device_record_and_string():
    d = GOLD["device"]
    g = device.New(d["minor"], d["uuid"])
    assert (g.MinorNumber, g.DeviceFilePath, g.UUID, g.State, g.PodName, g.Namespace) == (
        d["minor"], d["device_file_path"], d["uuid"], d["state"], "", "")
    assert g.String() == d["string"]
    g.PodName, g.Namespace, g.State = 'we"ird<&>\\', "ns", device.GPU_ALLOCATED_STATE
    o = ref.device_new(d["minor"], d["uuid"])
    o.pod_name, o.namespace, o.state = g.PodName, "ns", "GPU_ALLOCATED_STATE"
    assert g.String() == ref.device_string(o)
    g.ResetState()
    assert g.String() == d["string"]


def test_device_group_path_and_permission_line(tmp_path):
    assert cgroup.GetDeviceGroupPath("/kubepods/podU/C") == "/sys/fs/cgroup/devices/kubepods/podU/C"
    env = MountEnv(devices_root=str(tmp_path))
    (tmp_path / "kubepods" / "podU" / "C").mkdir(parents=True)
    g = device.New(3, "GPU-x")
    cgroup.AddGPUDevicePermission("/kubepods/podU/C", g, env)
    cgroup.RemoveGPUDevicePermission("/kubepods/podU/C", g, env)
    line = GOLD["permission"]["line"] + "\n"  # `echo` appends a newline
    assert (tmp_path / "kubepods/podU/C/devices.allow").read_text() == line
    assert (tmp_path / "kubepods/podU/C/devices.deny").read_text() == line
    with pytest.raises(ProbeError) as e:
        cgroup.AddGPUDevicePermission("/missing", g, env)
    assert e.value.code == _lib.GMP_ERR_SYS


@pytest.mark.parametrize("v", GOLD["nsenter"])
def test_build_command(v):
    names = {"cgroup": "Cgroup", "cgroup_file": "CgroupFile", "follow_context": "FollowContext", "gid": "GID", "ipc": "IPC",
             "ipc_file": "IPCFile", "mount": "Mount", "mount_file": "MountFile", "net": "Net", "net_file": "NetFile",
             "no_fork": "NoFork", "pid": "PID", "pid_file": "PIDFile", "preserve_credentials": "PreserveCredentials",
             "root_directory": "RootDirectory", "target": "Target", "uid": "UID", "user": "User", "user_file": "UserFile",
             "uts": "UTS", "uts_file": "UTSFile", "working_directory": "WorkingDirectory"}
    cfg = namespace.Config(**{names[k]: (bool(x) if isinstance(getattr(namespace.Config(), names[k]), bool) else x)
                              for k, x in v["config"].items()})
    if "error" in v:
        with pytest.raises(ProbeError) as e:
            cfg.buildCommand()
        assert e.value.code == _lib.GMP_ERR_TARGET and e.value.detail == v["error"]
    else:
        n_tail = 1 + len(v["args"])  # buildCommand stops before program/args (namespace.go:162 vs :59-60)
        assert cfg.buildCommand() == v["want"][:-n_tail]


@pytest.mark.parametrize("v", GOLD["cgroup_procs"])
def test_get_cgroup_pids(v, tmp_path):
    d = tmp_path / "kubepods" / "x"
    d.mkdir(parents=True)
    (d / "cgroup.procs").write_text(v["content"])
    assert cgroup.GetCgroupPIDs("/kubepods/x", MountEnv(devices_root=str(tmp_path))) == v["want"]


@pytest.mark.parametrize("v", GOLD["can_mount"])
def test_can_mount(v):
    assert util.CanMount(v["type"], v["entire"]) == v["want"]


@pytest.mark.parametrize("v", GOLD["mount_type"], ids=lambda v: v["name"])
def test_get_mount_type(v):
    assert make_allocator(v["owners"]).GetMountType(v["pod"], v["ns"]) == v["want"]


@pytest.mark.parametrize("v", GOLD["remove_gpu"], ids=lambda v: v["name"])
def test_get_remove_gpu(v):
    a = make_allocator(v["owners"])
    got = a.GetRemoveGPU(v["pod"], v["ns"], v["uuids"])
    assert [g.UUID for g in got] == [GOLD["gpus"][i]["uuid"] for i in v["want"]]


def test_slave_pod_name():
    s = GOLD["slave_pod"]
    assert allocator.newGPUSlavePodName(s["owner"], bytes(s["rand"])) == s["name"]
    assert len(allocator.newGPUSlavePodName("x")) == len("x-slave-pod-") + 6


def test_update_gpu_status_matches_oracle():
    u = [g["uuid"] for g in GOLD["gpus"]]
    listing = [
        ("a", "default", [("main", [("nvidia.com/gpu", [u[0]]), ("example.com/fpga", ["x"])])]),
        ("a-slave-pod-aaaaaa", "gpu-pool", [("gpu-container", [("nvidia.com/gpu", [u[1], u[2]])])]),
        ("b", "default", [("main", [])]),
    ]
    a = make_allocator({"3": ["stale", "ns"]})
    a.UpdateGPUStatus(listing)
    want = ref.update_gpu_status([ref.device_new(g["minor"], g["uuid"]) for g in GOLD["gpus"]], listing)
    assert [(g.PodName, g.Namespace, g.State) for g in a.GPUList] == [(w.pod_name, w.namespace, w.state) for w in want]
    assert [g.UUID for g in a.GetPodGPUResources("a", "default")] == u[:3]
    assert a.GetMountType("a", "default") == "entire-mount"
    with pytest.raises(ProbeError) as e:
        a.UpdateGPUStatus([("z", "default", [("c", [("nvidia.com/gpu", ["GPU-unknown"])])])])
    assert e.value.code == _lib.GMP_ERR_NO_DEVICE
    assert all(g.State == "GPU_FREE_STATE" for g in a.GPUList)  # reset happened before the failure, as in the reference


# ---- randomised parity with the oracle ---------------------------------------------------------------------------
quantity = st.one_of(
    st.none(),
    st.sampled_from(["0", "1", "2", "100m", "250m", "500m", "0.5", "1.5", "1000m", "1Gi", "512Mi", "1024Mi", "1G", "1e3",
                     "1073741824", "2Gi", "128974848", "129e6", "123Mi", "1k", "0m"]))
container = st.builds(dict, cpu_request=quantity, memory_request=quantity, cpu_limit=quantity, memory_limit=quantity)
uid_text = st.text(alphabet="0123456789abcdef-", min_size=1, max_size=36)


@settings(max_examples=300, deadline=None)
@given(containers=st.lists(container, min_size=0, max_size=4), uid=uid_text, driver=st.sampled_from(["systemd", "cgroupfs"]))
def test_random_pods_qos_and_cgroup_path_match_oracle(containers, uid, driver):
    product_pod = pod_with(containers, uid=uid)
    oracle_pod = ref.Pod(uid=uid, containers=[ref.Resources(**c) for c in containers])
    assert cgroup.GetPodQOS(product_pod) == ref.pod_qos(oracle_pod)
    try:
        want = ref.get_cgroup_name(driver, oracle_pod, C)
    except ref.OracleError as e:
        with pytest.raises(ProbeError) as pe:
            cgroup.GetCgroupName(driver, product_pod, C)
        assert pe.value.code == {ref.E_PANIC_NAME: _lib.GMP_ERR_NAME, ref.E_SLICE: _lib.GMP_ERR_NAME}[e.code]
    else:
        assert cgroup.GetCgroupName(driver, product_pod, C) == want


name_part = st.text(alphabet="abcxyz-.", min_size=0, max_size=6)


@settings(max_examples=300, deadline=None)
@given(parts=st.lists(name_part, min_size=0, max_size=5))
def test_random_cgroup_names_match_oracle(parts):
    assert cgroup.CgroupName(parts).ToCgroupfs() == ref.to_cgroupfs(parts)
    try:
        want = ref.to_systemd(parts)
    except ref.OracleError:
        with pytest.raises(ProbeError):
            cgroup.CgroupName(parts).ToSystemd()
    else:
        assert cgroup.CgroupName(parts).ToSystemd() == want


owner = st.sampled_from([("", ""), ("a", "default"), ("a", "other"), ("a-slave-pod-111111", "gpu-pool"),
                         ("a-slave-pod-222222", "gpu-pool"), ("xa-slave-pod-333333", "gpu-pool"),
                         ("a-slave-pod-444444", "default"), ("b", "default"), ("b-slave-pod-555555", "gpu-pool")])


@settings(max_examples=300, deadline=None)
@given(owners=st.lists(owner, min_size=4, max_size=4),
       uuids=st.lists(st.sampled_from([g["uuid"] for g in GOLD["gpus"]] + ["x", "y"]), max_size=4),
       pod=st.sampled_from(["a", "b", "xa"]))
def test_random_ownership_selection_matches_oracle(owners, uuids, pod):
    a = make_allocator({str(i): list(o) for i, o in enumerate(owners) if o[0]})
    o_gpus = []
    for g, (name, ns) in zip(GOLD["gpus"], owners):
        d = ref.device_new(g["minor"], g["uuid"])
        if name:
            d.pod_name, d.namespace, d.state = name, ns, "GPU_ALLOCATED_STATE"
        o_gpus.append(d)
    assert a.GetMountType(pod, "default") == ref.mount_type_string(ref.get_mount_type(o_gpus, pod, "default"))
    assert [g.MinorNumber for g in a.GetPodGPUResources(pod, "default")] == ref.pod_gpu_resources(o_gpus, pod, "default")
    assert [g.MinorNumber for g in a.GetRemoveGPU(pod, "default", uuids)] == ref.get_remove_gpu(o_gpus, pod, "default", uuids)


def test_device_string_follows_go_on_non_ascii_bytes():
    """json.Marshal walks a string rune by rune (encoding/json encode.go): well-formed UTF-8 is copied, every byte of an
    ill-formed sequence becomes \\ufffd, U+2028 / U+2029 are escaped, DEL is left alone.  Expected strings written out by
    hand from those rules; oracle and library must both produce them."""
    import ctypes

    from gpumounter_b200 import lib
    from gpumounter_b200._model import gmp_gpu
    from oracle import gmo_ref_ctypes as ref

    cases = [
        (b"caf\xc3\xa9 \xe2\x82\xac \xf0\x9f\x98\x80", "café € 😀"),
        (b"a\xe2\x80\xa8b\xe2\x80\xa9c", "a\\u2028b\\u2029c"),
        (b"\xff", "\\ufffd"),
        (b"\xed\xa0\x80", "\\ufffd\\ufffd\\ufffd"),            # a surrogate is not UTF-8
        (b"\xc0\xaf", "\\ufffd\\ufffd"),                        # overlong '/'
        (b"\xf4\x90\x80\x80", "\\ufffd\\ufffd\\ufffd\\ufffd"),  # above U+10FFFF
        (b"x\xe2\x82", "x\\ufffd\\ufffd"),                      # cut off mid-rune
        (b"\x7f\x1f", "\x7f\\u001f"),
    ]
    for raw, want in cases:
        g = gmp_gpu()
        assert lib().gmp_device_new(3, b"GPU-x", ctypes.byref(g)) == 0
        g.pod_name = raw
        buf = ctypes.create_string_buffer(4096)
        assert lib().gmp_device_string(ctypes.byref(g), buf, 4096) == 0
        expected = ('{"MinorNumber":3,"DeviceFilePath":"/dev/nvidia3","UUID":"GPU-x","State":"GPU_FREE_STATE","PodName":"'
                    + want + '","Namespace":""}').encode()
        assert buf.value == expected, (raw, buf.value)
        og = ref.gmo_gpu()
        ref._lib().gmo_device_new(3, b"GPU-x", ctypes.byref(og))
        og.pod_name = raw
        obuf = ctypes.create_string_buffer(4096)
        assert ref._lib().gmo_device_string(ctypes.byref(og), obuf, 4096) >= 0
        assert obuf.value == expected, (raw, obuf.value)
