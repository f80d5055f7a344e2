"""The effectful mount path: the product's native MountGPU / UnmountGPU against the oracle's fork/exec restatement of
the reference, on a cluster-less target (own mount namespace + private /dev tmpfs + device-cgroup directory).
CPU only; needs root, unshare and nsenter (skipped otherwise)."""

import os
import subprocess
import time

import pytest

from gpumounter_b200 import ProbeError, _lib, cgroup, device, lib, namespace, util
from gpumounter_b200._model import MountEnv, Pod
from oracle import gmo_ref_ctypes as ref
from tests import harness
from tests.conftest import FAKE_UUIDS

reason = harness.can_run()
pytestmark = pytest.mark.skipif(reason is not None, reason=reason or "")


def product_pod():
    return Pod("gpu-pod", uid=harness.UID, container_id="docker://" + harness.CID)


def oracle_pod():
    return ref.Pod(uid=harness.UID, container_id="docker://" + harness.CID)


def env_of(t: harness.Target) -> MountEnv:
    return MountEnv(devices_root=t.devices_root, dev_prefix=t.dev_prefix, cgroup_driver="cgroupfs")


def oracle_kw(t: harness.Target):
    return dict(devices_root=t.devices_root, dev_prefix=t.dev_prefix, cgroup_driver="cgroupfs")


def test_harness_paths_agree_with_both_implementations():
    with harness.Target() as t:
        assert cgroup.GetCgroupName("cgroupfs", product_pod(), harness.CID) == t.cgroup_path
        assert ref.get_cgroup_name("cgroupfs", oracle_pod(), harness.CID) == t.cgroup_path
        assert cgroup.GetCgroupPIDs(t.cgroup_path, env_of(t)) == [str(t.pid)]


def test_mount_creates_the_same_node_and_permission_line_as_the_reference_path(fake_nvml):
    fake_nvml()
    with harness.Target() as a, harness.Target() as b:
        gpu = device.New(3, FAKE_UUIDS[0])
        util.MountGPU(product_pod(), gpu, env_of(a))                       # product: write(2) + setns + mknod(2)
        ref.mount_gpu(oracle_pod(), ref.device_new(3, FAKE_UUIDS[0]), **oracle_kw(b))  # oracle: sh/echo + nsenter/sh/mknod
        assert a.node(3) == b.node(3) == (195, 3, 0o666)
        assert a.nodes() == b.nodes() == ["nvidia3"]
        assert a.host_sees() == b.host_sees() == []                         # only inside the target's mount namespace
        assert a.read("devices.allow") == b.read("devices.allow") == "c 195:3 rw\n"
        # unmount, not busy
        util.UnmountGPU(product_pod(), gpu, False, env_of(a))
        ref.unmount_gpu(oracle_pod(), ref.device_new(3, FAKE_UUIDS[0]), False, [], **oracle_kw(b))
        assert a.node(3) is None and b.node(3) is None
        assert a.read("devices.deny") == b.read("devices.deny") == "c 195:3 rw\n"


def test_mount_twice_fails_like_mknod_does(fake_nvml):
    fake_nvml()
    with harness.Target() as a, harness.Target() as b:
        gpu = device.New(5, FAKE_UUIDS[1])
        util.MountGPU(product_pod(), gpu, env_of(a))
        with pytest.raises(ProbeError) as e:
            util.MountGPU(product_pod(), gpu, env_of(a))
        assert e.value.code == _lib.GMP_ERR_SYS and "File exists" in e.value.detail
        ref.mount_gpu(oracle_pod(), ref.device_new(5, FAKE_UUIDS[1]), **oracle_kw(b))
        with pytest.raises(ref.OracleError) as oe:
            ref.mount_gpu(oracle_pod(), ref.device_new(5, FAKE_UUIDS[1]), **oracle_kw(b))
        assert oe.value.code == ref.E_EXEC and "File exists" in str(oe.value)
        # removing a node that is not there fails in both, too (rm without -f)
        with pytest.raises(ProbeError):
            util.UnmountGPU(product_pod(), device.New(6, FAKE_UUIDS[2]), False, env_of(a))
        with pytest.raises(ref.OracleError):
            ref.unmount_gpu(oracle_pod(), ref.device_new(6, FAKE_UUIDS[2]), False, [], **oracle_kw(b))


def test_empty_cgroup_and_missing_cgroup(fake_nvml):
    fake_nvml()
    with harness.Target() as t:
        gpu = device.New(1, FAKE_UUIDS[0])
        open(os.path.join(t.cgroup_dir, "cgroup.procs"), "w").write(str(t.pid))  # no trailing newline -> dropped
        with pytest.raises(ProbeError) as e:
            util.MountGPU(product_pod(), gpu, env_of(t))
        assert e.value.code == _lib.GMP_ERR_EMPTY_CGROUP       # Go: index out of range panic at util.go:50
        with pytest.raises(ref.OracleError) as oe:
            ref.mount_gpu(oracle_pod(), ref.device_new(1, FAKE_UUIDS[0]), **oracle_kw(t))
        assert oe.value.code == ref.E_PANIC_INDEX
        assert t.read("devices.allow") == "c 195:1 rw\n"        # both wrote the permission before failing, as the reference does
        env = env_of(t)
        env.devices_root = t.base + "/nowhere"
        with pytest.raises(ProbeError) as e:
            util.MountGPU(product_pod(), gpu, env)
        assert e.value.code == _lib.GMP_ERR_SYS


def test_unsupported_cgroup_driver_is_reported_first(fake_nvml, monkeypatch):
    fake_nvml()
    monkeypatch.delenv("CGROUP_DRIVER", raising=False)
    with harness.Target() as t:
        env = MountEnv(devices_root=t.devices_root, dev_prefix=t.dev_prefix)  # driver from $CGROUP_DRIVER: unset
        with pytest.raises(ProbeError) as e:
            util.MountGPU(product_pod(), device.New(0, FAKE_UUIDS[0]), env)
        assert e.value.code == _lib.GMP_ERR_CGROUP_DRIVER and e.value.detail == "unsupported cgroup driver: "
        assert t.read("devices.allow") is None
        monkeypatch.setenv("CGROUP_DRIVER", "cgroupfs")
        util.MountGPU(product_pod(), device.New(0, FAKE_UUIDS[0]), env)
        assert t.node(0) == (195, 0, 0o666)


def test_busy_gpu_is_refused_without_force_and_killed_with_force(fake_nvml):
    with harness.Target() as a, harness.Target() as b:
        pa, pb = a.spawn_member(), b.spawn_member()
        fake_nvml(procs=[(FAKE_UUIDS[0], pa.pid, 1 << 20, "compute"), (FAKE_UUIDS[0], 999999, 5, "graphics")])
        gpu = device.New(2, FAKE_UUIDS[0])
        ogpu = ref.device_new(2, FAKE_UUIDS[0])
        util.MountGPU(product_pod(), gpu, env_of(a))
        ref.mount_gpu(oracle_pod(), ogpu, **oracle_kw(b))
        assert util.GetPodGPUProcesses(product_pod(), gpu, env_of(a)) == [str(pa.pid)]
        assert util.GetPodGPUProcesses(product_pod(), device.New(4, FAKE_UUIDS[3]), env_of(a)) is None
        with pytest.raises(ProbeError) as e:
            util.UnmountGPU(product_pod(), gpu, False, env_of(a))
        assert e.value.code == _lib.GMP_ERR_BUSY
        with pytest.raises(ref.OracleError) as oe:
            ref.unmount_gpu(oracle_pod(), ogpu, False, [999999, pb.pid], **oracle_kw(b))
        assert oe.value.code == ref.E_BUSY
        assert a.node(2) == b.node(2) == (195, 2, 0o666) and a.read("devices.deny") is None and b.read("devices.deny") is None
        # force: permission removed, node removed, the pod's GPU process gets SIGTERM — in both
        util.UnmountGPU(product_pod(), gpu, True, env_of(a))
        ref.unmount_gpu(oracle_pod(), ogpu, True, [999999, pb.pid], **oracle_kw(b))
        assert a.node(2) is None and b.node(2) is None
        for p in (pa, pb):
            assert p.wait(timeout=5) == -15
        assert a.proc.poll() is None and b.proc.poll() is None  # the container's main process is untouched


def test_batched_mount_and_unmount_of_eight_gpus(fake_nvml):
    gpus_cfg = fake_nvml()
    with harness.Target() as a, harness.Target() as b:
        gpus = [device.New(m, u) for u, m, _ in gpus_cfg]
        assert util.MountGPUs(product_pod(), gpus, env_of(a)) == 8
        for u, m, _ in gpus_cfg:  # the reference's serial loop, server.go:80-95
            ref.mount_gpu(oracle_pod(), ref.device_new(m, u), **oracle_kw(b))
        assert a.nodes() == b.nodes() == [f"nvidia{i}" for i in range(8)]
        assert [a.node(i) for i in range(8)] == [b.node(i) for i in range(8)] == [(195, i, 0o666) for i in range(8)]
        assert a.read("devices.allow") == b.read("devices.allow")  # the last line written wins in a plain file
        # the busy check asks NVML once per GPU (ref: util.go:100-109, nvidia.go:58-63: init .. shutdown each time);
        # batched, the eight queries share one driver attach
        import ctypes
        from tests.conftest import FAKE_NVML
        cold = ctypes.CDLL(FAKE_NVML).fake_nvml_cold_inits
        cold.restype = ctypes.c_uint
        before = cold()
        assert util.UnmountGPUs(product_pod(), gpus, False, env_of(a)) == 8
        assert cold() - before == 1
        assert a.nodes() == []
        # a failure part-way reports how far it got and leaves earlier GPUs mounted (no rollback, as in the reference)
        util.MountGPU(product_pod(), gpus[2], env_of(a))
        done = None
        try:
            util.MountGPUs(product_pod(), gpus, env_of(a))
        except ProbeError as e:
            done = e
        assert done is not None and done.code == _lib.GMP_ERR_SYS
        assert a.nodes() == sorted(f"nvidia{g.MinorNumber}" for g in gpus[:3])


def test_namespace_calls_directly(fake_nvml):
    fake_nvml()
    with harness.Target() as t:
        cfg = namespace.Config(Mount=True, Target=t.pid)
        gpu = device.New(7, FAKE_UUIDS[7])
        env = env_of(t)
        namespace.AddGPUDeviceFile(cfg, gpu, env)
        assert t.node(7) == (195, 7, 0o666)
        out, err = cfg.Execute("sh", "-c", f"ls {t.devdir}")   # the reference's generic nsenter path still works
        assert out.split() == ["nvidia7"] and err == ""
        namespace.RemoveGPUDeviceFile(cfg, gpu, env)
        assert t.node(7) is None
        with pytest.raises(ProbeError) as e:
            namespace.AddGPUDeviceFile(namespace.Config(Mount=True, Target=0), gpu, env)
        assert e.value.code == _lib.GMP_ERR_TARGET
        with pytest.raises(ProbeError) as e:
            namespace.AddGPUDeviceFile(namespace.Config(Mount=True, Target=2 ** 22 + 12345), gpu, env)
        assert e.value.code == _lib.GMP_ERR_SYS
        p = t.spawn_member()
        namespace.KillRunningGPUProcesses(cfg, [str(p.pid)])
        assert p.wait(timeout=5) == -15
        with pytest.raises(ProbeError):
            namespace.KillRunningGPUProcesses(cfg, [str(p.pid)])  # already gone: kill(1) would fail as well


REAL = "/sys/fs/cgroup/devices"


@pytest.mark.skipif(not (os.path.isdir(REAL) and os.access(REAL, os.W_OK)), reason="no writable cgroup-v1 devices hierarchy")
def test_real_devices_cgroup_allow_and_deny(fake_nvml):
    fake_nvml()
    try:
        t = harness.Target(real_cgroup=True).start()
    except OSError as e:
        pytest.skip(f"cannot create a devices cgroup: {e}")
    try:
        with open(os.path.join(t.cgroup_dir, "devices.deny"), "w") as f:
            f.write("a")  # start from an empty allow-list so that the effect is visible in devices.list
        assert t.read("devices.list") == ""
        gpu = device.New(3, FAKE_UUIDS[0])
        util.MountGPU(product_pod(), gpu, env_of(t))
        assert t.read("devices.list") == "c 195:3 rw\n"
        assert t.node(3) == (195, 3, 0o666)
        util.UnmountGPU(product_pod(), gpu, False, env_of(t))
        assert t.read("devices.list") == ""
        ref.mount_gpu(oracle_pod(), ref.device_new(3, FAKE_UUIDS[0]), **oracle_kw(t))   # the reference path, same cgroup
        assert t.read("devices.list") == "c 195:3 rw\n"
        ref.unmount_gpu(oracle_pod(), ref.device_new(3, FAKE_UUIDS[0]), False, [], **oracle_kw(t))
        assert t.read("devices.list") == "" and t.node(3) is None
    finally:
        with open("/sys/fs/cgroup/devices/cgroup.procs", "w") as f:
            f.write(f"{t.pid}\n")
        t.stop()


def test_probe_in_pod_runs_inside_the_targets_cgroup_and_mount_namespace(fake_nvml, tmp_path):
    """A stand-in 'probe' (a static shell-free C-less script is not possible with fexecve of a script whose
    interpreter must exist in the pod; the harness target shares the host root, so /bin/sh works): it reports
    where it runs, which is what this test is about."""
    fake_nvml()
    fake_probe = tmp_path / "fake-probe"
    fake_probe.write_text(
        "#!/bin/sh\n"
        "printf '{\"status\":0,\"args\":\"%s\",\"mnt\":\"%s\",\"nodes\":\"%s\",\"self\":%s}\\n' "
        "\"$*\" \"$(readlink /proc/self/ns/mnt)\" \"$(ls \"$GMP_TEST_DEVDIR\" | tr '\\n' ' ')\" \"$$\"\n")
    fake_probe.chmod(0o755)
    with harness.Target() as t:
        gpus = [device.New(3, FAKE_UUIDS[0]), device.New(2, FAKE_UUIDS[1])]
        util.MountGPUs(product_pod(), gpus, env_of(t))
        os.environ["GMP_TEST_DEVDIR"] = t.devdir
        try:
            rc, rep = util.ProbeInPod(product_pod(), gpus, ["--bytes", "1M"], env_of(t), probe_exe=str(fake_probe))
        finally:
            del os.environ["GMP_TEST_DEVDIR"]
        assert rc == 0 and rep["status"] == 0
        assert rep["args"] == f"--uuid {FAKE_UUIDS[0]}:3 --uuid {FAKE_UUIDS[1]}:2 --bytes 1M"
        assert rep["mnt"] == os.readlink(f"/proc/{t.pid}/ns/mnt") != os.readlink("/proc/self/ns/mnt")
        assert rep["nodes"].split() == ["nvidia2", "nvidia3"]          # sees the pod's /dev, not the host's
        assert str(rep["self"]) in t.read("cgroup.procs").split()       # joined the pod's device cgroup first
        # failures of the set-up are named, the probe's own verdict is passed through
        with pytest.raises(ProbeError) as e:
            util.ProbeInPod(product_pod(), gpus, [], env_of(t), probe_exe=str(tmp_path / "missing"))
        assert e.value.code == _lib.GMP_ERR_SYS and "open the probe binary" in e.value.detail
        bad = tmp_path / "bad-probe"
        bad.write_text("#!/bin/sh\necho '{\"status\":-7}'\nexit 7\n")
        bad.chmod(0o755)
        rc, rep = util.ProbeInPod(product_pod(), gpus, [], env_of(t), probe_exe=str(bad), raise_on_error=False)
        assert rc == _lib.GMP_ERR_VERIFY and rep["status"] == -7


def test_unmount_is_verified_from_inside_the_pod(fake_nvml):
    """RemoveGPU-side check with the real gpumounter-probe binary in --gone mode: runs in the target's device cgroup and
    mount namespace and must see the nodes gone; a node that survived is GMP_ERR_STILL_MOUNTED."""
    fake_nvml()
    with harness.Target() as t:
        gpus = [device.New(3, FAKE_UUIDS[0]), device.New(2, FAKE_UUIDS[1])]
        env = env_of(t)
        util.MountGPUs(product_pod(), gpus, env)
        rc, rep = util.VerifyUnmountedInPod(product_pod(), gpus, env, dev_dir=t.devdir, raise_on_error=False)
        assert rc == _lib.GMP_ERR_STILL_MOUNTED and rep["status"] == rc and "nvidia3" in rep["detail"]
        util.UnmountGPUs(product_pod(), gpus[:1], False, env)
        rc, rep = util.VerifyUnmountedInPod(product_pod(), gpus, env, dev_dir=t.devdir, raise_on_error=False)
        assert rc == _lib.GMP_ERR_STILL_MOUNTED and "nvidia2" in rep["detail"]      # the second one is still there
        assert util.VerifyUnmountedInPod(product_pod(), gpus[:1], env, dev_dir=t.devdir)[0] == 0
        util.UnmountGPUs(product_pod(), gpus[1:], False, env)
        rc, rep = util.VerifyUnmountedInPod(product_pod(), gpus, env, dev_dir=t.devdir)
        assert rc == 0 and rep == {"status": 0, "verdict": "ok", "detail": "", "checked": 2}
        # the host's view of that directory was empty all along: the check really ran inside the target
        assert t.host_sees() == []


@pytest.mark.skipif(not (os.path.isdir(REAL) and os.access(REAL, os.W_OK)), reason="no writable cgroup-v1 devices hierarchy")
def test_real_v1_cgroup_gates_open_until_the_gpu_is_mounted(fake_nvml):
    """A real cgroup-v1 devices group set up like a container without GPUs (default-deny + the standard devices + the
    NVIDIA control node): the kernel refuses open() of a GPU node until MountGPU has written devices.allow, lets it through
    afterwards (ENXIO: no driver on this box), and refuses again after UnmountGPU."""
    import errno

    fake_nvml()
    try:
        t = harness.Target(real_cgroup=True, container_whitelist=True).start()
    except OSError as e:
        pytest.skip(f"cannot create a devices cgroup: {e}")
    try:
        assert "c 195:255 rw" in t.read("devices.list") and "a *:* rwm" not in t.read("devices.list")
        gpu = device.New(3, FAKE_UUIDS[0])
        os.mknod(t.node_path_in_target(3), 0o666 | 0o020000, os.makedev(195, 3))
        assert t.try_open(3) == errno.EPERM
        os.unlink(t.node_path_in_target(3))
        util.MountGPU(product_pod(), gpu, env_of(t))
        assert t.try_open(3) in (errno.ENXIO, errno.ENODEV)
        assert "c 195:3 rw\n" in t.read("devices.list")
        util.UnmountGPU(product_pod(), gpu, False, env_of(t))
        os.mknod(t.node_path_in_target(3), 0o666 | 0o020000, os.makedev(195, 3))
        assert t.try_open(3) == errno.EPERM
        os.unlink(t.node_path_in_target(3))
    finally:
        with open("/sys/fs/cgroup/devices/cgroup.procs", "w") as f:
            f.write(f"{t.pid}\n")
        t.stop()


def test_check_nodes_in_pod_without_executing_anything(fake_nvml):
    """Stage 1 of the resident hook for images where the probe binary cannot start: the helper itself opens the nodes
    from inside the pod's cgroup and mount namespace."""
    gpus_cfg = fake_nvml()
    with harness.Target() as t:
        env, pod = env_of(t), product_pod()
        gpus = [device.New(m, u) for u, m, _ in gpus_cfg[:2]]
        rc, idx = util.CheckNodesInPod(pod, gpus, env, raise_on_error=False)
        assert rc == _lib.GMP_ERR_NODE_CTL and "nvidiactl" in lib().gmp_last_error().decode()
        open(f"/proc/{t.pid}/root{t.devdir}/nvidiactl", "w").close()
        assert util.CheckNodesInPod(pod, gpus, env, raise_on_error=False)[0] == _lib.GMP_ERR_NODE_UVM
        open(f"/proc/{t.pid}/root{t.devdir}/nvidia-uvm", "w").close()
        assert util.CheckNodesInPod(pod, gpus, env, raise_on_error=False) == (_lib.GMP_ERR_NODE_GPU, 0)     # not mounted yet
        util.MountGPU(pod, gpus[0], env)
        rc, idx = util.CheckNodesInPod(pod, gpus, env, raise_on_error=False)
        # gpus[0] now has a node; with no driver on this box open() says ENXIO, which is still "GPU node unusable" for
        # index 0 - on a GPU box it opens and the verdict moves on to the unmounted gpus[1]
        assert rc == _lib.GMP_ERR_NODE_GPU and idx in (0, 1)
        assert util.CheckNodesInPod(pod, [], env) == (0, -1)                                                # control nodes only
        # every check joined the pod's cgroup first (the scratch cgroup.procs is a plain file: each join is a line)
        assert len(t.read("cgroup.procs").split()) == 1 + 5
    with pytest.raises(ProbeError) as e:
        util.CheckNodesInPod(product_pod(), [], MountEnv(devices_root="/nonexistent", cgroup_driver="cgroupfs"))
    assert e.value.code == _lib.GMP_ERR_SYS


@pytest.mark.skipif(not (os.path.isdir(REAL) and os.access(REAL, os.W_OK)), reason="no writable cgroup-v1 devices hierarchy")
def test_check_nodes_in_pod_sees_the_kernels_verdict(fake_nvml):
    fake_nvml()
    try:
        t = harness.Target(real_cgroup=True, container_whitelist=True).start()
    except OSError as e:
        pytest.skip(f"cannot create a devices cgroup: {e}")
    try:
        env, pod, gpu = env_of(t), product_pod(), device.New(3, FAKE_UUIDS[0])
        for name in ("nvidiactl", "nvidia-uvm"):
            open(f"/proc/{t.pid}/root{t.devdir}/{name}", "w").close()
        os.mknod(t.node_path_in_target(3), 0o666 | 0o020000, os.makedev(195, 3))      # a node, but no permission
        assert util.CheckNodesInPod(pod, [gpu], env, raise_on_error=False) == (_lib.GMP_ERR_NODE_DENIED, 0)
        os.unlink(t.node_path_in_target(3))
        util.MountGPU(pod, gpu, env)
        rc, idx = util.CheckNodesInPod(pod, [gpu], env, raise_on_error=False)
        assert rc in (0, _lib.GMP_ERR_NODE_GPU) and "Operation not permitted" not in lib().gmp_last_error().decode()
    finally:
        with open("/sys/fs/cgroup/devices/cgroup.procs", "w") as f:
            f.write(f"{t.pid}\n")
        t.stop()


def test_verify_gone_in_pod_without_executing_anything(fake_nvml):
    gpus_cfg = fake_nvml()
    with harness.Target() as t:
        env, pod = env_of(t), product_pod()
        gpus = [device.New(m, u) for u, m, _ in gpus_cfg[:3]]
        util.MountGPUs(pod, gpus, env)
        assert util.VerifyGoneInPod(pod, gpus, env, raise_on_error=False) == (_lib.GMP_ERR_STILL_MOUNTED, 0)
        util.UnmountGPUs(pod, gpus[:2], False, env)
        rc, idx = util.VerifyGoneInPod(pod, gpus, env, raise_on_error=False)
        assert (rc, idx) == (_lib.GMP_ERR_STILL_MOUNTED, 2) and f"nvidia{gpus[2].MinorNumber}" in lib().gmp_last_error().decode()
        assert util.VerifyGoneInPod(pod, gpus[:2], env) == (0, -1)
        util.UnmountGPUs(pod, gpus[2:], False, env)
        assert util.VerifyGoneInPod(pod, gpus, env) == (0, -1)
        # the same verdict as the probe binary's --gone mode, run inside
        rc, rep = util.VerifyUnmountedInPod(pod, gpus, env, dev_dir=t.devdir, raise_on_error=False)
        assert rc == 0 and rep["status"] == 0


def test_in_pod_probe_is_killed_when_it_overruns_its_deadline(fake_nvml, tmp_path):
    """A GPU that hangs must not hang the worker with it: gmp_mount_env.probe_timeout_ms bounds the in-pod probe."""
    import time

    fake_nvml()
    slow = tmp_path / "slow-probe"
    slow.write_text("#!/bin/sh\necho $$ > %s\nexec sleep 30\n" % (tmp_path / "pid"))
    slow.chmod(0o755)
    quick = tmp_path / "quick-probe"
    quick.write_text("#!/bin/sh\nsleep 0.2\necho '{\"status\":0}'\n")
    quick.chmod(0o755)
    with harness.Target() as t:
        env = env_of(t)
        env.probe_timeout_ms = 400
        gpu = device.New(3, FAKE_UUIDS[0])
        t0 = time.monotonic()
        with pytest.raises(ProbeError) as e:
            util.ProbeInPod(product_pod(), [gpu], [], env, probe_exe=str(slow))
        took = time.monotonic() - t0
        assert e.value.code == _lib.GMP_ERR_TIMEOUT and "400 ms" in e.value.detail and 0.35 < took < 5
        pid = int((tmp_path / "pid").read_text())
        time.sleep(0.05)
        assert not os.path.exists(f"/proc/{pid}") or open(f"/proc/{pid}/stat").read().split()[2] == "Z"   # killed, not left running
        rc, rep = util.ProbeInPod(product_pod(), [gpu], [], env, probe_exe=str(quick))     # a probe within the deadline is untouched
        assert rc == 0 and rep == {"status": 0}


def test_pod_like_target_has_its_own_dev_and_only_mounted_gpus_open(fake_nvml):
    """The stand-in the GPU runs use (tools/standin_pod.py, pod_like): tmpfs ON /dev of the target's mount namespace,
    populated like a container started with NVIDIA_VISIBLE_DEVICES=none, in a real devices cgroup with a container's
    whitelist.  With the reference's own names (dev_prefix unset: /dev/nvidia<N>) the mount lands in THAT /dev and
    nowhere else, and the kernel lets exactly the mounted minors through."""
    import errno

    why = harness.can_run_pod_like()
    if why:
        pytest.skip(why)
    gpus_cfg = fake_nvml()
    t = harness.Target(pod_like=True).start()
    try:
        env = MountEnv(devices_root=t.devices_root, cgroup_driver="cgroupfs")
        assert {"null", "zero", "urandom", "tty", "nvidiactl"} <= set(t.nodes()) and t.gpu_nodes() == []
        recs = [device.New(m, u) for u, m, _ in gpus_cfg[:3]]
        minors = [m for _, m, _ in gpus_cfg[:3]]
        other = next(m for m in range(8) if m not in minors)
        before_host = sorted(x for x in os.listdir("/dev") if x.startswith("nvidia"))
        assert util.MountGPUs(product_pod(), recs, env) == 3
        assert t.gpu_nodes() == sorted(f"nvidia{m}" for m in minors)
        assert sorted(x for x in os.listdir("/dev") if x.startswith("nvidia")) == before_host  # the host's /dev is untouched
        for m in minors:
            assert t.node(m) == (195, m, 0o666)
            assert t.try_open(m) in (errno.ENXIO, errno.ENODEV)  # through the filter; no driver behind it on this box
        os.mknod(t.node_path_in_target(other), 0o666 | 0o020000, os.makedev(195, other))
        assert t.try_open(other) == errno.EPERM                   # a node alone opens nothing
        os.unlink(t.node_path_in_target(other))
        # a process started the way the pod's own processes are sees the pod's /dev and sits in the pod's cgroup
        p = t.spawn_in_pod(["sh", "-c", "ls /dev; grep devices /proc/self/cgroup"], stdout=subprocess.PIPE, text=True)
        out = p.communicate(timeout=30)[0]
        assert all(f"nvidia{m}" in out.split() for m in minors) and t.cgroup_path in out
        assert util.UnmountGPUs(product_pod(), recs, False, env) == 3
        assert t.gpu_nodes() == [] and "195:%d" % minors[0] not in t.read("devices.list")
        rc, _ = util.VerifyGoneInPod(product_pod(), recs, env, raise_on_error=False)
        assert rc == 0
    finally:
        with open("/sys/fs/cgroup/devices/cgroup.procs", "w") as f:
            f.write(f"{t.pid}\n")
        t.stop()


def test_in_pod_probe_runs_with_the_pod_users_identity_and_no_capabilities(fake_nvml, tmp_path):
    """What gmp_probe_in_pod execs resolves its loader and libraries from the POD's root, so it must not hold more than the
    tenant does: identity of the pod's first process, empty capability sets (bounding included), no_new_privs - and the
    worker that spawned it keeps all of its own.  GMP_EXT_PROBE_KEEP_PRIVILEGES is the explicit opt-out."""
    from gpumounter_b200._model import EXT_PROBE_KEEP_PRIVILEGES, MountEnv as Env

    fake_nvml()
    fake_probe = tmp_path / "who-am-i"
    fake_probe.write_text("#!/bin/sh\n"
                          "s() { grep \"^$1:\" /proc/self/status | tr -s '\\t ' ' ' | cut -d' ' -f2; }\n"
                          "printf '{\"status\":0,\"uid\":%s,\"gid\":%s,\"eff\":\"%s\",\"prm\":\"%s\",\"bnd\":\"%s\",\"nnp\":%s,\"groups\":\"%s\"}\\n' "
                          "\"$(s Uid)\" \"$(s Gid)\" \"$(s CapEff)\" \"$(s CapPrm)\" \"$(s CapBnd)\" \"$(s NoNewPrivs)\" "
                          "\"$(grep '^Groups:' /proc/self/status | cut -f2- | tr -d '\\n')\"\n")
    fake_probe.chmod(0o755)

    def caps_of_self():
        return {k: v for k, v in (line.split(":\t") for line in open("/proc/self/status").read().splitlines()
                                  if line.startswith(("CapEff", "CapPrm", "CapBnd", "Uid", "NoNewPrivs")))}

    mine = caps_of_self()
    with harness.Target() as t:
        gpus = [device.New(3, FAKE_UUIDS[0])]
        util.MountGPUs(product_pod(), gpus, env_of(t))
        # a pod whose first process runs as an ordinary user: the probe runs as that user
        member = subprocess.Popen(["setpriv", "--reuid", "12345", "--regid", "23456", "--clear-groups", "sleep", "600"])
        try:
            with open(os.path.join(t.cgroup_dir, "cgroup.procs"), "w") as f:  # scratch cgroup file: make it the FIRST pid
                f.write(f"{member.pid}\n{t.pid}\n")
            for _ in range(100):  # setpriv has exec'ed sleep once the uid shows
                if f"Uid:\t12345" in open(f"/proc/{member.pid}/status").read():
                    break
                time.sleep(0.01)
            rc, rep = util.ProbeInPod(product_pod(), gpus, [], env_of(t), probe_exe=str(fake_probe))
            assert rc == 0
            assert (rep["uid"], rep["gid"], rep["groups"].strip()) == (12345, 23456, "")
            assert int(rep["eff"], 16) == int(rep["prm"], 16) == int(rep["bnd"], 16) == 0 and rep["nnp"] == 1
        finally:
            member.kill()
            member.wait()
        # first process is root (the usual container): still root by name, but without a single capability
        with open(os.path.join(t.cgroup_dir, "cgroup.procs"), "w") as f:
            f.write(f"{t.pid}\n")
        rc, rep = util.ProbeInPod(product_pod(), gpus, [], env_of(t), probe_exe=str(fake_probe))
        assert rc == 0 and rep["uid"] == 0
        assert int(rep["eff"], 16) == int(rep["prm"], 16) == int(rep["bnd"], 16) == 0 and rep["nnp"] == 1
        # the worker itself is untouched by what its helper did to itself (raw system calls, not glibc's setxid broadcast)
        assert caps_of_self() == mine
        # opt-out: round 1's behaviour
        e = env_of(t)
        keep = Env(devices_root=e.devices_root, dev_prefix=e.dev_prefix, cgroup_driver=e.cgroup_driver,
                   extensions=EXT_PROBE_KEEP_PRIVILEGES)
        rc, rep = util.ProbeInPod(product_pod(), gpus, [], keep, probe_exe=str(fake_probe))
        assert rc == 0 and int(rep["eff"], 16) == int(mine["CapEff"], 16) and rep["nnp"] == 0


def test_fanout_probe_starts_one_process_per_gpu_each_in_a_narrowed_child_group(fake_nvml, tmp_path):
    """gmp_probe_in_pod_fanout on a real cgroup-v1 devices group: every per-GPU process sits in a child group that holds
    the NVIDIA control node and ITS GPU only, all of them run inside the pod's mount namespace, the whole-pod process
    runs in the pod's own group afterwards, the merged report carries all of it, and the child groups are gone again."""
    why = harness.can_run_pod_like()
    if why:
        pytest.skip(why)
    gpus_cfg = fake_nvml()
    fake_probe = tmp_path / "fake-probe"
    fake_probe.write_text(
        "#!/bin/sh\n"
        "now=$(date +%s%N)\n"
        "printf '{\"status\":0,\"args\":\"%s\",\"cgroup\":\"%s\",\"list\":\"%s\",\"mnt\":\"%s\",\"hbm\":[{\"uuid\":\"x\"}],"
        "\"devices\":[{\"uuid\":\"x\"}],\"ms_cuda_init\":1.5,\"mono_start_ns\":%s,\"mono_usable_ns\":%s}\\n' "
        "\"$*\" \"$(grep devices /proc/self/cgroup | cut -d: -f3)\" "
        "\"$(cat /sys/fs/cgroup/devices$(grep devices /proc/self/cgroup | cut -d: -f3)/devices.list | grep '195:' | tr '\\n' ';')\" "
        "\"$(readlink /proc/self/ns/mnt)\" \"$now\" \"$(( now + 1000 ))\"\n")
    fake_probe.chmod(0o755)
    t = harness.Target(pod_like=True).start()
    try:
        env = MountEnv(devices_root=t.devices_root, cgroup_driver="cgroupfs")
        recs = [device.New(m, u) for u, m, _ in gpus_cfg[:3]]
        util.MountGPUs(product_pod(), recs, env)
        rc, rep = util.ProbeInPodFanout(product_pod(), recs, ["--bytes", "1M"], env, probe_exe=str(fake_probe))
        assert rc == 0 and rep["status"] == 0 and rep["fanout"] is True and rep["n_gpus"] == 3
        assert len(rep["reports"]) == 3 and len(rep["hbm"]) == 3
        pod_group = t.cgroup_dir[len("/sys/fs/cgroup/devices"):]
        for i, (r, g) in enumerate(zip(rep["reports"], recs)):
            assert r["args"] == f"--uuid {g.UUID}:{g.MinorNumber} --bytes 1M --skip-p2p"
            assert r["cgroup"].startswith(pod_group + "/gmp-probe-") and r["cgroup"].endswith(f"-{i}")
            assert sorted(x for x in r["list"].split(";") if x) == sorted([f"c 195:{g.MinorNumber} rw", "c 195:255 rw"])
            assert r["mnt"] == os.readlink(f"/proc/{t.pid}/ns/mnt")
        w = rep["whole_pod"]
        assert w["cgroup"] == pod_group and w["args"].endswith("--bytes 1M --skip-hbm") and w["args"].count("--uuid") == 3
        assert rep["mono_usable_ns"] == max(r["mono_usable_ns"] for r in rep["reports"])
        assert rep["mono_start_ns"] == min(r["mono_start_ns"] for r in rep["reports"])
        assert [d for d in os.listdir(t.cgroup_dir) if d.startswith("gmp-probe-")] == []
        # a probe that fails on one GPU fails the call, names the GPU, and the whole-pod step is not run
        bad = tmp_path / "bad-probe"
        bad.write_text("#!/bin/sh\ncase \"$*\" in *:%d*) echo '{\"status\":-7}'; exit 7;; esac\necho '{\"status\":0}'\n" % recs[1].MinorNumber)
        bad.chmod(0o755)
        rc, rep = util.ProbeInPodFanout(product_pod(), recs, [], env, probe_exe=str(bad), raise_on_error=False)
        assert rc == _lib.GMP_ERR_VERIFY and rep["status"] == -7 and recs[1].UUID in rep["detail"] and rep["whole_pod"] is None
        assert rep["mono_usable_ns"] == 0
        assert [d for d in os.listdir(t.cgroup_dir) if d.startswith("gmp-probe-")] == []
        util.UnmountGPUs(product_pod(), recs, False, env)
    finally:
        with open("/sys/fs/cgroup/devices/cgroup.procs", "w") as f:
            f.write(f"{t.pid}\n")
        t.stop()


def test_fanout_probe_deadline_kills_every_child_and_removes_the_child_groups(fake_nvml, tmp_path):
    """A GPU that hangs must not hang AddGPU: probe_timeout_ms bounds the per-GPU phase, every probe process is killed,
    reaped, and the child groups are removed."""
    why = harness.can_run_pod_like()
    if why:
        pytest.skip(why)
    gpus_cfg = fake_nvml()
    slow = tmp_path / "slow-probe"
    slow.write_text("#!/bin/sh\nexec sleep 30\n")
    slow.chmod(0o755)
    t = harness.Target(pod_like=True).start()
    try:
        env = MountEnv(devices_root=t.devices_root, cgroup_driver="cgroupfs", probe_timeout_ms=300)
        recs = [device.New(m, u) for u, m, _ in gpus_cfg[:2]]
        util.MountGPUs(product_pod(), recs, env)
        t0 = time.monotonic()
        rc, rep = util.ProbeInPodFanout(product_pod(), recs, [], env, probe_exe=str(slow), raise_on_error=False)
        assert rc == _lib.GMP_ERR_TIMEOUT and time.monotonic() - t0 < 5
        assert [d for d in os.listdir(t.cgroup_dir) if d.startswith("gmp-probe-")] == []
        assert t.read("cgroup.procs").split() == [str(t.pid)]          # nobody of ours is left in the pod's group
        util.UnmountGPUs(product_pod(), recs, False, env)
    finally:
        with open("/sys/fs/cgroup/devices/cgroup.procs", "w") as f:
            f.write(f"{t.pid}\n")
        t.stop()


def test_fanout_in_a_pod_that_holds_a_wildcard_cannot_narrow_and_says_so_to_its_probes(fake_nvml, tmp_path):
    """A privileged pod ("a *:* rwm" in its devices group) cannot be narrowed per GPU: every per-GPU probe will see all of
    the pod's GPUs, so it is started with --host-view (more devices visible than named is expected there)."""
    why = harness.can_run_pod_like()
    if why:
        pytest.skip(why)
    gpus_cfg = fake_nvml()
    echo = tmp_path / "echo-probe"
    echo.write_text("#!/bin/sh\nprintf '{\"status\":0,\"args\":\"%s\",\"mono_start_ns\":1,\"mono_usable_ns\":2}\\n' \"$*\"\n")
    echo.chmod(0o755)
    t = harness.Target(real_cgroup=True).start()   # a real devices group left at the parent's allow-everything
    try:
        assert "a *:* rwm" in t.read("devices.list")
        env = MountEnv(devices_root=t.devices_root, dev_prefix=t.dev_prefix, cgroup_driver="cgroupfs")
        recs = [device.New(m, u) for u, m, _ in gpus_cfg[:2]]
        util.MountGPUs(product_pod(), recs, env)
        rc, rep = util.ProbeInPodFanout(product_pod(), recs, [], env, probe_exe=str(echo))
        assert rc == 0 and all(r["args"].endswith("--skip-p2p --host-view") for r in rep["reports"])
        assert "--host-view" not in rep["whole_pod"]["args"]
        util.UnmountGPUs(product_pod(), recs, False, env)
    finally:
        t.stop()
