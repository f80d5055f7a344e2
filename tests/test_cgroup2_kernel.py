"""cgroup-v2 extension against the real kernel (header section 7; no reference counterpart): device programs are loaded,
attached, read back, edited and ENFORCED on a scratch cgroup of this machine's unified hierarchy.  Skipped where there
is no writable cgroup2 mount or bpf(2) is refused.  The "runtime's" program is always attached by the independent
Python side (oracle/gmo_devfilter.py), the library under test only ever edits it."""

import errno
import os

import pytest

from gpumounter_b200 import ProbeError, _lib, cgroup2 as c, device, util
from gpumounter_b200._model import EXT_CGROUP_V2, EXT_CRI_RUNTIMES, MountEnv, Pod
from oracle import gmo_devfilter as o
from tests import harness

why = harness.can_run_cgroup2()
pytestmark = pytest.mark.skipif(why is not None, reason=f"cgroup2 kernel path not testable here: {why}")

NO_DRIVER = (errno.ENXIO, errno.ENODEV)  # open() got past the device filter and found no driver behind the node


def env_of(t, ext=EXT_CGROUP_V2):
    return MountEnv(cgroup2_root=t.devices_root, dev_prefix=t.dev_prefix, cgroup_driver="cgroupfs", extensions=ext)


def product_pod(cid="docker://" + harness.CID):
    return Pod("gpu-pod", uid=harness.UID, container_id=cid)


def attached(t):
    ids, flags = c.device_programs(t.cgroup_dir)
    return [c.read_program(i) for i in ids], flags


def test_runtime_program_reads_back_exactly_and_decodes(fake_nvml):
    with harness.Target(cgroup2=True) as t:
        progs, flags = attached(t)
        assert flags == 2 and len(progs) == 1                      # BPF_F_ALLOW_MULTI, as runc attaches
        assert progs[0] == o.runc_program(o.RUNC_DEFAULT_RULES)    # translated form == loaded form for this program type
        rules, default = c.decode(progs[0])
        assert [(r.type, r.major, r.minor, r.access, r.allow) for r in rules] == o.RUNC_DEFAULT_RULES and default is False


def test_mount_and_unmount_on_the_unified_hierarchy_are_enforced_by_the_kernel(fake_nvml):
    gpus_cfg = fake_nvml()
    with harness.Target(cgroup2=True) as t:
        env, pod = env_of(t), product_pod()
        baseline, _ = attached(t)
        gpus = [device.New(m, u) for u, m, _ in gpus_cfg]
        # a node alone is not access: made by hand (the runtime's wildcard 'm' rule allows mknod), open is refused
        os.mknod(t.node_path_in_target(3), 0o666 | 0o020000, os.makedev(195, 3))
        assert t.try_open(3) == errno.EPERM
        os.unlink(t.node_path_in_target(3))
        assert util.MountGPUs(pod, gpus[:2], env) == 2
        m0, m1 = gpus[0].MinorNumber, gpus[1].MinorNumber
        assert t.node(m0) == (195, m0, 0o666) and t.node(m1) == (195, m1, 0o666)
        assert t.try_open(m0) in NO_DRIVER and t.try_open(m1) in NO_DRIVER       # the filter lets them through now
        other = gpus[2].MinorNumber
        os.mknod(t.node_path_in_target(other), 0o666 | 0o020000, os.makedev(195, other))
        assert t.try_open(other) == errno.EPERM                                  # ... and only them
        os.unlink(t.node_path_in_target(other))
        assert t.try_open(3 if 3 not in (m0, m1) else 4) == errno.ENOENT
        progs, flags = attached(t)
        assert flags == 2 and len(progs) == 1
        rules, _ = c.decode(progs[0])
        assert rules[-2:] == [c.gpu_rule(m0), c.gpu_rule(m1)] and len(rules) == len(o.RUNC_DEFAULT_RULES) + 2
        # unmount one: its node goes, its rule goes, the other GPU keeps working
        assert util.UnmountGPUs(pod, gpus[:1], False, env) == 1
        assert t.node(m0) is None and t.try_open(m1) in NO_DRIVER
        os.mknod(t.node_path_in_target(m0), 0o666 | 0o020000, os.makedev(195, m0))
        assert t.try_open(m0) == errno.EPERM
        os.unlink(t.node_path_in_target(m0))
        # unmount the other: the program is byte for byte what the runtime attached ("devices.list back to baseline")
        assert util.UnmountGPUs(pod, gpus[1:2], False, env) == 1
        assert attached(t)[0] == baseline and t.nodes() == []


def test_churn_does_not_grow_the_program(fake_nvml):
    gpus_cfg = fake_nvml()
    with harness.Target(cgroup2=True) as t:
        env, pod = env_of(t), product_pod()
        gpus = [device.New(m, u) for u, m, _ in gpus_cfg]
        baseline, _ = attached(t)
        fds0 = len(os.listdir("/proc/self/fd"))
        for k in list(range(1, 9)) + list(range(7, 0, -1)):
            util.MountGPUs(pod, gpus[:k], env)
            assert len(c.decode(attached(t)[0][0])[0]) == len(o.RUNC_DEFAULT_RULES) + k
            util.UnmountGPUs(pod, gpus[:k], False, env)
            assert attached(t)[0] == baseline
        assert len(os.listdir("/proc/self/fd")) == fds0


def test_grant_twice_is_idempotent_and_revoke_of_a_runtime_granted_device_denies(fake_nvml):
    fake_nvml()
    runtime = o.RUNC_DEFAULT_RULES + [(o.DEV_CHAR, 195, 5, o.ACC_ALL, True)]   # the pod was STARTED with GPU 5
    with harness.Target(cgroup2=True, runtime_rules=runtime) as t:
        os.mknod(t.node_path_in_target(5), 0o666 | 0o020000, os.makedev(195, 5))
        assert t.try_open(5) in NO_DRIVER
        before = attached(t)[0]
        assert c.update_devices(t.cgroup_dir, [c.gpu_rule(7)], True) == 1
        once = attached(t)[0]
        assert c.update_devices(t.cgroup_dir, [c.gpu_rule(7)], True) == 1
        assert attached(t)[0] == once != before
        assert c.update_devices(t.cgroup_dir, [c.gpu_rule(7)], False) == 1
        assert attached(t)[0] == before
        # revoking what the runtime itself granted cannot be done by removing a rule of ours: a deny rule goes on top
        assert c.update_devices(t.cgroup_dir, [c.gpu_rule(5)], False) == 1
        rules, _ = c.decode(attached(t)[0][0])
        assert rules[-1] == c.Rule(c.DEV_CHAR, 195, 5, c.ACC_READ | c.ACC_WRITE, False)
        assert t.try_open(5) == errno.EPERM and t.try_open(5, os.O_RDONLY) == errno.EPERM
        # and granting it again takes the deny rule away instead of stacking another allow on it
        assert c.update_devices(t.cgroup_dir, [c.gpu_rule(5)], True) == 1
        rules, _ = c.decode(attached(t)[0][0])
        assert rules[-1] == c.gpu_rule(5) and len(rules) == len(runtime) + 1
        assert t.try_open(5) in NO_DRIVER


def test_program_in_another_layout_is_wrapped_not_rewritten(fake_nvml):
    """A runtime that is not runc (crun, systemd: first match wins, 64-bit moves, JSET access test).  The library cannot
    turn it into a rule list; it keeps it verbatim and puts its own rules in front.  Same kernel-enforced behaviour."""
    gpus_cfg = fake_nvml()
    original = o.first_match_program(o.RUNC_DEFAULT_RULES)
    with harness.Target(cgroup2=True, runtime_program=original) as t:
        env, pod = env_of(t), product_pod()
        gpus = [device.New(m, u) for u, m, _ in gpus_cfg]
        assert attached(t)[0] == [original]
        with pytest.raises(ProbeError):
            c.decode(original)
        assert util.MountGPUs(pod, gpus[:3], env) == 3
        now = attached(t)[0][0]
        assert now[0] == (0xB4, 0, 0, 0, 0x474D5046) and now[1] == (0xB4, 0, 0, 0, 3) and now[-len(original):] == original
        for g in gpus[:3]:
            assert t.try_open(g.MinorNumber) in NO_DRIVER
        other = gpus[3].MinorNumber
        os.mknod(t.node_path_in_target(other), 0o666 | 0o020000, os.makedev(195, other))
        assert t.try_open(other) == errno.EPERM
        os.unlink(t.node_path_in_target(other))
        assert util.UnmountGPUs(pod, gpus[:2], False, env) == 2
        assert t.try_open(gpus[2].MinorNumber) in NO_DRIVER and attached(t)[0][0][1] == (0xB4, 0, 0, 0, 1)
        assert util.UnmountGPUs(pod, gpus[2:3], False, env) == 1
        assert attached(t)[0] == [original] and t.nodes() == []


def test_two_programs_on_one_cgroup_are_both_edited(fake_nvml):
    """BPF_F_ALLOW_MULTI: every attached program must say yes.  runc's program plus a second one in another layout."""
    gpus_cfg = fake_nvml()
    with harness.Target(cgroup2=True) as t:
        second = o.first_match_program(o.RUNC_DEFAULT_RULES)
        fd = o.kernel_load(second)
        o.kernel_attach(t.cgroup_dir, fd)
        os.close(fd)
        before = attached(t)
        assert len(before[0]) == 2
        u, m, _ = gpus_cfg[0]
        gpu = device.New(m, u)
        util.MountGPU(product_pod(), gpu, env_of(t))
        assert t.try_open(m) in NO_DRIVER and len(attached(t)[0]) == 2
        util.UnmountGPU(product_pod(), gpu, False, env_of(t))
        assert sorted(attached(t)[0]) == sorted(before[0]) and t.nodes() == []


def test_program_that_cannot_be_reloaded_leaves_everything_untouched(fake_nvml):
    gpus_cfg = fake_nvml()
    with harness.Target(cgroup2=True) as t:
        # a second program that calls a helper (bpf_get_current_uid_gid = 15): its translated form is not loadable
        fd = o.kernel_load([(0x85, 0, 0, 0, 15), (0xB7, 0, 0, 0, 1), (0x95, 0, 0, 0, 0)])
        o.kernel_attach(t.cgroup_dir, fd)
        os.close(fd)
        before = attached(t)
        assert len(before[0]) == 2
        u, m, _ = gpus_cfg[0]
        with pytest.raises(ProbeError) as e:
            util.MountGPU(product_pod(), device.New(m, u), env_of(t))
        assert e.value.code == _lib.GMP_ERR_UNSUPPORTED and "re-loaded" in e.value.detail
        assert attached(t) == before and t.nodes() == []          # nothing half done: no rule edited, no node made


def test_cgroup_without_a_device_program_is_unrestricted(fake_nvml):
    gpus_cfg = fake_nvml()
    with harness.Target(cgroup2=True) as t:
        bare = os.path.join(t.devices_root, "bare")
        os.mkdir(bare)
        try:
            assert c.device_programs(bare) == ([], 0)
            assert c.update_devices(bare, [c.gpu_rule(0)], True) == 0
            assert c.device_programs(bare) == ([], 0)
        finally:
            os.rmdir(bare)
        with pytest.raises(ProbeError) as e:
            c.device_programs(os.path.join(t.devices_root, "missing"))
        assert e.value.code == _lib.GMP_ERR_SYS


def test_in_pod_probe_joins_the_v2_cgroup(fake_nvml, tmp_path):
    """gmp_probe_in_pod writes the child into <cgroup2_root><path>/cgroup.procs before it execs the probe: the script
    below stands in for gpumounter-probe and reports which v2 cgroup it runs in and what open() says."""
    gpus_cfg = fake_nvml()
    with harness.Target(cgroup2=True) as t:
        env, pod = env_of(t), product_pod()
        u, m, _ = gpus_cfg[0]
        gpu = device.New(m, u)
        util.MountGPU(pod, gpu, env)
        script = tmp_path / "fake-probe"
        script.write_text("#!/bin/sh\n"
                          "cg=$(grep '^0::' /proc/self/cgroup | cut -d: -f3)\n"
                          f"if sh -c ': <> {t.devdir}/nvidia{m}' 2>/dev/null; then r=opened; else r=refused; fi\n"
                          "printf '{\"status\":0,\"cgroup\":\"%s\",\"open\":\"%s\"}\\n' \"$cg\" \"$r\"\n")
        script.chmod(0o755)
        rc, rep = util.ProbeInPod(pod, [gpu], [], env, probe_exe=str(script), raise_on_error=False)
        assert rc == 0 and rep["cgroup"].endswith(t.cgroup_path), rep


@pytest.mark.parametrize("cid, driver, leaf", [
    ("containerd://" + harness.CID, "cgroupfs", harness.CID),
    ("containerd://" + harness.CID, "systemd", f"cri-containerd-{harness.CID}.scope"),
    ("cri-o://" + harness.CID, "cgroupfs", f"crio-{harness.CID}"),
    ("cri-o://" + harness.CID, "systemd", f"crio-{harness.CID}.scope"),
    ("docker://" + harness.CID, "systemd", f"docker-{harness.CID}.scope"),
])
def test_cri_runtime_container_ids(cid, driver, leaf):
    from gpumounter_b200 import cgroup

    pod = product_pod(cid)
    on = cgroup.GetCgroupNameOfPod(pod, MountEnv(cgroup_driver=driver, extensions=EXT_CRI_RUNTIMES))
    assert on.rsplit("/", 1)[1] == leaf
    off = cgroup.GetCgroupNameOfPod(pod, MountEnv(cgroup_driver=driver))
    # without the extension: the reference's behaviour, which only knows how to strip "docker://" (ref: util.go:23)
    ref_id = cid.replace("docker://", "", 1)
    assert off.endswith("/" + ref_id if driver == "cgroupfs" else f"/docker-{ref_id}.scope")
    assert on.rsplit("/", 1)[0] == off[:len(on.rsplit("/", 1)[0])]   # same pod-level directory either way


def test_concurrent_editors_of_one_cgroup_do_not_lose_updates(fake_nvml):
    """Two workers (threads here) granting and revoking different GPUs on the same cgroup at once.  The replacement names
    the program it read (BPF_F_REPLACE): whoever loses the race gets ENOENT from the kernel, re-reads and tries again,
    so no grant is overwritten by a stale read-modify-write."""
    import threading

    fake_nvml()
    with harness.Target(cgroup2=True) as t:
        baseline = attached(t)[0]
        errors = []

        def editor(minors):
            try:
                for _ in range(40):
                    for m in minors:
                        c.update_devices(t.cgroup_dir, [c.gpu_rule(m)], True)
                    for m in minors:
                        c.update_devices(t.cgroup_dir, [c.gpu_rule(m)], False)
                for m in minors:
                    c.update_devices(t.cgroup_dir, [c.gpu_rule(m)], True)
            except Exception as e:  # noqa: BLE001
                errors.append(e)

        threads = [threading.Thread(target=editor, args=(ms,)) for ms in ([0, 1], [2, 3], [4, 5], [6, 7])]
        for th in threads:
            th.start()
        for th in threads:
            th.join()
        assert errors == []
        rules, _ = c.decode(attached(t)[0][0])
        assert sorted(r.minor for r in rules if r.major == 195) == list(range(8))      # every editor's last grant is there
        assert c.update_devices(t.cgroup_dir, [c.gpu_rule(m) for m in range(8)], False) == 1
        assert attached(t)[0] == baseline


def test_real_probe_binary_sees_the_v2_verdict_from_inside_the_pod(fake_nvml):
    """gpumounter-probe --nodes-only, fexecve'd inside the target (v2 cgroup joined, mount namespace entered): a node
    whose device the filter refuses is GMP_ERR_NODE_DENIED; once mounted, the open gets past the filter (and finds no
    driver on this box: GMP_ERR_NODE_GPU with ENXIO's text, where a GPU box would go on to CUDA)."""
    gpus_cfg = fake_nvml()
    with harness.Target(cgroup2=True) as t:
        env, pod = env_of(t), product_pod()
        u, m, _ = gpus_cfg[0]
        gpu = device.New(m, u)
        for name in ("nvidiactl", "nvidia-uvm"):
            open(f"/proc/{t.pid}/root{t.devdir}/{name}", "w").close()
        os.mknod(t.node_path_in_target(m), 0o666 | 0o020000, os.makedev(195, m))   # a node, but no grant
        rc, rep = util.ProbeInPod(pod, [gpu], ["--dev-dir", t.devdir, "--nodes-only"], env, raise_on_error=False)
        assert rc == _lib.GMP_ERR_NODE_DENIED and rep["nodes_status"] == _lib.GMP_ERR_NODE_DENIED, rep
        assert "Operation not permitted" in rep["detail"]
        os.unlink(t.node_path_in_target(m))
        util.MountGPU(pod, gpu, env)
        rc, rep = util.ProbeInPod(pod, [gpu], ["--dev-dir", t.devdir, "--nodes-only"], env, raise_on_error=False)
        assert rc in (_lib.GMP_ERR_NODE_GPU, _lib.GMP_ERR_DRIVER_LIB, 0), rep
        assert "Operation not permitted" not in rep["detail"]
        util.UnmountGPU(pod, gpu, False, env)
        rc, gone = util.VerifyUnmountedInPod(pod, [gpu], env, dev_dir=t.devdir, raise_on_error=False)
        assert rc == 0 and gone["status"] == 0


def test_detect_cgroup_version(tmp_path):
    from gpumounter_b200 import cgroup

    assert cgroup.DetectCgroupVersion() in (1, 2)                       # this machine, whatever it is
    (tmp_path / "devices").mkdir()
    assert cgroup.DetectCgroupVersion(MountEnv(devices_root=str(tmp_path / "devices"))) == 1
    v2 = o.cgroup2_mount()
    assert cgroup.DetectCgroupVersion(MountEnv(devices_root=str(tmp_path / "absent"), cgroup2_root=v2)) == 2
    with pytest.raises(ProbeError) as e:
        cgroup.DetectCgroupVersion(MountEnv(devices_root=str(tmp_path / "absent"), cgroup2_root=str(tmp_path)))
    assert e.value.code == _lib.GMP_ERR_UNAVAILABLE


def test_interpreters_compute_what_the_kernel_computes(tmp_path):
    """Random ALU / JMP / LDX device programs are attached to a scratch cgroup; a child inside it calls mknod(2) for a
    handful of (type, major, minor) and the kernel's verdict (success / EPERM) is compared with gmp_devfilter_eval and with
    the Python interpreter.  Programs the verifier refuses are skipped.  This is what pins both interpreters."""
    import json
    import random
    import subprocess
    import sys
    import uuid

    rng = random.Random(20240611)
    alu = [0x00, 0x10, 0x20, 0x40, 0x50, 0x60, 0x70, 0xA0, 0xB0, 0xC0]
    jmp = [0x10, 0x20, 0x30, 0x40, 0x50, 0x60, 0x70, 0xA0, 0xB0, 0xC0, 0xD0]
    imms = [0, 1, 2, 7, 16, 31, 195, 0xFFFF, 0x7FFFFFFF, -1, -2]

    def program():
        n = rng.randint(1, 16)
        prog = [(0x61, 2, 1, 0, 0), (0x61, 3, 1, 4, 0), (0x61, 4, 1, 8, 0), (0xB7, 0, 0, 0, rng.choice(imms)), (0xB7, 5, 0, 0, rng.choice(imms))]
        for i in range(n):
            kind = rng.randint(0, 9)
            dst, src = rng.choice([0, 2, 3, 4, 5]), rng.choice([0, 2, 3, 4, 5])
            if kind <= 5:
                cls, op, use_reg = rng.choice([0x04, 0x07]), rng.choice(alu), rng.random() < 0.5
                imm = rng.choice(imms)
                if op in (0x60, 0x70, 0xC0) and not use_reg:
                    imm = rng.choice([0, 1, 7, 16, 31])                      # the verifier refuses shifts >= the width
                prog.append((cls | op | (0x08 if use_reg else 0), dst, src if use_reg else 0, 0, 0 if use_reg else imm))
            elif kind == 6:
                prog.append((rng.choice([0x84, 0x87]), dst, 0, 0, 0))
            elif kind == 7:
                prog.append(rng.choice([(0x61, dst, 1, 0, 0), (0x61, dst, 1, 4, 0), (0x61, dst, 1, 8, 0), (0x69, dst, 1, 0, 0),
                                        (0x69, dst, 1, 2, 0), (0x71, dst, 1, 1, 0)]))
            else:
                cls, op, use_reg = rng.choice([0x05, 0x06]), rng.choice(jmp), rng.random() < 0.5
                prog.append((cls | op | (0x08 if use_reg else 0), dst, src if use_reg else 0, rng.randint(0, n - 1 - i),
                             0 if use_reg else rng.choice(imms)))
        prog += [(0x57, 0, 0, 0, 1), (0x95, 0, 0, 0, 0)]   # r0 &= 1: the verifier wants a return value it can prove is 0 or 1
        return prog

    requests = [(t, ma, mi) for t in (c.DEV_CHAR, c.DEV_BLOCK) for ma, mi in [(1, 3), (195, 0), (195, 255), (0, 1), (4095, 1048575), (7, 16)]]
    # (not 0:0: a character device 0:0 is overlayfs's whiteout, which vfs_mknod lets through without asking the filter)
    child = (
        "import json, os, stat, sys\n"
        "cg, base, reqs = sys.argv[1], sys.argv[2], json.loads(sys.argv[3])\n"
        "open(cg + '/cgroup.procs', 'w').write(str(os.getpid()))\n"
        "out = []\n"
        "for k, (t, ma, mi) in enumerate(reqs):\n"
        "    path = f'{base}/n{k}'\n"
        "    try:\n"
        "        os.mknod(path, (stat.S_IFCHR if t == 2 else stat.S_IFBLK) | 0o600, os.makedev(ma, mi)); os.unlink(path); out.append(0)\n"
        "    except OSError as e:\n"
        "        out.append(e.errno)\n"
        "print(json.dumps(out))\n")
    root = o.cgroup2_mount()
    checked = skipped = 0
    while checked < 120 and checked + skipped < 600:
        prog = program()
        try:
            fd = o.kernel_load(prog)
        except OSError:
            skipped += 1
            continue
        cg = f"{root}/gmp-fuzz-{uuid.uuid4().hex[:8]}"
        os.mkdir(cg)
        try:
            o.kernel_attach(cg, fd)
            os.close(fd)
            out = subprocess.run([sys.executable, "-c", child, cg, str(tmp_path), json.dumps(requests)], capture_output=True, text=True,
                                 timeout=60)
            assert out.returncode == 0, out.stderr
            verdicts = json.loads(out.stdout)
        finally:
            os.rmdir(cg)
        for (t, ma, mi), err in zip(requests, verdicts):
            assert err in (0, errno.EPERM), err
            kernel_allows = err == 0
            assert c.evaluate(prog, t, c.ACC_MKNOD, ma, mi) == kernel_allows, (prog, t, ma, mi)
            assert o.interpret(prog, t, c.ACC_MKNOD, ma, mi) == kernel_allows, (prog, t, ma, mi)
        checked += 1
    assert checked == 120, (checked, skipped)


def test_check_nodes_in_pod_on_the_unified_hierarchy(fake_nvml):
    gpus_cfg = fake_nvml()
    with harness.Target(cgroup2=True) as t:
        env, pod = env_of(t), product_pod()
        u, m, _ = gpus_cfg[0]
        gpu = device.New(m, u)
        for name in ("nvidiactl", "nvidia-uvm"):
            open(f"/proc/{t.pid}/root{t.devdir}/{name}", "w").close()
        os.mknod(t.node_path_in_target(m), 0o666 | 0o020000, os.makedev(195, m))
        assert util.CheckNodesInPod(pod, [gpu], env, raise_on_error=False) == (_lib.GMP_ERR_NODE_DENIED, 0)   # the v2 program says no
        os.unlink(t.node_path_in_target(m))
        util.MountGPU(pod, gpu, env)
        rc, idx = util.CheckNodesInPod(pod, [gpu], env, raise_on_error=False)
        assert rc in (0, _lib.GMP_ERR_NODE_GPU)          # past the filter (ENXIO here: no driver), never "denied"
        util.UnmountGPU(pod, gpu, False, env)


def test_an_ancestor_that_refuses_the_gpu_is_reported_before_anything_changes(fake_nvml):
    """The kernel runs the device programs of every level and all must agree.  v1 refuses the write to a child's
    devices.allow when the parent lacks the permission; on v2 the ancestors' programs are run on the request first."""
    gpus_cfg = fake_nvml()
    with harness.Target(cgroup2=True) as t:
        env, pod = env_of(t), product_pod()
        u, m, _ = gpus_cfg[0]
        gpu = device.New(m, u)
        pod_level = os.path.dirname(t.cgroup_dir)                       # .../pod<uid>: an ancestor of the container
        o.kernel_attach_rules(pod_level, o.RUNC_DEFAULT_RULES)          # knows nothing of 195:*
        before = attached(t)
        with pytest.raises(ProbeError) as e:
            util.MountGPU(pod, gpu, env)
        assert e.value.code == _lib.GMP_ERR_NODE_DENIED and "ancestor cgroup refuses" in e.value.detail
        assert attached(t) == before and t.nodes() == []
        # the ancestor learns about the GPUs (its own program is edited the same way): now the container-level grant works
        assert c.update_devices(pod_level, [c.Rule(c.DEV_CHAR, 195, -1, c.ACC_ALL)], True) == 1
        util.MountGPU(pod, gpu, env)
        assert t.try_open(m) in NO_DRIVER
        util.UnmountGPU(pod, gpu, False, env)
        assert attached(t) == before
