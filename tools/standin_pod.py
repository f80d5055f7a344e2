"""Cluster-less stand-in for a pod's container (test + bench infrastructure; imports nothing from oracle/ or tests/).

A `Target` is a process in its OWN mount namespace.  Two shapes:

* scratch (default): a private tmpfs on a scratch "dev" directory, so a device node created for it is visible only
  through /proc/<pid>/root — enough for the mount mechanics, but CUDA in that "pod" would still open the host's /dev.
* `pod_like=True`: the private tmpfs is mounted ON /dev of the target's mount namespace and populated the way a
  container runtime populates it for a pod started with NVIDIA_VISIBLE_DEVICES=none (ref: docs/guide/QuickStart.md:9,
  26-28): the standard character devices and NVIDIA's control nodes (nvidiactl, nvidia-uvm, nvidia-uvm-tools,
  nvidia-caps/*), and NO /dev/nvidia<N>.  Its device cgroup is a real cgroup-v1 `devices` group carrying the whitelist
  such a container gets (default deny; mknod of anything; read/write only for the standard devices and the control
  nodes).  A process that joins that group and that mount namespace is then in the pod's position: the only way for
  it to reach a GPU is the node + the `devices.allow` line the mount path under test creates
  (ref: pkg/util/namespace/namespace.go:167-177, pkg/util/cgroup/cgroup.go:143-155).  No CUDA_VISIBLE_DEVICES involved.

The device cgroup is otherwise a scratch directory tree with a hand-written cgroup.procs (always available) or a real
cgroup-v1 `devices` group (when /sys/fs/cgroup/devices is writable).  tests/harness.py adds the cgroup-v2 shape on top.
"""

from __future__ import annotations

import errno
import os
import shutil
import stat
import subprocess
import sys
import tempfile
import uuid as uuidlib
from typing import Dict, List, Optional, Tuple

UID = "0a1b2c3d-1111-2222-3333-444455556666"
CID = "c0ffee" * 10 + "beef"  # 64 hex chars

V1_DEVICES_ROOT = "/sys/fs/cgroup/devices"

# what runc's default device list grants every container: null zero full random urandom tty console ptmx, pts/*
STANDARD_NODES = [("null", 1, 3), ("zero", 1, 5), ("full", 1, 7), ("random", 1, 8), ("urandom", 1, 9), ("tty", 5, 0)]
STANDARD_WHITELIST = ["c *:* m", "b *:* m", "c 1:3 rwm", "c 1:5 rwm", "c 1:7 rwm", "c 1:8 rwm", "c 1:9 rwm", "c 5:0 rwm",
                      "c 5:1 rwm", "c 5:2 rwm", "c 136:* rwm"]
NVIDIA_MAJOR = 195  # ref: pkg/device/nvidia.go:37
NVIDIA_CTL_MINOR = 255


def can_run() -> Optional[str]:
    if os.geteuid() != 0:
        return "needs root"
    for tool in ("unshare", "nsenter", "mknod", "sh"):
        if shutil.which(tool) is None:
            return f"needs {tool}"
    if subprocess.run(["unshare", "-m", "true"], capture_output=True).returncode != 0:
        return "unshare -m not permitted"
    return None


def can_run_pod_like() -> Optional[str]:
    why = can_run()
    if why:
        return why
    if not os.path.isdir(V1_DEVICES_ROOT) or not os.access(os.path.join(V1_DEVICES_ROOT, "cgroup.procs"), os.W_OK):
        return f"no writable cgroup-v1 devices hierarchy at {V1_DEVICES_ROOT}"
    probe_dir = os.path.join(V1_DEVICES_ROOT, f"gmp-can-{uuidlib.uuid4().hex[:8]}")
    try:
        os.mkdir(probe_dir)
        os.rmdir(probe_dir)
    except OSError as e:
        return f"cannot create a devices cgroup: {e}"
    return None


def nvidia_dynamic_majors() -> Dict[str, int]:
    """Dynamic character majors the NVIDIA driver registered (nvidia-uvm, nvidia-caps, ...), from /proc/devices."""
    out: Dict[str, int] = {}
    try:
        with open("/proc/devices") as f:
            for row in f:
                parts = row.split()
                if len(parts) == 2 and parts[1].startswith("nvidia-") and parts[0].isdigit() and int(parts[0]) != NVIDIA_MAJOR:
                    out[parts[1]] = int(parts[0])
    except OSError:
        pass
    return out


def host_control_nodes() -> List[Tuple[str, int, int]]:
    """(path relative to /dev, major, minor) of the control nodes nvidia-container-runtime gives every pod, as they
    exist on this host.  On a driver-less box: nvidiactl alone, with the constant numbers, so that the mechanics can
    still be exercised (open() then answers ENXIO)."""
    found: List[Tuple[str, int, int]] = []
    for rel in ("nvidiactl", "nvidia-uvm", "nvidia-uvm-tools"):
        try:
            st = os.stat(os.path.join("/dev", rel))
        except OSError:
            continue
        if stat.S_ISCHR(st.st_mode):
            found.append((rel, os.major(st.st_rdev), os.minor(st.st_rdev)))
    try:
        for name in sorted(os.listdir("/dev/nvidia-caps")):
            st = os.stat(os.path.join("/dev/nvidia-caps", name))
            if stat.S_ISCHR(st.st_mode):
                found.append((f"nvidia-caps/{name}", os.major(st.st_rdev), os.minor(st.st_rdev)))
    except OSError:
        pass
    if not any(rel == "nvidiactl" for rel, _, _ in found):
        found.insert(0, ("nvidiactl", NVIDIA_MAJOR, NVIDIA_CTL_MINOR))
    return found


class Target:
    def __init__(self, real_cgroup: bool = False, qos_dir: str = "besteffort", container_whitelist: bool = False,
                 pod_like: bool = False):
        self.base = tempfile.mkdtemp(prefix="gmp_target_")
        self.pod_like = pod_like
        if pod_like:
            real_cgroup = True
            container_whitelist = True
            self.devdir = "/dev"
        else:
            self.devdir = os.path.join(self.base, "dev")
            os.mkdir(self.devdir)
        self.real_cgroup = real_cgroup
        self.cgroup_path = f"/kubepods/{qos_dir}/pod{UID}/{CID}" if qos_dir else f"/kubepods/pod{UID}/{CID}"
        self.proc = None
        self.extra: List[subprocess.Popen] = []
        self.cgroup2 = False
        # real cgroup-v1 only: start from what a container runtime gives a pod WITHOUT GPUs (default-deny, the standard
        # character devices, mknod of anything, the NVIDIA control devices) instead of the parent's allow-everything
        self.container_whitelist = container_whitelist
        self.whitelist_refused: List[str] = []  # lines the parent cgroup did not let us grant (it lacks them itself)
        if real_cgroup:
            self.devices_root = f"{V1_DEVICES_ROOT}/gmp-{uuidlib.uuid4().hex[:8]}"
        else:
            self.devices_root = os.path.join(self.base, "cgroup_devices")
        self.cgroup_dir = self.devices_root + self.cgroup_path

    # ---- lifecycle ------------------------------------------------------------------------------------------
    def _prepare_cgroup(self) -> None:
        """Hook for subclasses (tests/harness.py attaches a cgroup-v2 device program here)."""
        if self.container_whitelist and self.real_cgroup:
            self._write_v1_whitelist()

    def _mount_script(self) -> str:
        if not self.pod_like:
            return f"mount -t tmpfs -o mode=755 tmpfs {self.devdir} && echo ready && exec sleep 600"
        lines = ["set -e", "mount -t tmpfs -o mode=755,size=4m tmpfs /dev"]
        for name, major, minor in STANDARD_NODES:
            lines.append(f"mknod -m 666 /dev/{name} c {major} {minor}")
        lines += ["mkdir -m 1777 /dev/shm", "mkdir /dev/pts", "ln -s /proc/self/fd /dev/fd",
                  "ln -s /proc/self/fd/0 /dev/stdin", "ln -s /proc/self/fd/1 /dev/stdout", "ln -s /proc/self/fd/2 /dev/stderr"]
        made_dirs = set()
        for rel, major, minor in host_control_nodes():
            d = os.path.dirname(rel)
            if d and d not in made_dirs:
                lines.append(f"mkdir -p /dev/{d}")
                made_dirs.add(d)
            lines.append(f"mknod -m 666 /dev/{rel} c {major} {minor}")
        lines += ["echo ready", "exec sleep 600"]
        return "\n".join(lines)

    def start(self) -> "Target":
        os.makedirs(self.cgroup_dir)
        self._prepare_cgroup()
        self.proc = subprocess.Popen(["unshare", "-m", "--propagation", "private", "sh", "-c", self._mount_script()],
                                     stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        line = self.proc.stdout.readline()
        if line.strip() != "ready":
            err = self.proc.stderr.read()
            self.stop()
            raise RuntimeError(f"target did not come up: {err}")
        self.register(self.proc.pid)
        return self

    def whitelist_lines(self) -> List[str]:
        lines = list(STANDARD_WHITELIST) + [f"c {NVIDIA_MAJOR}:{NVIDIA_CTL_MINOR} rw"]  # ... /dev/nvidiactl
        for _name, major in sorted(nvidia_dynamic_majors().items()):                     # nvidia-uvm, nvidia-caps, ...
            lines.append(f"c {major}:* rw")
        return lines

    def _write_v1_whitelist(self) -> None:
        def put(name, line):
            with open(os.path.join(self.cgroup_dir, name), "w") as f:
                f.write(line)

        put("devices.deny", "a")
        for line in self.whitelist_lines():
            try:
                put("devices.allow", line)
            except OSError as e:
                # a child may only be granted what its parent holds; on a box that is itself a restricted container some
                # of the standard lines are not ours to give
                if e.errno not in (errno.EPERM, errno.EACCES):
                    raise
                self.whitelist_refused.append(line)

    def register(self, pid: int) -> None:
        procs = os.path.join(self.cgroup_dir, "cgroup.procs")
        if self.real_cgroup:
            with open(procs, "w") as f:
                f.write(f"{pid}\n")
        else:
            with open(procs, "a") as f:
                f.write(f"{pid}\n")

    def spawn_member(self) -> subprocess.Popen:
        """Another process of the 'container' (same cgroup), e.g. one that holds the GPU."""
        p = subprocess.Popen(["sleep", "600"])
        self.extra.append(p)
        self.register(p.pid)
        return p

    def spawn_in_pod(self, argv: List[str], **popen_kw) -> subprocess.Popen:
        """A process started the way the container's own processes are: it joins the target's device cgroup first and
        then its mount namespace, and only then execs `argv` (which resolves inside that namespace)."""
        procs = os.path.join(self.cgroup_dir, "cgroup.procs")
        ns = f"/proc/{self.pid}/ns/mnt"
        mode = "w" if self.real_cgroup else "a"
        boot = ("import os, sys\n"
                f"open({procs!r}, {mode!r}).write(str(os.getpid()) + '\\n')\n"
                f"fd = os.open({ns!r}, os.O_RDONLY)\n"
                "os.setns(fd, os.CLONE_NEWNS)\n"
                "os.close(fd)\n"
                "os.execvp(sys.argv[1], sys.argv[1:])\n")
        p = subprocess.Popen([sys.executable, "-c", boot] + list(argv), **popen_kw)
        self.extra.append(p)
        return p

    def stop(self) -> None:
        for p in self.extra + ([self.proc] if self.proc else []):
            try:
                p.kill()
                p.wait(timeout=5)
            except Exception:
                pass
            for pipe in (p.stdin, p.stdout, p.stderr):  # or the descriptors live on until the Popen object is collected
                try:
                    if pipe is not None:
                        pipe.close()
                except Exception:
                    pass
        if self.real_cgroup:
            self._remove_cgroup_dirs()
        shutil.rmtree(self.base, ignore_errors=True)

    def _remove_cgroup_dirs(self) -> None:
        d = self.cgroup_dir
        while d.startswith(self.devices_root) and len(d) >= len(self.devices_root):
            try:
                os.rmdir(d)
            except OSError:
                pass
            if d == self.devices_root:
                break
            d = os.path.dirname(d)

    def __enter__(self):
        return self.start()

    def __exit__(self, *exc):
        self.stop()

    # ---- views ------------------------------------------------------------------------------------------------
    @property
    def pid(self) -> int:
        return self.proc.pid

    @property
    def dev_prefix(self) -> str:
        return os.path.join(self.devdir, "nvidia")

    def node_path_in_target(self, minor: int) -> str:
        return f"/proc/{self.pid}/root{self.devdir}/nvidia{minor}"

    def node(self, minor: int):
        """(major, minor, mode) of the node as the target sees it, or None."""
        try:
            st = os.stat(self.node_path_in_target(minor))
        except FileNotFoundError:
            return None
        assert stat.S_ISCHR(st.st_mode), "not a character device"
        return os.major(st.st_rdev), os.minor(st.st_rdev), stat.S_IMODE(st.st_mode)

    def nodes(self) -> List[str]:
        return sorted(os.listdir(f"/proc/{self.pid}/root{self.devdir}"))

    def gpu_nodes(self) -> List[str]:
        """The /dev/nvidia<N> nodes only (a pod-like /dev also holds the standard and the control nodes)."""
        return [x for x in self.nodes() if x.startswith("nvidia") and x[6:].isdigit()]

    def try_open(self, minor: int, flags: int = os.O_RDWR) -> int:
        """errno of open(2) on the target's node for `minor`, attempted by a fresh process that first joins the target's
        cgroup (so the cgroup's device filter judges it): 0 = opened, EPERM = the filter refused, ENXIO/ENODEV = the
        filter let it through and no driver answered (this box has none), ENOENT = no node."""
        code = (f"import os\nopen({os.path.join(self.cgroup_dir, 'cgroup.procs')!r}, 'w').write(str(os.getpid()))\n"
                f"try:\n    os.close(os.open({self.node_path_in_target(minor)!r}, {flags}))\n    print(0)\n"
                f"except OSError as e:\n    print(e.errno)\n")
        out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=30)
        return int(out.stdout.strip())

    def host_sees(self) -> List[str]:
        return sorted(os.listdir(self.devdir))

    def read(self, name: str) -> Optional[str]:
        try:
            with open(os.path.join(self.cgroup_dir, name)) as f:
                return f.read()
        except FileNotFoundError:
            return None
