"""Cluster-less stand-in for a pod's container (test + bench infrastructure).

A `Target` is a process in its OWN mount namespace with a private tmpfs mounted on a scratch "dev" directory, so a
device node created for it is visible only through /proc/<pid>/root — the same situation as a container whose /dev
the host cannot see.  Its device cgroup is either a scratch directory tree with a hand-written cgroup.procs (always
available), a real cgroup-v1 `devices` group (when /sys/fs/cgroup/devices is writable), or - `cgroup2=True` - a real
cgroup-v2 directory with the device program a container runtime would have attached at start (runc's default allow
list, built and attached by oracle/gmo_devfilter.py, not by the library under test).
"""

from __future__ import annotations

import os
import shutil
import stat
import subprocess
import tempfile
import uuid as uuidlib
from typing import List, Optional

UID = "0a1b2c3d-1111-2222-3333-444455556666"
CID = "c0ffee" * 10 + "beef"  # 64 hex chars


def can_run() -> Optional[str]:
    if os.geteuid() != 0:
        return "needs root"
    for tool in ("unshare", "nsenter", "mknod", "sh"):
        if shutil.which(tool) is None:
            return f"needs {tool}"
    if subprocess.run(["unshare", "-m", "true"], capture_output=True).returncode != 0:
        return "unshare -m not permitted"
    return None


def can_run_cgroup2() -> Optional[str]:
    from oracle import gmo_devfilter

    return can_run() or (gmo_devfilter.kernel_available() or None)


class Target:
    def __init__(self, real_cgroup: bool = False, qos_dir: str = "besteffort", cgroup2: bool = False, runtime_rules=None,
                 runtime_program=None, container_whitelist: bool = False):
        self.base = tempfile.mkdtemp(prefix="gmp_target_")
        self.devdir = os.path.join(self.base, "dev")
        os.mkdir(self.devdir)
        self.real_cgroup = real_cgroup
        self.cgroup_path = f"/kubepods/{qos_dir}/pod{UID}/{CID}" if qos_dir else f"/kubepods/pod{UID}/{CID}"
        self.proc = None
        self.extra: List[subprocess.Popen] = []
        self.cgroup2 = cgroup2
        self.runtime_rules = runtime_rules
        self.runtime_program = runtime_program  # explicit instructions instead of runc's layout of runtime_rules
        # real cgroup-v1 only: start from what a container runtime gives a pod WITHOUT GPUs (default-deny, the standard
        # character devices, mknod of anything, the NVIDIA control devices) instead of the parent's allow-everything
        self.container_whitelist = container_whitelist
        if cgroup2:
            from oracle import gmo_devfilter

            self.real_cgroup = True
            self.devices_root = f"{gmo_devfilter.cgroup2_mount()}/gmp-{uuidlib.uuid4().hex[:8]}"
        elif real_cgroup:
            self.devices_root = f"/sys/fs/cgroup/devices/gmp-{uuidlib.uuid4().hex[:8]}"
        else:
            self.devices_root = os.path.join(self.base, "cgroup_devices")
        self.cgroup_dir = self.devices_root + self.cgroup_path

    # ---- lifecycle ------------------------------------------------------------------------------------------
    def start(self) -> "Target":
        os.makedirs(self.cgroup_dir)
        if self.cgroup2:
            from oracle import gmo_devfilter

            rules = gmo_devfilter.RUNC_DEFAULT_RULES if self.runtime_rules is None else self.runtime_rules
            if self.runtime_program is not None:
                fd = gmo_devfilter.kernel_load(self.runtime_program)
                try:
                    gmo_devfilter.kernel_attach(self.cgroup_dir, fd)
                finally:
                    os.close(fd)
            else:
                gmo_devfilter.kernel_attach_rules(self.cgroup_dir, rules)
        if self.container_whitelist and self.real_cgroup and not self.cgroup2:
            self._write_v1_whitelist()
        script = f"mount -t tmpfs -o mode=755 tmpfs {self.devdir} && echo ready && exec sleep 600"
        self.proc = subprocess.Popen(["unshare", "-m", "--propagation", "private", "sh", "-c", script],
                                     stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        line = self.proc.stdout.readline()
        if line.strip() != "ready":
            err = self.proc.stderr.read()
            self.stop()
            raise RuntimeError(f"target did not come up: {err}")
        self.register(self.proc.pid)
        return self

    def _write_v1_whitelist(self) -> None:
        def put(name, line):
            with open(os.path.join(self.cgroup_dir, name), "w") as f:
                f.write(line)

        put("devices.deny", "a")
        lines = ["c *:* m", "b *:* m", "c 1:3 rwm", "c 1:5 rwm", "c 1:7 rwm", "c 1:8 rwm", "c 1:9 rwm", "c 5:0 rwm", "c 5:1 rwm",
                 "c 5:2 rwm", "c 136:* rwm", "c 195:255 rw"]          # ... /dev/nvidiactl
        try:
            with open("/proc/devices") as f:                          # nvidia-uvm, nvidia-caps: dynamic majors
                for row in f:
                    parts = row.split()
                    if len(parts) == 2 and parts[1].startswith("nvidia-") and parts[0].isdigit():
                        lines.append(f"c {parts[0]}:* rw")
        except OSError:
            pass
        for line in lines:
            put("devices.allow", line)

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

    def stop(self) -> None:
        for p in self.extra + ([self.proc] if self.proc else []):
            try:
                p.kill()
                p.wait(timeout=5)
            except Exception:
                pass
        if self.real_cgroup:
            d = self.cgroup_dir
            while d.startswith(self.devices_root) and len(d) >= len(self.devices_root):
                try:
                    os.rmdir(d)
                except OSError:
                    pass
                if d == self.devices_root:
                    break
                d = os.path.dirname(d)
        shutil.rmtree(self.base, ignore_errors=True)

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

    def try_open(self, minor: int, flags: int = os.O_RDWR) -> int:
        """errno of open(2) on the target's node for `minor`, attempted by a fresh process that first joins the target's
        cgroup (so the cgroup's device filter judges it): 0 = opened, EPERM = the filter refused, ENXIO/ENODEV = the
        filter let it through and no driver answered (this box has none), ENOENT = no node."""
        code = (f"import os\nopen({os.path.join(self.cgroup_dir, 'cgroup.procs')!r}, 'w').write(str(os.getpid()))\n"
                f"try:\n    os.close(os.open({self.node_path_in_target(minor)!r}, {flags}))\n    print(0)\n"
                f"except OSError as e:\n    print(e.errno)\n")
        import sys

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
