"""Cluster-less stand-in for a pod's container, test side.

The stand-in itself lives in tools/standin_pod.py (bench.py and tools/mount_latency.py use it too and must not depend on
the test package).  This module adds the one shape that needs the oracle: `cgroup2=True`, a real cgroup-v2 directory
with the device program a container runtime would have attached at start (runc's default allow list, built and
attached by oracle/gmo_devfilter.py, not by the library under test).
"""

from __future__ import annotations

import uuid as uuidlib
from typing import Optional

from tools import standin_pod
from tools.standin_pod import CID, UID, can_run, can_run_pod_like  # noqa: F401  (re-exported for the tests)


def can_run_cgroup2() -> Optional[str]:
    from oracle import gmo_devfilter

    return can_run() or (gmo_devfilter.kernel_available() or None)


class Target(standin_pod.Target):
    def __init__(self, real_cgroup: bool = False, qos_dir: str = "besteffort", cgroup2: bool = False, runtime_rules=None,
                 runtime_program=None, container_whitelist: bool = False, pod_like: bool = False):
        super().__init__(real_cgroup=real_cgroup or cgroup2, qos_dir=qos_dir,
                         container_whitelist=container_whitelist and not cgroup2, pod_like=pod_like)
        self.cgroup2 = cgroup2
        self.runtime_rules = runtime_rules
        self.runtime_program = runtime_program  # explicit instructions instead of runc's layout of runtime_rules
        if cgroup2:
            from oracle import gmo_devfilter

            self.devices_root = f"{gmo_devfilter.cgroup2_mount()}/gmp-{uuidlib.uuid4().hex[:8]}"
            self.cgroup_dir = self.devices_root + self.cgroup_path

    def _prepare_cgroup(self) -> None:
        if not self.cgroup2:
            return super()._prepare_cgroup()
        from oracle import gmo_devfilter

        rules = gmo_devfilter.RUNC_DEFAULT_RULES if self.runtime_rules is None else self.runtime_rules
        if self.runtime_program is not None:
            fd = gmo_devfilter.kernel_load(self.runtime_program)
            try:
                gmo_devfilter.kernel_attach(self.cgroup_dir, fd)
            finally:
                import os

                os.close(fd)
        else:
            gmo_devfilter.kernel_attach_rules(self.cgroup_dir, rules)
