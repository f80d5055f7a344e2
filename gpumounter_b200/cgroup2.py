"""cgroup-v2 device filter over the C ABI (section 7 of include/gpumounter_probe.h).  No reference counterpart: the
reference writes devices.allow / devices.deny of the v1 controller (ref: pkg/util/cgroup/cgroup.go:143-169), which the
unified hierarchy does not have; there the runtime's BPF_PROG_TYPE_CGROUP_DEVICE program decides."""

from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import List, Sequence, Tuple

from ._lib import check, lib

DEV_BLOCK, DEV_CHAR = 1, 2
ACC_MKNOD, ACC_READ, ACC_WRITE, ACC_ALL = 1, 2, 4, 7
NVIDIA_MAJOR = 195  # ref: pkg/device/nvidia.go:37


class gmp_dev_rule(ctypes.Structure):
    _fields_ = [("type", ctypes.c_int32), ("major", ctypes.c_int32), ("minor", ctypes.c_int32),
                ("access", ctypes.c_uint32), ("allow", ctypes.c_int32)]


class gmp_bpf_insn(ctypes.Structure):
    _fields_ = [("code", ctypes.c_uint8), ("regs", ctypes.c_uint8), ("off", ctypes.c_int16), ("imm", ctypes.c_int32)]


@dataclass(frozen=True)
class Rule:
    type: int = DEV_CHAR
    major: int = -1
    minor: int = -1
    access: int = ACC_ALL
    allow: bool = True

    def to_c(self) -> gmp_dev_rule:
        return gmp_dev_rule(self.type, self.major, self.minor, self.access, 1 if self.allow else 0)


Insn = Tuple[int, int, int, int, int]  # (code, dst, src, off, imm)


def _c_rules(rules: Sequence[Rule]):
    return (gmp_dev_rule * max(1, len(rules)))(*[r.to_c() for r in rules])


def _c_insns(insns: Sequence[Insn]):
    return (gmp_bpf_insn * max(1, len(insns)))(*[gmp_bpf_insn(c, (d & 0xF) | ((s & 0xF) << 4), o, i) for c, d, s, o, i in insns])


def _py_insns(arr, n) -> List[Insn]:
    return [(arr[k].code, arr[k].regs & 0xF, arr[k].regs >> 4, arr[k].off, arr[k].imm) for k in range(n)]


def build(rules: Sequence[Rule], default_allow: bool = False) -> List[Insn]:
    arr, n = (gmp_bpf_insn * 8192)(), ctypes.c_int()
    check(lib().gmp_devfilter_build(_c_rules(rules), len(rules), int(default_allow), arr, 8192, ctypes.byref(n)), "gmp_devfilter_build")
    return _py_insns(arr, n.value)


def decode(insns: Sequence[Insn]) -> Tuple[List[Rule], bool]:
    arr, n, d = (gmp_dev_rule * 2048)(), ctypes.c_int(), ctypes.c_int()
    check(lib().gmp_devfilter_decode(_c_insns(insns), len(insns), arr, 2048, ctypes.byref(n), ctypes.byref(d)), "gmp_devfilter_decode")
    return [Rule(arr[k].type, arr[k].major, arr[k].minor, arr[k].access, bool(arr[k].allow)) for k in range(n.value)], bool(d.value)


def evaluate(insns: Sequence[Insn], type: int, access: int, major: int, minor: int) -> bool:
    f = lib().gmp_devfilter_eval
    f.argtypes = [ctypes.POINTER(gmp_bpf_insn), ctypes.c_int, ctypes.c_int, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32,
                  ctypes.POINTER(ctypes.c_int)]
    ok = ctypes.c_int()
    check(f(_c_insns(insns), len(insns), type, access, major, minor, ctypes.byref(ok)), "gmp_devfilter_eval")
    return bool(ok.value)


def plan(insns: Sequence[Insn], rules: Sequence[Rule], grant: bool) -> List[Insn]:
    """The program gmp_cgroup2_update_devices would put in place of `insns` (no kernel involved)."""
    arr, n = (gmp_bpf_insn * 16384)(), ctypes.c_int()
    check(lib().gmp_devfilter_plan(_c_insns(insns), len(insns), _c_rules(rules), len(rules), int(grant), arr, 16384, ctypes.byref(n)),
          "gmp_devfilter_plan")
    return _py_insns(arr, n.value)


def device_programs(cgroup_dir: str) -> Tuple[List[int], int]:
    ids, n, flags = (ctypes.c_uint32 * 64)(), ctypes.c_int(), ctypes.c_uint32()
    check(lib().gmp_cgroup2_device_programs(cgroup_dir.encode(), ids, 64, ctypes.byref(n), ctypes.byref(flags)), "gmp_cgroup2_device_programs")
    return list(ids[:n.value]), flags.value


def read_program(prog_id: int) -> List[Insn]:
    arr, n = (gmp_bpf_insn * 8192)(), ctypes.c_int()
    check(lib().gmp_cgroup2_read_program(ctypes.c_uint32(prog_id), arr, 8192, ctypes.byref(n)), "gmp_cgroup2_read_program")
    return _py_insns(arr, n.value)


def attach_rules(cgroup_dir: str, rules: Sequence[Rule], default_allow: bool = False) -> None:
    check(lib().gmp_cgroup2_attach_rules(cgroup_dir.encode(), _c_rules(rules), len(rules), int(default_allow)), "gmp_cgroup2_attach_rules")


def update_devices(cgroup_dir: str, rules: Sequence[Rule], grant: bool) -> int:
    n = ctypes.c_int()
    check(lib().gmp_cgroup2_update_devices(cgroup_dir.encode(), _c_rules(rules), len(rules), int(grant), ctypes.byref(n)), "gmp_cgroup2_update_devices")
    return n.value


def gpu_rule(minor: int) -> Rule:
    """What `c 195:<minor> rw` says (ref: cgroup.go:145, nvidia.go:37-38): no mknod permission."""
    return Rule(DEV_CHAR, NVIDIA_MAJOR, minor, ACC_READ | ACC_WRITE, True)
