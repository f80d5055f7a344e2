"""TEST INFRASTRUCTURE ONLY (imported by tests/ and the test harness; never by gpumounter_b200/).

Independent restatement, in plain Python, of the cgroup-v2 device filter that section 7 of
include/gpumounter_probe.h implements in C++ (gpumounter_b200/csrc/gmp_devfilter.cpp):

  * `decide`      - what a rule list means: the LAST matching rule wins, else the default;
  * `runc_program`- the instruction sequence opencontainers/runc emits for a rule list
                    (libcontainer/cgroups/devices/devicefilter.go: program.init / appendRule / finalize, rules walked
                    from last to first), v1.0.0-rc93 and later;
  * `interpret`   - a small eBPF interpreter (ALU, ALU64, JMP, JMP32, LDX from the 12-byte context);
  * `kernel_*`    - raw bpf(2) through ctypes, so that the harness can play the container runtime's part (load and
                    attach a program at "container start") without going through the library under test.

Parity: runc is a third-party dependency of container runtimes, not of the reference, and is not present on this
machine: PARITY UNPINNED against runc itself.  What pins the restatement instead: (1) the block layout and instruction
counts of runc's own devicefilter_test.go expectations as remembered (prologue 0-5, a wildcard-'m' block is 6
instructions, a (c,1,3,rwm) block 5), (2) the Linux kernel on this box, which verifies, attaches and ENFORCES the
programs in tests/test_cgroup2_kernel.py (open() of a denied device returns EPERM, of a granted one succeeds; and for
120 random ALU/JMP/LDX programs the kernel's mknod(2) verdict equals `interpret` and the library's gmp_devfilter_eval).
"""

from __future__ import annotations

import ctypes
import os
from typing import List, Sequence, Tuple

DEV_BLOCK, DEV_CHAR = 1, 2
ACC_MKNOD, ACC_READ, ACC_WRITE, ACC_ALL = 1, 2, 4, 7

# (type, major, minor, access, allow); major/minor -1 = wildcard
Rule = Tuple[int, int, int, int, bool]
Insn = Tuple[int, int, int, int, int]  # (code, dst, src, off, imm)

# opcodes, from the eBPF instruction set (class | op | source)
LDX_W, AND32_K, RSH32_K, MOV32_K, MOV32_X, JNE_K, JNE_X, EXIT = 0x61, 0x54, 0x74, 0xB4, 0xBC, 0x55, 0x5D, 0x95


def decide(rules: Sequence[Rule], default_allow: bool, type_: int, access: int, major: int, minor: int) -> bool:
    for t, ma, mi, acc, allow in reversed(rules):
        if t == type_ and ma in (-1, major) and mi in (-1, minor) and (access & ~acc) == 0:
            return allow
    return default_allow


def runc_program(rules: Sequence[Rule], default_allow: bool = False) -> List[Insn]:
    prog: List[Insn] = [
        (LDX_W, 2, 1, 0, 0), (AND32_K, 2, 0, 0, 0xFFFF),   # r2 = type
        (LDX_W, 3, 1, 0, 0), (RSH32_K, 3, 0, 0, 16),       # r3 = access
        (LDX_W, 4, 1, 4, 0),                               # r4 = major
        (LDX_W, 5, 1, 8, 0),                               # r5 = minor
    ]
    for t, ma, mi, acc, allow in reversed(rules):
        block: List[List[int]] = [[JNE_K, 2, 0, None, t]]
        if acc != ACC_ALL:
            block += [[MOV32_X, 1, 3, 0, 0], [AND32_K, 1, 0, 0, acc], [JNE_X, 1, 3, None, 0]]
        if ma >= 0:
            block.append([JNE_K, 4, 0, None, ma])
        if mi >= 0:
            block.append([JNE_K, 5, 0, None, mi])
        block += [[MOV32_K, 0, 0, 0, 1 if allow else 0], [EXIT, 0, 0, 0, 0]]
        for k, ins in enumerate(block):  # "goto next block": relative to the instruction after the jump
            if ins[3] is None:
                ins[3] = len(block) - (k + 1)
        prog += [tuple(i) for i in block]
    prog += [(MOV32_K, 0, 0, 0, 1 if default_allow else 0), (EXIT, 0, 0, 0, 0)]
    return prog


class Unsupported(Exception):
    pass


def _s32(v: int) -> int:
    v &= 0xFFFFFFFF
    return v - (1 << 32) if v & 0x80000000 else v


def _s64(v: int) -> int:
    v &= (1 << 64) - 1
    return v - (1 << 64) if v >> 63 else v


def interpret(prog: Sequence[Insn], type_: int, access: int, major: int, minor: int, max_steps: int = 65536) -> bool:
    ctx = ((access << 16) | (type_ & 0xFFFF)).to_bytes(4, "little") + (major & 0xFFFFFFFF).to_bytes(4, "little") + \
        (minor & 0xFFFFFFFF).to_bytes(4, "little")
    reg = [0] * 11
    is_ctx = [False] * 11
    is_ctx[1] = True
    pc = 0
    for _ in range(max_steps):
        if not 0 <= pc < len(prog):
            raise IndexError(f"control reaches {pc}")
        code, dst, src, off, imm = prog[pc]
        cls = code & 0x07
        if cls in (0x04, 0x07):  # ALU, ALU64
            wide = cls == 0x07
            mask = (1 << 64) - 1 if wide else 0xFFFFFFFF
            operand = (reg[src] if code & 0x08 else _s32(imm)) & mask  # an immediate is sign-extended to the width
            d = reg[dst] & mask
            op = code & 0xF0
            sh = 63 if wide else 31
            if op == 0x00: d = d + operand
            elif op == 0x10: d = d - operand
            elif op == 0x20: d = d * operand
            elif op == 0x30: d = d // operand if operand else 0
            elif op == 0x40: d = d | operand
            elif op == 0x50: d = d & operand
            elif op == 0x60: d = d << (operand & sh)
            elif op == 0x70: d = d >> (operand & sh)
            elif op == 0x80: d = -d
            elif op == 0x90: d = d % operand if operand else d
            elif op == 0xA0: d = d ^ operand
            elif op == 0xB0: d = operand
            elif op == 0xC0: d = (_s64(d) if wide else _s32(d)) >> (operand & sh)
            else: raise Unsupported(hex(code))
            keeps = op == 0xB0 and bool(code & 0x08) and wide and is_ctx[src]
            reg[dst] = d & mask
            is_ctx[dst] = keeps
            pc += 1
        elif cls == 0x01:  # LDX
            if (code & 0xE0) != 0x60 or not is_ctx[src]:
                raise Unsupported("load not from the context")
            size = {0x00: 4, 0x08: 2, 0x10: 1, 0x18: 8}[code & 0x18]
            if size == 8 or off < 0 or off + size > 12:
                raise IndexError("context load out of range")
            reg[dst] = int.from_bytes(ctx[off:off + size], "little")
            is_ctx[dst] = False
            pc += 1
        elif cls in (0x05, 0x06):  # JMP, JMP32
            wide = cls == 0x05
            op = code & 0xF0
            if op == 0x90 and wide:
                return (reg[0] & 0xFFFFFFFF) != 0
            if op == 0x80:
                raise Unsupported("call")
            if op == 0x00:
                pc += 1 + (off if wide else imm)
                continue
            mask = (1 << 64) - 1 if wide else 0xFFFFFFFF
            a = reg[dst] & mask
            b = (reg[src] if code & 0x08 else _s32(imm)) & mask
            sa, sb = (_s64(a), _s64(b)) if wide else (_s32(a), _s32(b))
            taken = {0x10: a == b, 0x20: a > b, 0x30: a >= b, 0x40: (a & b) != 0, 0x50: a != b, 0x60: sa > sb,
                     0x70: sa >= sb, 0xA0: a < b, 0xB0: a <= b, 0xC0: sa < sb, 0xD0: sa <= sb}.get(op)
            if taken is None:
                raise Unsupported(hex(code))
            pc += 1 + (off if taken else 0)
        else:
            raise Unsupported(hex(code))
    raise Unsupported("step limit")


# What runc allows every container by default (libcontainer/specconv AllowedDevices): mknod of anything, the standard
# character devices, pseudo terminals.  The harness attaches this at "container start".
RUNC_DEFAULT_RULES: List[Rule] = [
    (DEV_CHAR, -1, -1, ACC_MKNOD, True), (DEV_BLOCK, -1, -1, ACC_MKNOD, True),
    (DEV_CHAR, 1, 3, ACC_ALL, True), (DEV_CHAR, 1, 8, ACC_ALL, True), (DEV_CHAR, 1, 7, ACC_ALL, True),
    (DEV_CHAR, 5, 0, ACC_ALL, True), (DEV_CHAR, 1, 5, ACC_ALL, True), (DEV_CHAR, 1, 9, ACC_ALL, True),
    (DEV_CHAR, 136, -1, ACC_ALL, True), (DEV_CHAR, 5, 2, ACC_ALL, True), (DEV_CHAR, 10, 200, ACC_ALL, True),
]


# ---- raw bpf(2) ------------------------------------------------------------------------------------------------------
_SYS_BPF = {"x86_64": 321, "aarch64": 280}.get(os.uname().machine, 321)
_BPF_PROG_LOAD, _BPF_PROG_ATTACH = 5, 8
_BPF_PROG_TYPE_CGROUP_DEVICE, _BPF_CGROUP_DEVICE, _BPF_F_ALLOW_MULTI = 15, 6, 2
_libc = ctypes.CDLL(None, use_errno=True)


def _bpf(cmd: int, attr: bytes) -> int:
    buf = ctypes.create_string_buffer(attr.ljust(160, b"\0"), 160)
    r = _libc.syscall(ctypes.c_long(_SYS_BPF), ctypes.c_int(cmd), buf, ctypes.c_uint(160))
    if r < 0:
        e = ctypes.get_errno()
        raise OSError(e, os.strerror(e))
    return r


def pack(prog: Sequence[Insn]) -> bytes:
    out = b""
    for code, dst, src, off, imm in prog:
        out += bytes([code, (dst & 0xF) | ((src & 0xF) << 4)]) + (off & 0xFFFF).to_bytes(2, "little") + \
            (imm & 0xFFFFFFFF).to_bytes(4, "little")
    return out


def kernel_load(prog: Sequence[Insn]) -> int:
    """BPF_PROG_LOAD of a cgroup-device program; returns the program fd."""
    insns = ctypes.create_string_buffer(pack(prog))
    lic = ctypes.create_string_buffer(b"Apache")
    attr = (_BPF_PROG_TYPE_CGROUP_DEVICE.to_bytes(4, "little") + len(prog).to_bytes(4, "little") +
            ctypes.addressof(insns).to_bytes(8, "little") + ctypes.addressof(lic).to_bytes(8, "little"))
    return _bpf(_BPF_PROG_LOAD, attr)


def kernel_attach(cgroup_dir: str, prog_fd: int, flags: int = _BPF_F_ALLOW_MULTI) -> None:
    cg = os.open(cgroup_dir, os.O_RDONLY | os.O_DIRECTORY)
    try:
        attr = (cg.to_bytes(4, "little") + prog_fd.to_bytes(4, "little") + _BPF_CGROUP_DEVICE.to_bytes(4, "little") +
                flags.to_bytes(4, "little"))
        _bpf(_BPF_PROG_ATTACH, attr)
    finally:
        os.close(cg)


def kernel_attach_rules(cgroup_dir: str, rules: Sequence[Rule], default_allow: bool = False) -> None:
    fd = kernel_load(runc_program(rules, default_allow))
    try:
        kernel_attach(cgroup_dir, fd)
    finally:
        os.close(fd)


def cgroup2_mount() -> str:
    """Mount point of a cgroup2 file system on this machine ('' if none)."""
    best = ""
    with open("/proc/mounts") as f:
        for line in f:
            parts = line.split()
            if len(parts) >= 3 and parts[2] == "cgroup2":
                best = parts[1]
    return best


def kernel_available() -> str:
    """'' when a device program can be loaded and a cgroup2 directory made here, else the reason."""
    root = cgroup2_mount()
    if not root:
        return "no cgroup2 mount"
    if os.geteuid() != 0 or not os.access(root, os.W_OK):
        return "cgroup2 mount not writable"
    try:
        os.close(kernel_load(runc_program([])))
    except OSError as e:
        return f"bpf(BPF_PROG_LOAD) refused: {e}"
    return ""


# ---- programs in OTHER layouts, to stand in for runtimes that are not runc ----------------------------------------------
def first_match_program(rules: Sequence[Rule], default_allow: bool = False) -> List[Insn]:
    """A device filter in the style crun and systemd use (both derive from systemd's bpf-devices.c): the same six loads,
    then one block per rule in LIST order (the FIRST matching rule wins), 64-bit moves for the verdict, the access test
    written as "(access & ~rule) != 0 -> next" with JSET.  Not a transcription of either project's emitter - the point is
    a valid, enforceable program whose layout gmp_devfilter_decode does not recognise."""
    MOV64_K, JSET_K = 0xB7, 0x45
    prog: List[Insn] = [
        (LDX_W, 2, 1, 0, 0), (AND32_K, 2, 0, 0, 0xFFFF), (LDX_W, 3, 1, 0, 0), (RSH32_K, 3, 0, 0, 16),
        (LDX_W, 4, 1, 4, 0), (LDX_W, 5, 1, 8, 0),
    ]
    for t, ma, mi, acc, allow in rules:
        block: List[List[int]] = [[JNE_K, 2, 0, None, t]]
        if acc != ACC_ALL:
            block.append([JSET_K, 3, 0, None, (~acc) & ACC_ALL])
        if ma >= 0:
            block.append([JNE_K, 4, 0, None, ma])
        if mi >= 0:
            block.append([JNE_K, 5, 0, None, mi])
        block += [[MOV64_K, 0, 0, 0, 1 if allow else 0], [EXIT, 0, 0, 0, 0]]
        for k, ins in enumerate(block):
            if ins[3] is None:
                ins[3] = len(block) - (k + 1)
        prog += [tuple(i) for i in block]
    prog += [(MOV64_K, 0, 0, 0, 1 if default_allow else 0), (EXIT, 0, 0, 0, 0)]
    return prog


def decide_first_match(rules: Sequence[Rule], default_allow: bool, type_: int, access: int, major: int, minor: int) -> bool:
    for t, ma, mi, acc, allow in rules:
        if t == type_ and ma in (-1, major) and mi in (-1, minor) and (access & ~acc) == 0:
            return allow
    return default_allow
