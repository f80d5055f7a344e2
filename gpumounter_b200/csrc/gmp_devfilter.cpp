// cgroup-v2 device filter: rules <-> eBPF program, an interpreter for it, and the bpf(2) calls that read,
// build and replace the programs attached to a pod's cgroup.  Section 7 of include/gpumounter_probe.h.
//
// No reference counterpart.  The reference grants a device by writing
// "c 195:<minor> rw" to devices.allow of the cgroup-v1 `devices` controller (ref: pkg/util/cgroup/cgroup.go:143-155);
// on the unified hierarchy that file does not exist and the decision is taken by BPF_PROG_TYPE_CGROUP_DEVICE programs
// the container runtime attached at start.  The instruction layout built and recognised here is the one runc emits
// (opencontainers/runc libcontainer/cgroups/devices/devicefilter.go, v1.0.0-rc93+; third-party, absent from the
// reference tree, restated from its published source):
//
//     prologue, 6 instructions:
//         r2 = *(u32*)(r1+0) & 0xFFFF      type
//         r3 = *(u32*)(r1+0) >> 16         access
//         r4 = *(u32*)(r1+4)               major
//         r5 = *(u32*)(r1+8)               minor
//     per rule, LAST rule of the list first:
//         if (r2 != type) goto next
//         [ w1 = w3; w1 &= access; if (r1 != r3) goto next ]     omitted when access == rwm
//         [ if (r4 != major) goto next ]                         omitted when major == -1
//         [ if (r5 != minor) goto next ]                         omitted when minor == -1
//         w0 = allow; exit
//     w0 = default_allow; exit

#include <errno.h>
#include <fcntl.h>
#include <linux/bpf.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <sys/file.h>
#include <sys/syscall.h>
#include <unistd.h>

#include <algorithm>
#include <new>
#include <string>
#include <vector>

#include "gpumounter_probe.h"

namespace gmp {
void set_error(const char* fmt, ...) __attribute__((format(printf, 1, 2)));
void clear_error();
}  // namespace gmp

namespace {

using gmp::clear_error;
using gmp::set_error;

static_assert(sizeof(gmp_bpf_insn) == sizeof(struct bpf_insn), "gmp_bpf_insn mirrors struct bpf_insn");
static_assert((int)GMP_DEV_ACC_MKNOD == (int)BPF_DEVCG_ACC_MKNOD && (int)GMP_DEV_ACC_READ == (int)BPF_DEVCG_ACC_READ &&
                  (int)GMP_DEV_ACC_WRITE == (int)BPF_DEVCG_ACC_WRITE,
              "access bits are the kernel's");
static_assert((int)GMP_DEV_BLOCK == (int)BPF_DEVCG_DEV_BLOCK && (int)GMP_DEV_CHAR == (int)BPF_DEVCG_DEV_CHAR,
              "type values are the kernel's");

constexpr uint8_t kLdxW = BPF_LDX | BPF_MEM | BPF_W;        // 0x61
constexpr uint8_t kAnd32K = BPF_ALU | BPF_AND | BPF_K;      // 0x54
constexpr uint8_t kRsh32K = BPF_ALU | BPF_RSH | BPF_K;      // 0x74
constexpr uint8_t kMov32K = BPF_ALU | BPF_MOV | BPF_K;      // 0xb4
constexpr uint8_t kMov32X = BPF_ALU | BPF_MOV | BPF_X;      // 0xbc
constexpr uint8_t kJneK = BPF_JMP | BPF_JNE | BPF_K;        // 0x55
constexpr uint8_t kJneX = BPF_JMP | BPF_JNE | BPF_X;        // 0x5d
constexpr uint8_t kExit = BPF_JMP | BPF_EXIT;               // 0x95
constexpr int kPrologue = 6;

gmp_bpf_insn insn(uint8_t code, int dst, int src, int off, int32_t imm) {
    gmp_bpf_insn i;
    i.code = code;
    i.regs = (uint8_t)((dst & 0xF) | ((src & 0xF) << 4));
    i.off = (int16_t)off;
    i.imm = imm;
    return i;
}
inline int dst_of(const gmp_bpf_insn& i) { return i.regs & 0xF; }
inline int src_of(const gmp_bpf_insn& i) { return (i.regs >> 4) & 0xF; }
inline bool same(const gmp_bpf_insn& a, const gmp_bpf_insn& b) {
    return a.code == b.code && a.regs == b.regs && a.off == b.off && a.imm == b.imm;
}

bool rule_ok(const gmp_dev_rule& r) {
    if (r.type != GMP_DEV_CHAR && r.type != GMP_DEV_BLOCK) {
        set_error("device rule: type %d is neither char (2) nor block (1)", r.type);
        return false;
    }
    if (r.access == 0 || (r.access & ~(uint32_t)GMP_DEV_ACC_ALL) != 0) {
        set_error("device rule: access bits 0x%x", r.access);
        return false;
    }
    if (r.major < -1 || r.minor < -1) {
        set_error("device rule: major %d minor %d", r.major, r.minor);
        return false;
    }
    return true;
}

int block_length(const gmp_dev_rule& r) {
    return 1 + (r.access != GMP_DEV_ACC_ALL ? 3 : 0) + (r.major >= 0 ? 1 : 0) + (r.minor >= 0 ? 1 : 0) + 2;
}

void emit_prologue(std::vector<gmp_bpf_insn>* p) {
    p->push_back(insn(kLdxW, 2, 1, 0, 0));
    p->push_back(insn(kAnd32K, 2, 0, 0, 0xFFFF));
    p->push_back(insn(kLdxW, 3, 1, 0, 0));
    p->push_back(insn(kRsh32K, 3, 0, 0, 16));
    p->push_back(insn(kLdxW, 4, 1, 4, 0));
    p->push_back(insn(kLdxW, 5, 1, 8, 0));
}

// `tmp` is the scratch register of the access test: r1 in runc's layout (the context pointer is dead after the prologue),
// r6 in the prefix put in front of a foreign program, which still needs r1.
void emit_block(const gmp_dev_rule& r, std::vector<gmp_bpf_insn>* p, int tmp = 1) {
    int left = block_length(r);  // a jump's offset counts from the instruction after it to the next block
    auto jump_off = [&left]() { return --left; };
    p->push_back(insn(kJneK, 2, 0, jump_off(), r.type));
    if (r.access != GMP_DEV_ACC_ALL) {
        p->push_back(insn(kMov32X, tmp, 3, 0, 0));
        --left;
        p->push_back(insn(kAnd32K, tmp, 0, 0, (int32_t)r.access));
        --left;
        p->push_back(insn(kJneX, tmp, 3, jump_off(), 0));
    }
    if (r.major >= 0) p->push_back(insn(kJneK, 4, 0, jump_off(), r.major));
    if (r.minor >= 0) p->push_back(insn(kJneK, 5, 0, jump_off(), r.minor));
    p->push_back(insn(kMov32K, 0, 0, 0, r.allow ? 1 : 0));
    p->push_back(insn(kExit, 0, 0, 0, 0));
}

int build(const gmp_dev_rule* rules, int n, int default_allow, std::vector<gmp_bpf_insn>* out) {
    for (int i = 0; i < n; ++i)
        if (!rule_ok(rules[i])) return GMP_ERR_INVALID;
    out->clear();
    emit_prologue(out);
    for (int i = n - 1; i >= 0; --i) emit_block(rules[i], out);
    out->push_back(insn(kMov32K, 0, 0, 0, default_allow ? 1 : 0));
    out->push_back(insn(kExit, 0, 0, 0, 0));
    return GMP_OK;
}

int unsupported(const char* why, int at) {
    set_error("device program is not in the canonical block form: %s at instruction %d", why, at);
    return GMP_ERR_UNSUPPORTED;
}

// One rule block starting at p[*at]; advances *at past it.
int parse_block(const gmp_bpf_insn* p, int n, int* at_io, int tmp, gmp_dev_rule* out) {
    int at = *at_io;
    gmp_dev_rule r;
    r.major = r.minor = -1;
    r.access = GMP_DEV_ACC_ALL;
    const int begin = at;
    if (at >= n || p[at].code != kJneK || dst_of(p[at]) != 2) return unsupported("block does not start with the type test", at);
    r.type = p[at].imm;
    if (r.type != GMP_DEV_CHAR && r.type != GMP_DEV_BLOCK) return unsupported("unknown device type", at);
    std::vector<int> jumps = {at};
    ++at;
    if (at + 3 <= n && p[at].code == kMov32X && dst_of(p[at]) == tmp && src_of(p[at]) == 3 && p[at + 1].code == kAnd32K &&
        dst_of(p[at + 1]) == tmp && p[at + 2].code == kJneX && dst_of(p[at + 2]) == tmp && src_of(p[at + 2]) == 3) {
        r.access = (uint32_t)p[at + 1].imm;
        if (r.access == 0 || (r.access & ~(uint32_t)GMP_DEV_ACC_ALL) != 0) return unsupported("access mask", at + 1);
        jumps.push_back(at + 2);
        at += 3;
    }
    if (at < n && p[at].code == kJneK && dst_of(p[at]) == 4) {
        if (p[at].imm < 0) return unsupported("major", at);
        r.major = p[at].imm;
        jumps.push_back(at++);
    }
    if (at < n && p[at].code == kJneK && dst_of(p[at]) == 5) {
        if (p[at].imm < 0) return unsupported("minor", at);
        r.minor = p[at].imm;
        jumps.push_back(at++);
    }
    if (at + 2 > n || p[at].code != kMov32K || dst_of(p[at]) != 0 || (p[at].imm != 0 && p[at].imm != 1) || p[at + 1].code != kExit)
        return unsupported("block does not end in 'w0 = 0|1; exit'", at);
    r.allow = p[at].imm;
    at += 2;
    for (int j : jumps)
        if (j + 1 + p[j].off != at) return unsupported("jump does not target the next block", j);
    if (at - begin != block_length(r)) return unsupported("block length", begin);
    *out = r;
    *at_io = at;
    return GMP_OK;
}

// Blocks in program order; the caller reverses them into list order.
int decode(const gmp_bpf_insn* p, int n, std::vector<gmp_dev_rule>* rules, int* default_allow) {
    std::vector<gmp_bpf_insn> pro;
    emit_prologue(&pro);
    if (n < kPrologue + 2) return unsupported("too short", 0);
    for (int i = 0; i < kPrologue; ++i)
        if (!same(p[i], pro[(size_t)i])) return unsupported("prologue differs", i);
    int at = kPrologue;
    rules->clear();
    for (;;) {
        if (at + 2 > n) return unsupported("runs off the end", at);
        if (p[at].code == kMov32K) {  // the final "w0 = default; exit"
            if (dst_of(p[at]) != 0 || (p[at].imm != 0 && p[at].imm != 1) || p[at + 1].code != kExit)
                return unsupported("malformed final block", at);
            if (at + 2 != n) return unsupported("instructions after the final exit", at + 2);
            *default_allow = p[at].imm;
            return GMP_OK;
        }
        gmp_dev_rule r;
        const int rc = parse_block(p, n, &at, 1, &r);
        if (rc != GMP_OK) return rc;
        rules->push_back(r);
    }
}

// ---- prefix in front of a program in some other layout (crun, systemd, hand-written) -------------------------------
// The foreign program is kept verbatim (its jumps are relative, so it can be moved); in front of it go
//     w0 = kWrapMagic; w0 = <number of rules>          two dead stores that mark and size the prefix
//     the six prologue loads                             (r1, the context pointer, is left alone)
//     one block per rule, LAST rule first, scratch r6    a match returns; no match falls through into the original
// Every verifiable program initialises r2..r5 itself before reading them, so leaving them loaded is harmless.
constexpr int32_t kWrapMagic = 0x474D5046;  // "GMPF"
constexpr int kWrapTmp = 6;

bool relocatable(const gmp_bpf_insn* p, int n, int* at) {
    for (int i = 0; i < n; ++i) {
        const uint8_t cls = BPF_CLASS(p[i].code);
        const bool call = (cls == BPF_JMP || cls == BPF_JMP32) && BPF_OP(p[i].code) == BPF_CALL;  // helper ids are rewritten
        const bool wide_imm = p[i].code == (BPF_LD | BPF_IMM | BPF_DW);                           // map references likewise
        if (call || wide_imm) {
            *at = i;
            return false;
        }
    }
    return true;
}

void wrap(const std::vector<gmp_dev_rule>& ours, const gmp_bpf_insn* original, int n, std::vector<gmp_bpf_insn>* out) {
    out->clear();
    if (!ours.empty()) {
        out->push_back(insn(kMov32K, 0, 0, 0, kWrapMagic));
        out->push_back(insn(kMov32K, 0, 0, 0, (int32_t)ours.size()));
        emit_prologue(out);
        for (size_t i = ours.size(); i-- > 0;) emit_block(ours[i], out, kWrapTmp);
    }
    out->insert(out->end(), original, original + n);
}

// Splits a program into (our prefix rules in list order, where the original starts).  No prefix: empty, 0.
int unwrap(const gmp_bpf_insn* p, int n, std::vector<gmp_dev_rule>* ours, int* original_at) {
    ours->clear();
    *original_at = 0;
    if (n < 2 || p[0].code != kMov32K || dst_of(p[0]) != 0 || p[0].imm != kWrapMagic) return GMP_OK;
    if (p[1].code != kMov32K || dst_of(p[1]) != 0 || p[1].imm <= 0 || p[1].imm > 4096) return unsupported("prefix header", 1);
    std::vector<gmp_bpf_insn> pro;
    emit_prologue(&pro);
    if (n < 2 + kPrologue) return unsupported("prefix too short", 2);
    for (int i = 0; i < kPrologue; ++i)
        if (!same(p[2 + i], pro[(size_t)i])) return unsupported("prefix prologue differs", 2 + i);
    int at = 2 + kPrologue;
    for (int k = 0; k < p[1].imm; ++k) {
        gmp_dev_rule r;
        const int rc = parse_block(p, n, &at, kWrapTmp, &r);
        if (rc != GMP_OK) return rc;
        ours->push_back(r);
    }
    std::reverse(ours->begin(), ours->end());
    if (at >= n) return unsupported("prefix with nothing behind it", at);
    *original_at = at;
    return GMP_OK;
}

}  // namespace

// ---- interpreter ------------------------------------------------------------------------------------
namespace {

struct Machine {
    uint64_t r[11] = {0};
    bool is_ctx[11] = {false};
};

bool alu(uint8_t op, bool is64, uint64_t* dst, uint64_t src) {
    uint64_t d = is64 ? *dst : (uint32_t)*dst;
    const uint64_t s = is64 ? src : (uint32_t)src;
    const unsigned shift_mask = is64 ? 63 : 31;
    switch (op) {
        case BPF_ADD: d += s; break;
        case BPF_SUB: d -= s; break;
        case BPF_MUL: d *= s; break;
        case BPF_DIV: d = s ? d / s : 0; break;
        case BPF_MOD: d = s ? d % s : d; break;
        case BPF_OR: d |= s; break;
        case BPF_AND: d &= s; break;
        case BPF_XOR: d ^= s; break;
        case BPF_LSH: d <<= (s & shift_mask); break;
        case BPF_RSH: d >>= (s & shift_mask); break;
        case BPF_ARSH:
            d = is64 ? (uint64_t)((int64_t)d >> (s & shift_mask)) : (uint64_t)(uint32_t)((int32_t)(uint32_t)d >> (s & shift_mask));
            break;
        case BPF_NEG: d = (uint64_t)(-(int64_t)d); break;
        case BPF_MOV: d = s; break;
        default: return false;
    }
    *dst = is64 ? d : (uint32_t)d;
    return true;
}

bool jump_taken(uint8_t op, bool is64, uint64_t a, uint64_t b, bool* ok) {
    *ok = true;
    if (!is64) {
        a = (uint32_t)a;
        b = (uint32_t)b;
    }
    const int64_t sa = is64 ? (int64_t)a : (int64_t)(int32_t)(uint32_t)a;
    const int64_t sb = is64 ? (int64_t)b : (int64_t)(int32_t)(uint32_t)b;
    switch (op) {
        case BPF_JEQ: return a == b;
        case BPF_JNE: return a != b;
        case BPF_JGT: return a > b;
        case BPF_JGE: return a >= b;
        case BPF_JLT: return a < b;
        case BPF_JLE: return a <= b;
        case BPF_JSET: return (a & b) != 0;
        case BPF_JSGT: return sa > sb;
        case BPF_JSGE: return sa >= sb;
        case BPF_JSLT: return sa < sb;
        case BPF_JSLE: return sa <= sb;
        default: *ok = false; return false;
    }
}

int eval(const gmp_bpf_insn* p, int n, int type, uint32_t access, uint32_t major, uint32_t minor, int* allowed) {
    // struct bpf_cgroup_dev_ctx { u32 access_type = access << 16 | type; u32 major; u32 minor; }
    const uint32_t ctx[3] = {(access << 16) | ((uint32_t)type & 0xFFFF), major, minor};
    Machine m;
    m.is_ctx[1] = true;
    int pc = 0;
    for (int steps = 0; steps < 65536; ++steps) {
        if (pc < 0 || pc >= n) {
            set_error("device program: control reaches instruction %d of %d", pc, n);
            return GMP_ERR_INVALID;
        }
        const gmp_bpf_insn& i = p[pc];
        const int d = dst_of(i), s = src_of(i);
        if (d > 10 || s > 10) {
            set_error("device program: register out of range at instruction %d", pc);
            return GMP_ERR_INVALID;
        }
        const uint8_t cls = BPF_CLASS(i.code);
        if (cls == BPF_ALU || cls == BPF_ALU64) {
            const bool is64 = cls == BPF_ALU64;
            const bool from_reg = BPF_SRC(i.code) == BPF_X;
            const uint64_t operand = from_reg ? m.r[s] : (uint64_t)(int64_t)i.imm;
            const bool keeps_ctx = BPF_OP(i.code) == BPF_MOV && from_reg && is64 && m.is_ctx[s];
            if (d == 10 || !alu(BPF_OP(i.code), is64, &m.r[d], operand)) {
                set_error("device program: unsupported ALU instruction 0x%02x at %d", i.code, pc);
                return GMP_ERR_UNSUPPORTED;
            }
            m.is_ctx[d] = keeps_ctx;
            ++pc;
        } else if (cls == BPF_LDX) {
            if (BPF_MODE(i.code) != BPF_MEM || !m.is_ctx[s]) {
                set_error("device program: load that is not from the context at instruction %d", pc);
                return GMP_ERR_UNSUPPORTED;
            }
            const int size = BPF_SIZE(i.code) == BPF_W ? 4 : BPF_SIZE(i.code) == BPF_H ? 2 : BPF_SIZE(i.code) == BPF_B ? 1 : 8;
            if (i.off < 0 || i.off + size > 12 || size == 8) {
                set_error("device program: context load out of range at instruction %d", pc);
                return GMP_ERR_INVALID;
            }
            uint64_t v = 0;
            memcpy(&v, reinterpret_cast<const unsigned char*>(ctx) + i.off, (size_t)size);  // little-endian host
            m.r[d] = v;
            m.is_ctx[d] = false;
            ++pc;
        } else if (cls == BPF_JMP || cls == BPF_JMP32) {
            const uint8_t op = BPF_OP(i.code);
            if (op == BPF_EXIT && cls == BPF_JMP) {
                *allowed = (uint32_t)m.r[0] != 0;  // the kernel looks at the 32-bit return value
                return GMP_OK;
            }
            if (op == BPF_CALL) {
                set_error("device program: call at instruction %d", pc);
                return GMP_ERR_UNSUPPORTED;
            }
            if (op == BPF_JA) {
                pc += 1 + (cls == BPF_JMP ? i.off : i.imm);
                continue;
            }
            bool ok = false;
            const uint64_t operand = BPF_SRC(i.code) == BPF_X ? m.r[s] : (uint64_t)(int64_t)i.imm;
            const bool taken = jump_taken(op, cls == BPF_JMP, m.r[d], operand, &ok);
            if (!ok) {
                set_error("device program: unsupported jump 0x%02x at %d", i.code, pc);
                return GMP_ERR_UNSUPPORTED;
            }
            pc += 1 + (taken ? i.off : 0);
        } else {
            set_error("device program: unsupported instruction class 0x%02x at %d", i.code, pc);
            return GMP_ERR_UNSUPPORTED;
        }
    }
    set_error("device program: step limit reached");
    return GMP_ERR_UNSUPPORTED;
}

}  // namespace

// ---- bpf(2) -----------------------------------------------------------------------------------------
namespace {

long sys_bpf(int cmd, union bpf_attr* attr) { return syscall(SYS_bpf, cmd, attr, sizeof(*attr)); }

struct Fd {
    int fd = -1;
    Fd() = default;
    explicit Fd(int f) : fd(f) {}
    Fd(const Fd&) = delete;
    Fd& operator=(const Fd&) = delete;
    Fd(Fd&& o) noexcept : fd(o.fd) { o.fd = -1; }
    ~Fd() {
        if (fd >= 0) close(fd);
    }
};

int open_cgroup(const char* dir, Fd* out) {
    const int fd = open(dir, O_RDONLY | O_DIRECTORY | O_CLOEXEC);
    if (fd < 0) {
        set_error("open %s: %s", dir, strerror(errno));
        return GMP_ERR_SYS;
    }
    out->fd = fd;
    return GMP_OK;
}

// effective = false: the programs attached to this very cgroup.  true: everything the kernel will run for a task in it,
// i.e. including what ancestors attached (every one of them has to say yes).
int query_programs(int cgroup_fd, std::vector<uint32_t>* ids, uint32_t* attach_flags, bool effective = false) {
    uint32_t buf[64];
    union bpf_attr a;
    memset(&a, 0, sizeof(a));
    a.query.target_fd = (uint32_t)cgroup_fd;
    a.query.attach_type = BPF_CGROUP_DEVICE;
    a.query.query_flags = effective ? BPF_F_QUERY_EFFECTIVE : 0;
    a.query.prog_ids = (uint64_t)(uintptr_t)buf;
    a.query.prog_cnt = 64;
    if (sys_bpf(BPF_PROG_QUERY, &a) != 0) {
        set_error("bpf(BPF_PROG_QUERY, BPF_CGROUP_DEVICE): %s", strerror(errno));
        return GMP_ERR_SYS;
    }
    ids->assign(buf, buf + a.query.prog_cnt);
    if (attach_flags) *attach_flags = a.query.attach_flags;
    return GMP_OK;
}

int program_fd_by_id(uint32_t id, Fd* out) {
    union bpf_attr a;
    memset(&a, 0, sizeof(a));
    a.prog_id = id;
    const long fd = sys_bpf(BPF_PROG_GET_FD_BY_ID, &a);
    if (fd < 0) {
        set_error("bpf(BPF_PROG_GET_FD_BY_ID, %u): %s", id, strerror(errno));
        return GMP_ERR_SYS;
    }
    out->fd = (int)fd;
    return GMP_OK;
}

int read_program(int prog_fd, std::vector<gmp_bpf_insn>* out) {
    struct bpf_prog_info info;
    union bpf_attr a;
    memset(&info, 0, sizeof(info));
    memset(&a, 0, sizeof(a));
    a.info.bpf_fd = (uint32_t)prog_fd;
    a.info.info_len = sizeof(info);
    a.info.info = (uint64_t)(uintptr_t)&info;
    if (sys_bpf(BPF_OBJ_GET_INFO_BY_FD, &a) != 0) {
        set_error("bpf(BPF_OBJ_GET_INFO_BY_FD): %s", strerror(errno));
        return GMP_ERR_SYS;
    }
    if (info.type != BPF_PROG_TYPE_CGROUP_DEVICE) {
        set_error("program %u is of type %u, not a cgroup device program", info.id, info.type);
        return GMP_ERR_INVALID;
    }
    const uint32_t bytes = info.xlated_prog_len;
    if (bytes == 0 || bytes % sizeof(gmp_bpf_insn) != 0) {
        // the kernel hides the instructions from callers without CAP_BPF/CAP_SYS_ADMIN (or under kptr_restrict)
        set_error("program %u: the kernel returned no instructions (xlated_prog_len = %u)", info.id, bytes);
        return GMP_ERR_SYS;
    }
    out->assign(bytes / sizeof(gmp_bpf_insn), gmp_bpf_insn{});
    memset(&info, 0, sizeof(info));
    info.xlated_prog_len = bytes;
    info.xlated_prog_insns = (uint64_t)(uintptr_t)out->data();
    memset(&a, 0, sizeof(a));
    a.info.bpf_fd = (uint32_t)prog_fd;
    a.info.info_len = sizeof(info);
    a.info.info = (uint64_t)(uintptr_t)&info;
    if (sys_bpf(BPF_OBJ_GET_INFO_BY_FD, &a) != 0) {
        set_error("bpf(BPF_OBJ_GET_INFO_BY_FD, instructions): %s", strerror(errno));
        return GMP_ERR_SYS;
    }
    if (info.xlated_prog_len != bytes || info.xlated_prog_insns == 0) {
        set_error("program %u: instructions not returned", info.id);
        return GMP_ERR_SYS;
    }
    return GMP_OK;
}

int load_program(const std::vector<gmp_bpf_insn>& insns, Fd* out) {
    static const char license[] = "Apache";  // what runc states; no GPL-only helper is called
    std::vector<char> log(1 << 16);
    union bpf_attr a;
    memset(&a, 0, sizeof(a));
    a.prog_type = BPF_PROG_TYPE_CGROUP_DEVICE;
    a.insns = (uint64_t)(uintptr_t)insns.data();
    a.insn_cnt = (uint32_t)insns.size();
    a.license = (uint64_t)(uintptr_t)license;
    long fd = sys_bpf(BPF_PROG_LOAD, &a);
    if (fd < 0) {
        const int first_errno = errno;
        a.log_buf = (uint64_t)(uintptr_t)log.data();  // once more, for the verifier's reason
        a.log_size = (uint32_t)log.size();
        a.log_level = 1;
        fd = sys_bpf(BPF_PROG_LOAD, &a);
        if (fd < 0) {
            log.back() = '\0';
            std::string tail = log.data();
            if (tail.size() > 300) tail = tail.substr(tail.size() - 300);
            set_error("bpf(BPF_PROG_LOAD, %zu instructions): %s %s", insns.size(), strerror(first_errno), tail.c_str());
            return GMP_ERR_SYS;
        }
    }
    out->fd = (int)fd;
    return GMP_OK;
}

// BPF_F_REPLACE names the program it expects to find: if another writer has swapped it out since it was read, the
// kernel answers ENOENT and nothing is attached - a compare-and-swap.  *lost_race tells that case apart.
int attach_program(int cgroup_fd, int prog_fd, uint32_t flags, int replace_fd, bool* lost_race = nullptr) {
    if (lost_race) *lost_race = false;
    union bpf_attr a;
    memset(&a, 0, sizeof(a));
    a.target_fd = (uint32_t)cgroup_fd;
    a.attach_bpf_fd = (uint32_t)prog_fd;
    a.attach_type = BPF_CGROUP_DEVICE;
    a.attach_flags = flags;
    if (replace_fd >= 0) {
        a.attach_flags |= BPF_F_REPLACE;
        a.replace_bpf_fd = (uint32_t)replace_fd;
    }
    if (sys_bpf(BPF_PROG_ATTACH, &a) != 0) {
        if (lost_race && replace_fd >= 0 && errno == ENOENT) *lost_race = true;
        set_error("bpf(BPF_PROG_ATTACH, flags 0x%x): %s", a.attach_flags, strerror(errno));
        return GMP_ERR_SYS;
    }
    return GMP_OK;
}

bool same_device(const gmp_dev_rule& a, const gmp_dev_rule& b) {
    return a.type == b.type && a.major == b.major && a.minor == b.minor;
}
bool same_rule(const gmp_dev_rule& a, const gmp_dev_rule& b) {
    return same_device(a, b) && a.access == b.access && (a.allow != 0) == (b.allow != 0);
}

// Appends `r` (the last matching rule wins).  Earlier rules on the same device whose access set is covered by r's can
// never be the last match again and are dropped, so mount/unmount cycles do not grow the program.
void append_rule(std::vector<gmp_dev_rule>* list, const gmp_dev_rule& r) {
    // already there, with nothing after it that speaks about this device: appending would only reorder
    for (size_t k = list->size(); k-- > 0;) {
        const gmp_dev_rule& o = (*list)[k];
        if (same_rule(o, r)) return;
        if (o.type == r.type && (o.major == -1 || r.major == -1 || o.major == r.major) &&
            (o.minor == -1 || r.minor == -1 || o.minor == r.minor))
            break;
    }
    list->erase(std::remove_if(list->begin(), list->end(),
                               [&r](const gmp_dev_rule& o) { return same_device(o, r) && (o.access & ~r.access) == 0; }),
                list->end());
    list->push_back(r);
}

// What the rule list decides for one request: the LAST matching rule wins, else the default.
bool decide(const std::vector<gmp_dev_rule>& list, bool default_allow, int type, int32_t major, int32_t minor, uint32_t q) {
    for (size_t k = list.size(); k-- > 0;) {
        const gmp_dev_rule& o = list[k];
        if (o.type == type && (o.major == -1 || o.major == major) && (o.minor == -1 || o.minor == minor) && (q & ~o.access) == 0)
            return o.allow != 0;
    }
    return default_allow;
}

void edit_rules(std::vector<gmp_dev_rule>* list, bool default_allow, const gmp_dev_rule* rules, int n, bool grant) {
    for (int i = 0; i < n; ++i) {
        gmp_dev_rule r = rules[i];
        r.allow = 1;
        if (grant) {
            append_rule(list, r);
            continue;
        }
        // revoke: take back exactly what a grant appended, if that is still the last word on this device ...
        for (size_t k = list->size(); k-- > 0;) {
            if (!same_device((*list)[k], r)) continue;
            if (same_rule((*list)[k], r)) list->erase(list->begin() + (long)k);
            break;
        }
        // ... and if some request within r.access still gets through (the runtime itself had granted the device, or a
        // wildcard rule covers it), put a deny rule on top - what writing devices.deny does on v1
        bool reachable = false;
        for (uint32_t q = 1; q <= GMP_DEV_ACC_ALL; ++q)
            if ((q & ~r.access) == 0 && r.major >= 0 && r.minor >= 0 && decide(*list, default_allow, r.type, r.major, r.minor, q))
                reachable = true;
        if (r.major < 0 || r.minor < 0) reachable = true;  // wildcard revoke: always explicit
        if (reachable) {
            r.allow = 0;
            append_rule(list, r);
        }
    }
}

}  // namespace

// ---- C ABI ------------------------------------------------------------------------------------------
namespace {

// A grant at the container's level cannot take effect if a program attached further up refuses the device: the kernel
// runs every effective program and all must agree.  The v1 controller says so itself (writing devices.allow of a child
// fails with EPERM when the parent lacks the permission); here the ancestors' programs are run on the request.  Programs
// the interpreter cannot run are not judged.
int ancestors_allow(int cgroup_fd, const std::vector<uint32_t>& own, const gmp_dev_rule* rules, int n) {
    std::vector<uint32_t> effective;
    if (query_programs(cgroup_fd, &effective, nullptr, true) != GMP_OK) {
        clear_error();  // older kernels: no effective query for this attach type; nothing to judge
        return GMP_OK;
    }
    for (uint32_t id : effective) {
        if (std::find(own.begin(), own.end(), id) != own.end()) continue;
        Fd fd;
        std::vector<gmp_bpf_insn> insns;
        if (program_fd_by_id(id, &fd) != GMP_OK || read_program(fd.fd, &insns) != GMP_OK) {
            clear_error();
            continue;
        }
        for (int i = 0; i < n; ++i) {
            if (rules[i].major < 0 || rules[i].minor < 0) continue;
            int allowed = 1;
            if (eval(insns.data(), (int)insns.size(), rules[i].type, rules[i].access, (uint32_t)rules[i].major,
                     (uint32_t)rules[i].minor, &allowed) != GMP_OK) {
                clear_error();
                continue;
            }
            if (!allowed) {
                set_error("device program %u of an ancestor cgroup refuses %c %d:%d: a grant at this level cannot take effect", id,
                          rules[i].type == GMP_DEV_CHAR ? 'c' : 'b', rules[i].major, rules[i].minor);
                return GMP_ERR_NODE_DENIED;
            }
        }
    }
    return GMP_OK;
}

template <typename F>
int guarded(F&& body) {  // no C++ exception crosses the C ABI
    clear_error();
    try {
        return body();
    } catch (const std::bad_alloc&) {
        set_error("out of memory");
        return GMP_ERR_NOMEM;
    } catch (...) {
        set_error("unexpected exception");
        return GMP_ERR_SYS;
    }
}

// Prefix rules in front of a foreign program: same editing as edit_rules(), except that "is the device still reachable"
// is answered by RUNNING the foreign program (its rules are not known), and that an answer it cannot give counts as yes.
void edit_prefix(std::vector<gmp_dev_rule>* ours, const gmp_bpf_insn* original, int n_original, const gmp_dev_rule* rules, int n,
                 bool grant) {
    for (int i = 0; i < n; ++i) {
        gmp_dev_rule r = rules[i];
        r.allow = 1;
        if (grant) {
            append_rule(ours, r);
            continue;
        }
        for (size_t k = ours->size(); k-- > 0;) {
            if (!same_device((*ours)[k], r)) continue;
            if (same_rule((*ours)[k], r)) ours->erase(ours->begin() + (long)k);
            break;
        }
        bool reachable = r.major < 0 || r.minor < 0;
        for (uint32_t q = 1; q <= GMP_DEV_ACC_ALL && !reachable; ++q) {
            if ((q & ~r.access) != 0) continue;
            bool by_prefix = false, decided = false;
            for (size_t k = ours->size(); k-- > 0 && !decided;) {
                const gmp_dev_rule& o = (*ours)[k];
                if (o.type == r.type && (o.major == -1 || o.major == r.major) && (o.minor == -1 || o.minor == r.minor) &&
                    (q & ~o.access) == 0) {
                    by_prefix = o.allow != 0;
                    decided = true;
                }
            }
            if (decided) {
                reachable = by_prefix;
                continue;
            }
            int allowed = 1;
            if (eval(original, n_original, r.type, q, (uint32_t)r.major, (uint32_t)r.minor, &allowed) != GMP_OK) allowed = 1;
            reachable = allowed != 0;
        }
        if (reachable) {
            r.allow = 0;
            append_rule(ours, r);
        }
    }
}

// What one attached program has to become.  *changed = false: it already says what is asked.
int plan_replacement(const std::vector<gmp_bpf_insn>& insns, const gmp_dev_rule* rules, int n, bool grant,
                     std::vector<gmp_bpf_insn>* fresh, bool* changed) {
    *changed = false;
    std::vector<gmp_dev_rule> ours;
    int original_at = 0;
    int rc = unwrap(insns.data(), (int)insns.size(), &ours, &original_at);
    if (rc != GMP_OK) return rc;
    if (original_at == 0) {  // no prefix of ours: runc's layout is edited in place, rule list to rule list
        std::vector<gmp_dev_rule> blocks;
        int default_allow = 0;
        if (decode(insns.data(), (int)insns.size(), &blocks, &default_allow) == GMP_OK) {
            std::vector<gmp_dev_rule> list(blocks.rbegin(), blocks.rend());
            const std::vector<gmp_dev_rule> before = list;
            edit_rules(&list, default_allow != 0, rules, n, grant);
            if (list.size() == before.size() && std::equal(list.begin(), list.end(), before.begin(), same_rule)) return GMP_OK;
            *changed = true;
            return build(list.data(), (int)list.size(), default_allow, fresh);
        }
        clear_error();
    }
    const gmp_bpf_insn* original = insns.data() + original_at;
    const int n_original = (int)insns.size() - original_at;
    int at = 0;
    if (!relocatable(original, n_original, &at)) {
        set_error("device program calls a helper or references a map (instruction %d): it cannot be read back and re-loaded", at);
        return GMP_ERR_UNSUPPORTED;
    }
    edit_prefix(&ours, original, n_original, rules, n, grant);
    wrap(ours, original, n_original, fresh);
    *changed = fresh->size() != insns.size() || !std::equal(fresh->begin(), fresh->end(), insns.begin(), same);
    return GMP_OK;
}

}  // namespace

extern "C" int gmp_devfilter_build(const gmp_dev_rule* rules, int n_rules, int default_allow, gmp_bpf_insn* insns, int cap,
                                   int* n_insns) {
    return guarded([&]() {
        if (n_insns == nullptr || n_rules < 0 || cap < 0 || (n_rules > 0 && rules == nullptr) || (cap > 0 && insns == nullptr)) {
            set_error("bad argument");
            return (int)GMP_ERR_INVALID;
        }
        std::vector<gmp_bpf_insn> p;
        const int rc = build(rules, n_rules, default_allow, &p);
        if (rc != GMP_OK) return rc;
        *n_insns = (int)p.size();
        if ((int)p.size() > cap) {
            set_error("%zu instructions, capacity %d", p.size(), cap);
            return (int)GMP_ERR_BUFFER_TOO_SMALL;
        }
        memcpy(insns, p.data(), p.size() * sizeof(gmp_bpf_insn));
        return (int)GMP_OK;
    });
}

extern "C" int gmp_devfilter_decode(const gmp_bpf_insn* insns, int n_insns, gmp_dev_rule* rules, int cap, int* n_rules,
                                    int* default_allow) {
    return guarded([&]() {
        if (insns == nullptr || n_insns <= 0 || n_rules == nullptr || default_allow == nullptr || cap < 0 ||
            (cap > 0 && rules == nullptr)) {
            set_error("bad argument");
            return (int)GMP_ERR_INVALID;
        }
        std::vector<gmp_dev_rule> blocks;
        const int rc = decode(insns, n_insns, &blocks, default_allow);
        if (rc != GMP_OK) return rc;
        *n_rules = (int)blocks.size();
        if ((int)blocks.size() > cap) {
            set_error("%zu rules, capacity %d", blocks.size(), cap);
            return (int)GMP_ERR_BUFFER_TOO_SMALL;
        }
        std::copy(blocks.rbegin(), blocks.rend(), rules);
        return (int)GMP_OK;
    });
}

extern "C" int gmp_devfilter_eval(const gmp_bpf_insn* insns, int n_insns, int type, uint32_t access, uint32_t major,
                                  uint32_t minor, int* allowed) {
    return guarded([&]() {
        if (insns == nullptr || n_insns <= 0 || allowed == nullptr) {
            set_error("bad argument");
            return (int)GMP_ERR_INVALID;
        }
        return eval(insns, n_insns, type, access, major, minor, allowed);
    });
}

extern "C" int gmp_devfilter_plan(const gmp_bpf_insn* insns, int n_insns, const gmp_dev_rule* rules, int n_rules, int grant,
                                  gmp_bpf_insn* out, int cap, int* n_out) {
    return guarded([&]() {
        if (insns == nullptr || n_insns <= 0 || n_out == nullptr || n_rules < 0 || (n_rules > 0 && rules == nullptr) || cap < 0 ||
            (cap > 0 && out == nullptr)) {
            set_error("bad argument");
            return (int)GMP_ERR_INVALID;
        }
        for (int i = 0; i < n_rules; ++i)
            if (!rule_ok(rules[i])) return (int)GMP_ERR_INVALID;
        const std::vector<gmp_bpf_insn> current(insns, insns + n_insns);
        std::vector<gmp_bpf_insn> fresh;
        bool changed = false;
        const int rc = plan_replacement(current, rules, n_rules, grant != 0, &fresh, &changed);
        if (rc != GMP_OK) return rc;
        const std::vector<gmp_bpf_insn>& result = changed ? fresh : current;
        *n_out = (int)result.size();
        if ((int)result.size() > cap) {
            set_error("%zu instructions, capacity %d", result.size(), cap);
            return (int)GMP_ERR_BUFFER_TOO_SMALL;
        }
        memcpy(out, result.data(), result.size() * sizeof(gmp_bpf_insn));
        return (int)GMP_OK;
    });
}

extern "C" int gmp_cgroup2_device_programs(const char* cgroup_dir, uint32_t* prog_ids, int cap, int* n, uint32_t* attach_flags) {
    return guarded([&]() {
        if (cgroup_dir == nullptr || n == nullptr || cap < 0 || (cap > 0 && prog_ids == nullptr)) {
            set_error("bad argument");
            return (int)GMP_ERR_INVALID;
        }
        Fd cg;
        int rc = open_cgroup(cgroup_dir, &cg);
        if (rc != GMP_OK) return rc;
        std::vector<uint32_t> ids;
        rc = query_programs(cg.fd, &ids, attach_flags);
        if (rc != GMP_OK) return rc;
        *n = (int)ids.size();
        if ((int)ids.size() > cap) {
            set_error("%zu programs, capacity %d", ids.size(), cap);
            return (int)GMP_ERR_BUFFER_TOO_SMALL;
        }
        std::copy(ids.begin(), ids.end(), prog_ids);
        return (int)GMP_OK;
    });
}

extern "C" int gmp_cgroup2_read_program(uint32_t prog_id, gmp_bpf_insn* insns, int cap, int* n_insns) {
    return guarded([&]() {
        if (n_insns == nullptr || cap < 0 || (cap > 0 && insns == nullptr)) {
            set_error("bad argument");
            return (int)GMP_ERR_INVALID;
        }
        Fd fd;
        int rc = program_fd_by_id(prog_id, &fd);
        if (rc != GMP_OK) return rc;
        std::vector<gmp_bpf_insn> p;
        rc = read_program(fd.fd, &p);
        if (rc != GMP_OK) return rc;
        *n_insns = (int)p.size();
        if ((int)p.size() > cap) {
            set_error("%zu instructions, capacity %d", p.size(), cap);
            return (int)GMP_ERR_BUFFER_TOO_SMALL;
        }
        memcpy(insns, p.data(), p.size() * sizeof(gmp_bpf_insn));
        return (int)GMP_OK;
    });
}

extern "C" int gmp_cgroup2_attach_rules(const char* cgroup_dir, const gmp_dev_rule* rules, int n_rules, int default_allow) {
    return guarded([&]() {
        if (cgroup_dir == nullptr || n_rules < 0 || (n_rules > 0 && rules == nullptr)) {
            set_error("bad argument");
            return (int)GMP_ERR_INVALID;
        }
        std::vector<gmp_bpf_insn> p;
        int rc = build(rules, n_rules, default_allow, &p);
        if (rc != GMP_OK) return rc;
        Fd cg, prog;
        rc = open_cgroup(cgroup_dir, &cg);
        if (rc != GMP_OK) return rc;
        rc = load_program(p, &prog);
        if (rc != GMP_OK) return rc;
        return attach_program(cg.fd, prog.fd, BPF_F_ALLOW_MULTI, -1);
    });
}

extern "C" int gmp_cgroup2_update_devices(const char* cgroup_dir, const gmp_dev_rule* rules, int n_rules, int grant,
                                          int* programs) {
    return guarded([&]() {
        if (programs) *programs = 0;
        if (cgroup_dir == nullptr || n_rules < 0 || (n_rules > 0 && rules == nullptr)) {
            set_error("bad argument");
            return (int)GMP_ERR_INVALID;
        }
        for (int i = 0; i < n_rules; ++i)
            if (!rule_ok(rules[i])) return (int)GMP_ERR_INVALID;
        Fd cg;
        int rc = open_cgroup(cgroup_dir, &cg);
        if (rc != GMP_OK) return rc;
        // Editors of this library take turns (advisory lock on the directory, dropped when cg closes); a writer that does
        // not know about the lock - the container runtime - is caught by the compare-and-swap below instead.
        if (flock(cg.fd, LOCK_EX) != 0) {
            set_error("flock %s: %s", cgroup_dir, strerror(errno));
            return (int)GMP_ERR_SYS;
        }
        std::vector<uint32_t> ids;
        for (int attempt = 0;; ++attempt) {
            uint32_t flags = 0;
            bool lost_race = false;
            rc = query_programs(cg.fd, &ids, &flags);
            if (rc != GMP_OK) return rc;
            if (grant != 0 && attempt == 0) {
                rc = ancestors_allow(cg.fd, ids, rules, n_rules);
                if (rc != GMP_OK) return rc;
            }
            // 1. plan and 2. load every replacement before 3. attaching any: a program that cannot be edited, or a
            //    replacement the verifier refuses, leaves the cgroup untouched
            struct Step {
                Fd old_fd, new_fd;
                bool changed = false;
            };
            std::vector<Step> steps(ids.size());
            for (size_t k = 0; k < ids.size(); ++k) {
                rc = program_fd_by_id(ids[k], &steps[k].old_fd);
                if (rc != GMP_OK) return rc;
                std::vector<gmp_bpf_insn> insns, fresh;
                rc = read_program(steps[k].old_fd.fd, &insns);
                if (rc != GMP_OK) return rc;
                rc = plan_replacement(insns, rules, n_rules, grant != 0, &fresh, &steps[k].changed);
                if (rc != GMP_OK) return rc;
                if (!steps[k].changed) continue;
                rc = load_program(fresh, &steps[k].new_fd);
                if (rc != GMP_OK) return rc;
            }
            const bool multi = (flags & BPF_F_ALLOW_MULTI) != 0;
            for (size_t k = 0; k < steps.size(); ++k) {
                if (!steps[k].changed) continue;
                rc = multi ? attach_program(cg.fd, steps[k].new_fd.fd, BPF_F_ALLOW_MULTI, steps[k].old_fd.fd, &lost_race)
                           : attach_program(cg.fd, steps[k].new_fd.fd, flags, -1);  // single-program attachment: attaching replaces
                if (rc == GMP_OK) continue;
                // an attach failed half way (not expected once the loads succeeded): put the earlier programs back
                const std::string why = gmp_last_error();
                for (size_t j = 0; j < k; ++j) {
                    if (!steps[j].changed) continue;
                    if (multi) (void)attach_program(cg.fd, steps[j].old_fd.fd, BPF_F_ALLOW_MULTI, steps[j].new_fd.fd);
                    else (void)attach_program(cg.fd, steps[j].old_fd.fd, flags, -1);
                }
                if (lost_race && attempt < 8) break;  // somebody else edited this cgroup in between: read it again
                set_error("%s (earlier replacements rolled back)", why.c_str());
                return rc;
            }
            if (!lost_race) break;
        }
        if (programs) *programs = (int)ids.size();
        return (int)GMP_OK;
    });
}
