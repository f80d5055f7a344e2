"""cgroup-v2 device filter (header section 7; no reference counterpart): the C++ builder, decoder
and interpreter against the independent Python restatement in oracle/gmo_devfilter.py and against hand-written
instruction lists."""

import ctypes

import pytest
from hypothesis import given, settings, strategies as st

from gpumounter_b200 import ProbeError, _lib, cgroup2 as c, lib
from oracle import gmo_devfilter as o

PROLOGUE = [
    (0x61, 2, 1, 0, 0), (0x54, 2, 0, 0, 0xFFFF),  # r2 = *(u32*)(r1+0); w2 &= 0xFFFF
    (0x61, 3, 1, 0, 0), (0x74, 3, 0, 0, 16),      # r3 = *(u32*)(r1+0); w3 >>= 16
    (0x61, 4, 1, 4, 0),                           # r4 = *(u32*)(r1+4)
    (0x61, 5, 1, 8, 0),                           # r5 = *(u32*)(r1+8)
]


def as_rules(tuples):
    return [c.Rule(*t) for t in tuples]


def as_tuples(rules):
    return [(r.type, r.major, r.minor, r.access, r.allow) for r in rules]


def test_empty_rule_list_is_prologue_plus_reject():
    # runc's TestDeviceFilter_Nil: eight instructions, "return 0 (reject)" at 6
    assert c.build([]) == PROLOGUE + [(0xB4, 0, 0, 0, 0), (0x95, 0, 0, 0, 0)]
    assert c.build([], default_allow=True)[-2:] == [(0xB4, 0, 0, 0, 1), (0x95, 0, 0, 0, 0)]


def test_hand_assembled_program():
    """Written out by hand from the layout in gmp_devfilter.cpp's header comment, not generated: the list
    [c *:* m allow, c 1:3 rwm allow, c 195:3 rw allow] is emitted last rule first."""
    rules = [c.Rule(c.DEV_CHAR, -1, -1, c.ACC_MKNOD), c.Rule(c.DEV_CHAR, 1, 3, c.ACC_ALL), c.gpu_rule(3)]
    expected = PROLOGUE + [
        # block 0: c 195:3 rw  (8 instructions; a jump's offset counts from the instruction after it)
        (0x55, 2, 0, 7, 2),      # if r2 != CHAR goto block 1
        (0xBC, 1, 3, 0, 0),      # w1 = w3
        (0x54, 1, 0, 0, 6),      # w1 &= READ|WRITE
        (0x5D, 1, 3, 4, 0),      # if r1 != r3 goto block 1   (the request has a bit outside rw)
        (0x55, 4, 0, 3, 195),    # if r4 != 195 goto block 1
        (0x55, 5, 0, 2, 3),      # if r5 != 3 goto block 1
        (0xB4, 0, 0, 0, 1),      # w0 = 1
        (0x95, 0, 0, 0, 0),      # exit
        # block 1: c 1:3 rwm  (5 instructions, no access test)
        (0x55, 2, 0, 4, 2), (0x55, 4, 0, 3, 1), (0x55, 5, 0, 2, 3), (0xB4, 0, 0, 0, 1), (0x95, 0, 0, 0, 0),
        # block 2: c *:* m  (6 instructions, no major/minor test)
        (0x55, 2, 0, 5, 2), (0xBC, 1, 3, 0, 0), (0x54, 1, 0, 0, 1), (0x5D, 1, 3, 2, 0), (0xB4, 0, 0, 0, 1), (0x95, 0, 0, 0, 0),
        # default: reject
        (0xB4, 0, 0, 0, 0), (0x95, 0, 0, 0, 0),
    ]
    assert len(expected) == 6 + 8 + 5 + 6 + 2
    assert c.build(rules) == expected == o.runc_program(as_tuples(rules))
    assert c.decode(expected) == (rules, False)
    # "rw" does not include mknod: the GPU rule must not let the pod create other nodes of that device
    assert c.evaluate(expected, c.DEV_CHAR, c.ACC_READ | c.ACC_WRITE, 195, 3)
    assert c.evaluate(expected, c.DEV_CHAR, c.ACC_MKNOD, 195, 3)          # ... but the runtime's wildcard 'm' rule does
    assert not c.evaluate(expected, c.DEV_CHAR, c.ACC_READ | c.ACC_MKNOD, 195, 3)
    assert not c.evaluate(expected, c.DEV_CHAR, c.ACC_READ, 195, 4)
    assert not c.evaluate(expected, c.DEV_BLOCK, c.ACC_READ, 195, 3)


def test_runc_default_allow_list_round_trips_and_means_what_it_says():
    prog = c.build(as_rules(o.RUNC_DEFAULT_RULES))
    assert prog == o.runc_program(o.RUNC_DEFAULT_RULES)
    rules, default = c.decode(prog)
    assert as_tuples(rules) == o.RUNC_DEFAULT_RULES and default is False
    for type_ in (c.DEV_BLOCK, c.DEV_CHAR):
        for access in range(1, 8):
            for major, minor in [(1, 3), (1, 5), (1, 4), (5, 2), (136, 0), (136, 77), (195, 0), (195, 255), (10, 200), (8, 0)]:
                want = o.decide(o.RUNC_DEFAULT_RULES, False, type_, access, major, minor)
                assert c.evaluate(prog, type_, access, major, minor) == want == o.interpret(prog, type_, access, major, minor)


rule_st = st.tuples(st.sampled_from([1, 2]), st.sampled_from([-1, 1, 5, 195, 0x7FFFFFFF]), st.sampled_from([-1, 0, 3, 7, 255, 65536]),
                    st.integers(1, 7), st.booleans())


@settings(max_examples=200, deadline=None)
@given(st.lists(rule_st, max_size=12), st.booleans())
def test_random_rule_lists_build_decode_and_evaluate_like_the_oracle(rules, default_allow):
    prog = c.build(as_rules(rules), default_allow)
    assert prog == o.runc_program(rules, default_allow)
    got, got_default = c.decode(prog)
    assert as_tuples(got) == rules and got_default == default_allow
    majors = sorted({r[1] for r in rules if r[1] >= 0} | {2})
    minors = sorted({r[2] for r in rules if r[2] >= 0} | {9})
    for type_ in (1, 2):
        for access in (1, 2, 3, 4, 6, 7):
            for major in majors:
                for minor in minors:
                    want = o.decide(rules, default_allow, type_, access, major, minor)
                    assert c.evaluate(prog, type_, access, major, minor) == want
                    assert o.interpret(prog, type_, access, major, minor) == want


def test_bad_rules_are_refused():
    for bad in (c.Rule(3, 1, 3, 7), c.Rule(c.DEV_CHAR, 1, 3, 0), c.Rule(c.DEV_CHAR, 1, 3, 8), c.Rule(c.DEV_CHAR, -2, 3, 7)):
        with pytest.raises(ProbeError) as e:
            c.build([bad])
        assert e.value.code == _lib.GMP_ERR_INVALID


def test_capacity_is_reported():
    rules = (c.gmp_dev_rule * 1)(c.gpu_rule(0).to_c())
    n = ctypes.c_int()
    buf = (c.gmp_bpf_insn * 4)()
    assert lib().gmp_devfilter_build(rules, 1, 0, buf, 4, ctypes.byref(n)) == _lib.GMP_ERR_BUFFER_TOO_SMALL
    assert n.value == 6 + 8 + 2


@pytest.mark.parametrize("mutate, why", [
    (lambda p: p[:3] + [(0x74, 3, 0, 0, 8)] + p[4:], "prologue"),                       # a different shift
    (lambda p: p[:6] + [(0x55, 2, 0, 3, 2)] + p[7:], "jump"),                           # a jump into the middle of a block
    (lambda p: p + [(0x95, 0, 0, 0, 0)], "after"),                                      # trailing instruction
    (lambda p: p[:-2], "end"),                                                          # no final block
    (lambda p: p[:6] + [(0x15, 2, 0, 4, 2)] + p[7:], "type test"),                      # JEQ instead of JNE
    (lambda p: p[:10] + [(0xB4, 0, 0, 0, 2)] + p[11:], "w0"),                           # returns 2
    (lambda p: [(0xB7, 0, 0, 0, 1), (0x95, 0, 0, 0, 0)], "short"),                      # "r0 = 1; exit" (allow everything)
])
def test_programs_outside_the_block_form_are_unsupported(mutate, why):
    prog = c.build([c.Rule(c.DEV_CHAR, 1, 3, c.ACC_ALL), c.Rule(c.DEV_CHAR, 1, 5, c.ACC_ALL)])
    with pytest.raises(ProbeError) as e:
        c.decode(mutate(list(prog)))
    assert e.value.code == _lib.GMP_ERR_UNSUPPORTED and "canonical block form" in e.value.detail, why


def test_interpreter_handles_general_programs_and_refuses_what_it_cannot_run():
    # a crun-style program does not use runc's layout; the interpreter still runs it:
    #   r2 = *(u32*)(r1+4); if r2 == 195 goto +2; r0 = 0; exit; r0 = 1; exit
    prog = [(0x61, 2, 1, 4, 0), (0x15, 2, 0, 2, 195), (0xB7, 0, 0, 0, 0), (0x95, 0, 0, 0, 0), (0xB7, 0, 0, 0, 1), (0x95, 0, 0, 0, 0)]
    assert c.evaluate(prog, 2, 2, 195, 0) and not c.evaluate(prog, 2, 2, 1, 3)
    assert o.interpret(prog, 2, 2, 195, 0) and not o.interpret(prog, 2, 2, 1, 3)
    # 64-bit ALU, a byte load, JSET, JA
    prog = [(0x71, 2, 1, 2, 0),            # r2 = *(u8*)(r1+2)  = low byte of the access bits
            (0x45, 2, 0, 1, 4),            # if r2 & WRITE goto +1
            (0x05, 0, 0, 2, 0),            # goto +2
            (0xB7, 0, 0, 0, 0), (0x95, 0, 0, 0, 0),   # writes: reject
            (0xB7, 0, 0, 0, 7), (0x95, 0, 0, 0, 0)]   # everything else: allow (any non-zero r0)
    for access in range(1, 8):
        assert c.evaluate(prog, 2, access, 1, 3) == (not access & 4) == o.interpret(prog, 2, access, 1, 3)
    for bad, code in [([(0x85, 0, 0, 0, 1), (0x95, 0, 0, 0, 0)], _lib.GMP_ERR_UNSUPPORTED),        # call
                      ([(0x61, 2, 1, 12, 0), (0x95, 0, 0, 0, 0)], _lib.GMP_ERR_INVALID),           # load past the context
                      ([(0x61, 2, 3, 0, 0), (0x95, 0, 0, 0, 0)], _lib.GMP_ERR_UNSUPPORTED),        # load from a scalar
                      ([(0x05, 0, 0, 5, 0), (0x95, 0, 0, 0, 0)], _lib.GMP_ERR_INVALID),            # jump out of the program
                      ([(0x05, 0, 0, -1, 0)], _lib.GMP_ERR_UNSUPPORTED),                           # endless loop: step limit
                      ([(0x18, 0, 0, 0, 0), (0, 0, 0, 0, 0), (0x95, 0, 0, 0, 0)], _lib.GMP_ERR_UNSUPPORTED)]:  # ld_imm64
        with pytest.raises(ProbeError) as e:
            c.evaluate(bad, 2, 2, 1, 3)
        assert e.value.code == code, bad


# ---- programs in a layout the decoder does not know: kept verbatim behind a prefix ----------------------------------
GPU_RW = c.ACC_READ | c.ACC_WRITE


def requests(extra=()):
    for type_ in (1, 2):
        for access in range(1, 8):
            for major, minor in [(1, 3), (1, 5), (195, 0), (195, 3), (195, 4), (136, 9), (8, 0), *extra]:
                yield type_, access, major, minor


def test_foreign_program_gets_a_prefix_and_comes_back_byte_for_byte():
    original = o.first_match_program(o.RUNC_DEFAULT_RULES)
    with pytest.raises(ProbeError):
        c.decode(original)                                      # not runc's layout
    granted = c.plan(original, [c.gpu_rule(3), c.gpu_rule(4)], True)
    assert granted[0] == (0xB4, 0, 0, 0, 0x474D5046) and granted[1] == (0xB4, 0, 0, 0, 2)   # marker, two rules
    assert granted[-len(original):] == original                 # the runtime's program, untouched, behind the prefix
    for type_, access, major, minor in requests():
        before = o.interpret(original, type_, access, major, minor)
        mine = type_ == c.DEV_CHAR and major == 195 and minor in (3, 4) and (access & ~GPU_RW) == 0
        assert c.evaluate(granted, type_, access, major, minor) == (mine or before) == o.interpret(granted, type_, access, major, minor)
    assert c.plan(granted, [c.gpu_rule(3)], True) == granted    # idempotent
    one = c.plan(granted, [c.gpu_rule(3)], False)
    assert one[1] == (0xB4, 0, 0, 0, 1) and not c.evaluate(one, c.DEV_CHAR, GPU_RW, 195, 3) and c.evaluate(one, c.DEV_CHAR, GPU_RW, 195, 4)
    assert c.plan(one, [c.gpu_rule(4)], False) == original      # last rule gone: the original, byte for byte
    assert c.plan(original, [c.gpu_rule(9)], False) == original # revoking what nobody granted changes nothing


def test_revoking_a_device_the_foreign_program_grants_puts_a_deny_in_the_prefix():
    original = o.first_match_program(o.RUNC_DEFAULT_RULES + [(o.DEV_CHAR, 195, -1, o.ACC_ALL, True)])   # all GPUs allowed
    assert o.interpret(original, c.DEV_CHAR, GPU_RW, 195, 2)
    denied = c.plan(original, [c.gpu_rule(2)], False)
    assert denied[1] == (0xB4, 0, 0, 0, 1) and denied[-len(original):] == original
    for access in range(1, 8):
        want = o.interpret(original, c.DEV_CHAR, access, 195, 2) and (access & ~GPU_RW) != 0    # rw requests denied, 'm' untouched
        assert c.evaluate(denied, c.DEV_CHAR, access, 195, 2) == want
        assert c.evaluate(denied, c.DEV_CHAR, access, 195, 3) == o.interpret(original, c.DEV_CHAR, access, 195, 3)
    assert c.plan(denied, [c.gpu_rule(2)], True)[-len(original) - 8:] == c.plan(original, [c.gpu_rule(2)], True)[-len(original) - 8:]


def test_programs_that_cannot_be_reloaded_are_refused():
    with_call = [(0x85, 0, 0, 0, 6), (0xB7, 0, 0, 0, 1), (0x95, 0, 0, 0, 0)]          # call bpf_trace_printk
    with_map = [(0x18, 1, 1, 0, 5), (0, 0, 0, 0, 0), (0xB7, 0, 0, 0, 1), (0x95, 0, 0, 0, 0)]   # ld_imm64 r1, map fd
    for prog in (with_call, with_map):
        with pytest.raises(ProbeError) as e:
            c.plan(prog, [c.gpu_rule(0)], True)
        assert e.value.code == _lib.GMP_ERR_UNSUPPORTED and "re-loaded" in e.value.detail


@settings(max_examples=150, deadline=None)
@given(st.lists(rule_st, max_size=8), st.booleans(), st.booleans(),
       st.lists(st.tuples(st.sampled_from([0, 3, 7, 255]), st.booleans()), min_size=1, max_size=6))
def test_random_grant_revoke_sequences_on_both_layouts(rules, default_allow, foreign, steps):
    """Any sequence of grants and revokes of GPU rules, on a runc-layout or a foreign program: after every step the
    planned program allows rw on exactly the granted GPUs (plus whatever the runtime's own rules allow for requests the
    GPU rules do not speak about), and once everything is revoked a device the runtime does not grant is closed."""
    prog = o.first_match_program(rules, default_allow) if foreign else o.runc_program(rules, default_allow)
    base = (lambda *a: o.decide_first_match(rules, default_allow, *a)) if foreign else (lambda *a: o.decide(rules, default_allow, *a))
    state = {}  # minor -> True (granted) / False (revoked)
    for minor, grant in steps:
        prog = c.plan(prog, [c.gpu_rule(minor)], grant)
        state[minor] = grant
        for m, g in state.items():
            for q in (c.ACC_READ, c.ACC_WRITE, GPU_RW):
                assert c.evaluate(prog, c.DEV_CHAR, q, 195, m) == g == o.interpret(prog, c.DEV_CHAR, q, 195, m)
        for type_, access, major, minor2 in requests(extra=[(195, 255), (195, 7)]):
            touched = type_ == c.DEV_CHAR and major == 195 and minor2 in state and (access & ~GPU_RW) == 0
            if not touched and not (type_ == c.DEV_CHAR and major == 195 and minor2 in state):
                assert c.evaluate(prog, type_, access, major, minor2) == base(type_, access, major, minor2)


# ---- the two interpreters against each other on arbitrary (forward-jumping) programs ---------------------------------
ALU_OPS = [0x00, 0x10, 0x20, 0x30, 0x40, 0x50, 0x60, 0x70, 0x90, 0xA0, 0xB0, 0xC0]      # add sub mul div or and lsh rsh mod xor mov arsh
JMP_OPS = [0x10, 0x20, 0x30, 0x40, 0x50, 0x60, 0x70, 0xA0, 0xB0, 0xC0, 0xD0]            # jeq jgt jge jset jne jsgt jsge jlt jle jslt jsle
imm_st = st.sampled_from([0, 1, 2, 7, 16, 31, 32, 63, 195, 0xFFFF, 0x7FFFFFFF, -1, -2, -0x80000000])


@st.composite
def random_program(draw):
    n = draw(st.integers(1, 24))
    prog = [(0x61, 2, 1, 0, 0), (0x61, 3, 1, 4, 0), (0x61, 4, 1, 8, 0), (0xB7, 0, 0, 0, draw(imm_st)), (0xB7, 5, 0, 0, draw(imm_st))]
    for i in range(n):
        kind = draw(st.integers(0, 9))
        dst, src = draw(st.integers(0, 5)), draw(st.integers(0, 5))
        if dst == 1 or src == 1:
            dst, src = 2, 3                                   # keep the context pointer out of the arithmetic
        if kind <= 5:
            cls = draw(st.sampled_from([0x04, 0x07]))
            op = draw(st.sampled_from(ALU_OPS))
            use_reg = draw(st.booleans())
            prog.append((cls | op | (0x08 if use_reg else 0), dst, src if use_reg else 0, 0, 0 if use_reg else draw(imm_st)))
        elif kind == 6:
            prog.append((draw(st.sampled_from([0x84, 0x87])), dst, 0, 0, 0))                      # neg
        elif kind == 7:
            size = draw(st.sampled_from([(0x61, 4), (0x69, 2), (0x71, 1)]))
            prog.append((size[0], dst, 1, draw(st.integers(0, 12 - size[1])), 0))                 # narrow context loads
        else:
            cls = draw(st.sampled_from([0x05, 0x06]))
            op = draw(st.sampled_from(JMP_OPS))
            use_reg = draw(st.booleans())
            off = draw(st.integers(0, n - 1 - i))                                                  # forward only: terminates
            prog.append((cls | op | (0x08 if use_reg else 0), dst, src if use_reg else 0, off, 0 if use_reg else draw(imm_st)))
    prog.append((0x95, 0, 0, 0, 0))
    return prog


@settings(max_examples=400, deadline=None)
@given(random_program(), st.sampled_from([1, 2]), st.integers(1, 7), st.sampled_from([0, 1, 195, 0xFFFFFFFF]), st.sampled_from([0, 3, 255, 0x80000000]))
def test_interpreters_agree_on_random_programs(prog, type_, access, major, minor):
    """gmp_devfilter_eval decides whether a revoke needs a deny rule in front of a foreign program, so it has to compute
    what the kernel would: checked here against the independent Python interpreter on random ALU/JMP/LDX programs."""
    want = o.interpret(prog, type_, access, major, minor)
    assert c.evaluate(prog, type_, access, major, minor) == want
