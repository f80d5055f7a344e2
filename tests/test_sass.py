"""What the compiler actually emitted for sm_100a, read back with cuobjdump (CPU only; cuobjdump ships with the toolkit).
Guards the properties the bandwidth numbers rest on: 256-bit global loads/stores in the default write/verify variants,
the TMA bulk-copy engine in the default copy kernel, no local-memory spills, sm_100a as the only architecture."""

import re
import shutil
import subprocess

import pytest

from gpumounter_b200 import _lib

pytestmark = pytest.mark.skipif(shutil.which("cuobjdump") is None, reason="cuobjdump not installed")


@pytest.fixture(scope="module")
def sass():
    out = subprocess.run(["cuobjdump", "-sass", _lib.lib_path()], capture_output=True, text=True, timeout=300).stdout
    functions = {}
    name = None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1)
            functions[name] = []
        elif name is not None:
            functions[name].append(line)
    return out, {k: "\n".join(v) for k, v in functions.items()}


def test_only_sm_100a_code_is_shipped(sass):
    out, _ = sass
    archs = set(re.findall(r"arch = (sm_\w+)", out))
    assert archs == {"sm_100a"}, archs


def test_vector_kernels_use_256_bit_global_accesses(sass):
    _, fn = sass
    writes = [k for k in fn if "k_hbm_write_pattern" in k]
    reads = [k for k in fn if "k_hbm_read_verify" in k and "segments" not in k]
    assert writes and reads
    wide_writes = [k for k in writes if "STG.E.ENL2.256" in fn[k]]
    wide_reads = [k for k in reads if "LDG.E.ENL2.256" in fn[k]]
    narrow_writes = [k for k in writes if "STG.E.128" in fn[k] and k not in wide_writes]
    # every pattern has both a 32-byte and a 16-byte flavour (the launch picks by alignment and by tuning knob)
    assert len(wide_writes) >= 5 and len(wide_reads) >= 5 and len(narrow_writes) >= 5
    for k in writes:
        assert not re.search(r"\b(LDL|STL)\b", fn[k]), f"{k} uses local memory"
    # The verify kernels keep their two mismatch counters in local memory because the out-of-line slow path takes them
    # by address; that is fine as long as the streaming loop never touches them: between the first 256-bit load and
    # the first call of the slow path there must be no local-memory instruction.
    for k in wide_reads:
        lines = [l for l in fn[k].splitlines() if re.search(r"/\*[0-9a-f]{4}\*/", l)]
        first_load = next(i for i, l in enumerate(lines) if "LDG.E.ENL2.256" in l)
        first_call = next(i for i, l in enumerate(lines) if i > first_load and "CALL" in l)
        hot = "\n".join(lines[first_load:first_call])
        if "256ELi2ELi2ELi0E" in k:   # <PAT, 256 threads, unroll 2, 32-byte lanes, default hint>: the default variant
            assert not re.search(r"\b(LDL|STL)\b", hot), f"{k}: local memory traffic inside the streaming loop"
            assert hot.count("LDG.E.ENL2.256") >= 2, k    # both loads of the unrolled tile issued before the first use


def test_copy_kernel_drives_the_tma_engine(sass):
    _, fn = sass
    bulk = [k for k in fn if "k_hbm_copy_bulk" in k]
    assert bulk
    for k in bulk:
        body = fn[k]
        assert "UBLKCP.S.G" in body and "UBLKCP.G.S" in body, k       # global -> shared and shared -> global bulk copies
        assert "SYNCS.ARRIVE.TRANS64" in body and "SYNCS.PHASECHK.TRANS64" in body, k   # mbarrier transaction counting
        assert "LDG" not in body and "STG" not in body, f"{k} moves data through registers"


def test_verify_slow_path_is_out_of_line(sass):
    """The mismatch bookkeeping must not sit in the streaming loop: it is a subroutine (CALL ... RET) of the kernel."""
    _, fn = sass
    for k, body in fn.items():
        if "k_hbm_read_verify" in k and "segments" not in k:
            assert "CALL.REL.NOINC" in body and "RET.REL.NODEC" in body, f"{k}: the __noinline__ slow path was inlined"


def test_peer_kernels(sass):
    """The all-to-all push writes with the same 256-bit streaming stores as the write kernel and indexes its per-peer
    arguments from the constant bank (no local-memory copy of the argument struct); the bulk-engine pattern write
    stages through shared memory and never stores to global from a register."""
    _, fn = sass
    a2a = [k for k in fn if "k_peer_alltoall_write" in k]
    assert len(a2a) == 1
    body = fn[a2a[0]]
    assert "STG.E.EF.ENL2.256" in body and not re.search(r"\b(LDL|STL)\b", body) and "LDC" in body
    bulk = [k for k in fn if "k_hbm_write_pattern_bulk" in k]
    assert bulk
    for k in bulk:
        assert "STS.128" in fn[k] and "UBLKCP.G.S" in fn[k] and "STG" not in fn[k], k


def test_allocation_free_verify_writes_its_slot_with_plain_stores(sass):
    """One CTA per segment, the segment's verdict stored into its own slot: no atomics in the streaming kernel (the one
    fold CTA afterwards is where the four atomics are)."""
    _, fn = sass
    seg = [k for k in fn if "k_hbm_read_verify_segments" in k]
    assert len(seg) == 5
    for k in seg:
        assert "LDG.E.128" in fn[k] and not re.search(r"\b(ATOMG|RED)\b", fn[k]), k
    fold = [k for k in fn if "k_noalloc_fold" in k]
    assert len(fold) == 1 and re.search(r"\b(ATOMG|RED)", fn[fold[0]])
