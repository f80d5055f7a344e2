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
