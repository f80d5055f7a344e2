"""The C-ABI library loads without a GPU and exports every symbol include/*.h declares."""

import ctypes
import glob
import os
import re

import pytest

import gpumounter_b200 as g
from gpumounter_b200 import _lib, probe

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = []
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        text = open(h).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names += re.findall(r"GMP_API[^;(]*?\b(gmp_[a-z0-9_]+)\s*\(", text)
    return sorted(set(names))


def test_header_declares_something():
    syms = declared_symbols()
    assert "gmp_probe_hbm" in syms and "gmp_hbm_read_verify" in syms and len(syms) >= 15


def test_every_declared_symbol_is_exported():
    handle = ctypes.CDLL(g.lib_path())
    missing = [s for s in declared_symbols() if not hasattr(handle, s)]
    assert not missing, f"declared in include/*.h but not exported: {missing}"


def test_no_undeclared_exports():
    import subprocess

    out = subprocess.run(["nm", "-D", "--defined-only", g.lib_path()], capture_output=True, text=True, check=True).stdout
    exported = sorted(l.split()[-1] for l in out.splitlines() if " T " in l and l.split()[-1].startswith("gmp_"))
    assert exported == declared_symbols()


def test_abi_version_and_strerror():
    assert probe.abi_version() == 1
    assert g.lib().gmp_strerror(0) == b"ok"
    assert g.lib().gmp_strerror(_lib.GMP_ERR_NVML_NOT_FOUND) == b"could not load NVML library"
    assert g.lib().gmp_strerror(-999) == b"unknown error"


def test_struct_sizes_match_header():
    # sizes computed by hand from the header layout (natural alignment, LP64)
    assert ctypes.sizeof(probe.gmp_device) == 96 + 4 + 32 + 4 + 4
    assert ctypes.sizeof(probe.gmp_verify_result) == 32
    assert ctypes.sizeof(probe.gmp_hbm_opts) == 40


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setenv("GPUMOUNTER_PROBE_LIB", str(tmp_path / "nope.so"))
    monkeypatch.setattr(_lib, "_lib", None)
    with pytest.raises(ImportError, match="no CPU fallback"):
        _lib.lib()


def test_argument_validation_without_gpu():
    with pytest.raises(g.ProbeError) as e:
        probe.hbm_write_pattern(0x1000, 64, pattern=9)
    assert e.value.code == _lib.GMP_ERR_INVALID
    with pytest.raises(g.ProbeError) as e:
        probe.hbm_write_pattern(0x1008, 64)  # misaligned
    assert e.value.code == _lib.GMP_ERR_INVALID
    with pytest.raises(g.ProbeError) as e:
        probe.tune("no_such_knob", 1)
    assert e.value.code == _lib.GMP_ERR_INVALID
    assert probe.tune("read_variant") >= 0


@pytest.mark.skipif(os.path.exists("/dev/nvidiactl"), reason="box has a GPU")
def test_no_cpu_fallback_on_a_gpu_less_box():
    for call in (probe.enumerate_cuda, lambda: probe.probe_hbm(nbytes=1 << 20)):
        with pytest.raises(g.ProbeError) as e:
            call()
        assert e.value.code == _lib.GMP_ERR_UNAVAILABLE


def test_header_is_plain_c(tmp_path):
    """The boundary is a C ABI: the header must compile as C99 (cgo compiles it as C), pedantically, on its own."""
    import subprocess

    src = tmp_path / "use_header.c"
    src.write_text('#include "gpumounter_probe.h"\n'
                   "int main(void) { gmp_report r; gmp_hbm_opts o; (void)r; (void)o; return sizeof(gmp_p2p_result) > 0 ? 0 : 1; }\n")
    for std in ("c99", "c11"):
        p = subprocess.run(["gcc", f"-std={std}", "-pedantic", "-Wall", "-Wextra", "-Werror", f"-I{ROOT}/include", "-c", str(src),
                            "-o", str(tmp_path / "use_header.o")], capture_output=True, text=True)
        assert p.returncode == 0, p.stderr[-2000:]


def test_python_struct_mirrors_match_the_compiled_sizes(tmp_path):
    """ctypes mirrors in probe.py / _model.py against sizeof() as the C compiler sees the header."""
    import subprocess

    from gpumounter_b200 import _model, cgroup2

    names = ["gmp_device", "gmp_verify_result", "gmp_hbm_opts", "gmp_hbm_result", "gmp_p2p_opts", "gmp_p2p_result",
             "gmp_probe_opts", "gmp_report", "gmp_gpu", "gmp_pod", "gmp_container_resources", "gmp_mount_env", "gmp_ns_config",
             "gmp_process_info", "gmp_gpu_health", "gmp_container_devices", "gmp_container_entry", "gmp_pod_resources",
             "gmp_dev_rule", "gmp_bpf_insn"]
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include "gpumounter_probe.h"\nint main(void) {\n' +
                   "".join(f'  printf("{n} %zu\\n", sizeof({n}));\n' for n in names) + "  return 0;\n}\n")
    exe = tmp_path / "sizes"
    subprocess.run(["gcc", f"-I{ROOT}/include", str(src), "-o", str(exe)], check=True)
    sizes = dict(l.split() for l in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.splitlines())
    for n in names:
        mirror = getattr(probe, n, None) or getattr(_model, n, None) or getattr(cgroup2, n)
        assert ctypes.sizeof(mirror) == int(sizes[n]), (n, ctypes.sizeof(mirror), sizes[n])
