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
