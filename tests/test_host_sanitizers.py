"""ASan+UBSan and TSan over the host-only half of the library (the reference builds and tests with neither; it has real
races).  The CUDA half cannot be sanitised here: compute-sanitizer is closed on the GPU pool."""

import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "gpumounter_b200", "csrc")
SRCS = [os.path.join(ROOT, "tests", "native", "host_sanitize.cpp"), os.path.join(CSRC, "gmp_mount.cpp"),
        os.path.join(CSRC, "gmp_nvml.cpp"), os.path.join(CSRC, "gmp_devfilter.cpp")]
CXX = shutil.which("g++") or "g++"


@pytest.mark.parametrize("name,flags", [("asan_ubsan", ["-fsanitize=address,undefined", "-fno-sanitize-recover=all"]),
                                         ("tsan", ["-fsanitize=thread"])])
def test_host_code_is_clean_under(name, flags, tmp_path):
    exe = str(tmp_path / f"host_{name}")
    build = subprocess.run([CXX, "-std=c++17", "-O1", "-g", "-fno-omit-frame-pointer", *flags, f"-I{ROOT}/include",
                            f"-I{CSRC}", "-I/usr/local/cuda/include", *SRCS, "-o", exe, "-ldl", "-lpthread"],
                           capture_output=True, text=True)
    if build.returncode != 0 and ("cannot find" in build.stderr or "unrecognized" in build.stderr):
        pytest.skip(f"toolchain lacks {name}: {build.stderr[-300:]}")
    assert build.returncode == 0, build.stderr[-3000:]
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=1:abort_on_error=0", TSAN_OPTIONS="halt_on_error=1",
               UBSAN_OPTIONS="print_stacktrace=1")
    run = subprocess.run([exe], capture_output=True, text=True, env=env, timeout=300)
    if run.returncode != 0 and "FATAL: ThreadSanitizer: unexpected memory mapping" in run.stderr:
        pytest.skip("TSan cannot map its shadow on this kernel (ASLR entropy)")
    assert run.returncode == 0 and "host sanitize ok" in run.stdout, run.stdout[-1000:] + run.stderr[-4000:]
