"""There is no Go toolchain here, so the cgo stub in INTEGRATION.md cannot be built as Go.  Its C half can: the preamble
(the comment block above `import "C"`) is extracted from the document, compiled with gcc against the shipped header and
driven from a small main against the built library - the same dlopen / dlsym / call sequence cgo would generate."""

import os
import re
import subprocess

from gpumounter_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def preamble():
    text = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    m = re.search(r"package probe\s*/\*\n(.*?)\*/\s*import \"C\"", text, re.S)
    assert m, "cgo preamble not found in INTEGRATION.md"
    lines = [l for l in m.group(1).splitlines() if not l.startswith("#cgo")]
    return "\n".join(lines)


def extension_snippet():
    text = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    m = re.search(r"```c\n(static gmp_mount_env g_env;.*?)```", text, re.S)
    assert m, "extension snippet not found in INTEGRATION.md"
    # it redefines call_mount with the environment: keep everything but that one line, checked separately below
    return m.group(1)


MAIN = r'''
#include <stdio.h>
#include <string.h>
int main(int argc, char** argv) {
    if (argc < 2 || gmp_load("/nonexistent/libgpumounter_probe.so") != -1) return 10;   /* library absent: the reference path stays */
    int rc = gmp_load(argv[1]);
    if (rc != 0) { fprintf(stderr, "gmp_load: %d\n", rc); return 11; }
    gmp_gpu g;
    if (call_device_new(3, "GPU-f61ffc1a-9e61-1c0e-2211-4f8f252fe7bc", &g) != 0) return 12;
    if (g.minor_number != 3 || strcmp(g.device_file_path, "/dev/nvidia3") != 0 || strcmp(g.state, "GPU_FREE_STATE") != 0) return 13;
    gmp_pod pod; memset(&pod, 0, sizeof(pod));
    pod.name = "p"; pod.ns = "default"; pod.uid = "0a1b2c3d-1111-2222-3333-444455556666"; pod.container_id = "docker://abc";
    int done = -1;
    rc = call_mount(&pod, &g, 1, &done);          /* $CGROUP_DRIVER unset: the reference's error, through the stub */
    printf("%d|%s|%s\n", rc, call_strerror(rc), call_last_error());
    char json[256]; int st = -1;
    rc = call_probe(&pod, &g, 1, "/nonexistent/gpumounter-probe", json, sizeof(json), &st);
    printf("%d\n", rc);
    rc = call_unmount(&pod, &g, 1, 0, &done);
    printf("%d\n", rc);
    return 0;
}
'''


def build_and_run(tmp_path, source, name):
    src = tmp_path / f"{name}.c"
    src.write_text(source)
    exe = tmp_path / name
    p = subprocess.run(["gcc", "-std=gnu11", "-Wall", "-Wextra", "-Werror", "-Wno-unused-function", f"-I{ROOT}/include", str(src),
                        "-o", str(exe), "-ldl"], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr[-3000:]
    env = {k: v for k, v in os.environ.items() if k != "CGROUP_DRIVER"}
    return subprocess.run([str(exe), _lib.lib_path()], capture_output=True, text=True, env=env, timeout=60)


def test_cgo_preamble_compiles_and_drives_the_library(tmp_path):
    out = build_and_run(tmp_path, preamble() + MAIN, "stub")
    assert out.returncode == 0, (out.returncode, out.stderr)
    first, second, third = out.stdout.strip().splitlines()
    code, verdict, detail = first.split("|")
    assert int(code) == _lib.GMP_ERR_CGROUP_DRIVER and detail == "unsupported cgroup driver: "   # ref: cgroup.go:81
    assert int(second) == _lib.GMP_ERR_CGROUP_DRIVER and int(third) == _lib.GMP_ERR_CGROUP_DRIVER


def test_extension_snippet_compiles_against_the_header(tmp_path):
    snippet = extension_snippet()
    base = preamble().replace(
        "static int call_mount(const gmp_pod* pod, const gmp_gpu* g, int n, int* done) { return p_mount_gpus(NULL, pod, g, n, done); }\n", "")
    assert base != preamble(), "call_mount line of the preamble changed: update this test"
    main = ("\nint main(void) { set_extensions(GMP_EXT_CRI_RUNTIMES | GMP_EXT_CGROUP_V2, 0); "
            "return (g_env.extensions == 3 && call_mount != 0) ? 0 : 1; }\n")
    src = tmp_path / "ext.c"
    src.write_text(base + "\n" + snippet + main)
    p = subprocess.run(["gcc", "-std=gnu11", "-Wall", "-Wextra", "-Werror", "-Wno-unused-function", "-Wno-address", f"-I{ROOT}/include",
                        str(src), "-o", str(tmp_path / "ext"), "-ldl"], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr[-3000:]
    assert subprocess.run([str(tmp_path / "ext")]).returncode == 0
