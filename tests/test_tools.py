"""The measurement tools keep working on a CPU box (they are what the profiles/ files come from)."""

import json
import os
import subprocess
import sys

import pytest

from tests import harness
from tests.conftest import FAKE_NVML

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_tool(*args, env=None, timeout=300):
    e = dict(os.environ)
    e.update(env or {})
    p = subprocess.run([sys.executable, *args], cwd=ROOT, capture_output=True, text=True, timeout=timeout, env=e)
    assert p.returncode == 0, p.stderr[-2000:]
    return json.loads(p.stdout.strip().splitlines()[-1])   # every tool prints one JSON object as its last line


def test_logic_latency_tool():
    out = run_tool("tools/logic_latency.py")
    assert set(out["rows"]) >= {"GetPodQOS", "GetCgroupName(systemd)", "NvidiaGPU.String", "devfilter plan (grant 1 GPU)"}
    assert all(0 <= row["library_us"] < 1000 for row in out["rows"].values())


def test_nvml_fd_leak_tool_against_the_fake(fake_nvml):
    fake_nvml()
    text = subprocess.run([sys.executable, "tools/nvml_fd_leak.py"], cwd=ROOT, capture_output=True, text=True, timeout=120,
                          env=dict(os.environ, GPUMOUNTER_NVML_LIB=FAKE_NVML))
    assert text.returncode == 0, text.stderr[-2000:]
    out = json.loads(text.stdout)
    assert all(v == [] for v in out.values()), out      # the fake leaks nothing in any of the four scenarios


@pytest.mark.skipif(harness.can_run() is not None, reason=str(harness.can_run()))
def test_mount_latency_tool_on_a_cpu_box():
    out = run_tool("tools/mount_latency.py", "--iters", "3", "--gpus", "1,8", "--churn-rounds", "1")
    assert out["mount"]["8"]["speedup_p50"] > 2 and out["churn_1_8_1"]["leaked_fds"] == 0 and out["churn_1_8_1"]["zombies"] == 0
    if "churn_1_8_1_cgroup2" in out:
        assert out["churn_1_8_1_cgroup2"]["leaked_fds"] == 0 and out["churn_1_8_1_cgroup2"]["leaked_nodes"] == 0


def test_gpu_side_tools_parse_and_import_without_a_gpu():
    """The scripts a GPU call runs (sweeps, ncu cases, the start-up timeline) must at least compile here; the ones that can say
    --help without touching CUDA do so.  A syntax error found on the GPU box costs GPU minutes."""
    import py_compile

    for tool in ("sweep_p2p.py", "sweep_hbm.py", "ncu_case.py", "ncu_p2p_case.py", "ctx_timeline.py", "mount_latency.py",
                 "standin_pod.py", "alloc_costs.py"):
        py_compile.compile(os.path.join(ROOT, "tools", tool), doraise=True)
    for tool in ("ctx_timeline.py", "mount_latency.py"):
        p = subprocess.run([sys.executable, os.path.join("tools", tool), "--help"], cwd=ROOT, capture_output=True, text=True, timeout=120)
        assert p.returncode == 0 and "usage" in p.stdout.lower(), p.stderr[-1500:]
    for script in ("run_1gpu_check.sh", "run_2gpu.sh", "run_8gpu.sh", "run_ncu.sh", "run_2gpu_ncu.sh", "pod_diag.sh", "sass_excerpt.sh"):
        assert subprocess.run(["bash", "-n", os.path.join(ROOT, "tools", script)]).returncode == 0, script


def test_sass_excerpt_tool_lists_the_shipped_kernels():
    import shutil

    if shutil.which("cuobjdump") is None:
        pytest.skip("cuobjdump not installed")
    out = subprocess.run(["bash", "tools/sass_excerpt.sh", "gpumounter_b200/gpumounter-probe"], cwd=ROOT, capture_output=True,
                         text=True, timeout=300).stdout
    assert "sm_100a" in out.splitlines()[1]
    for kernel, mnemonic in (("k_hbm_copy_bulk<2, 114688>", "UBLKCP.G.S"), ("k_peer_alltoall_write", "STG.E.EF.ENL2.256"),
                             ("k_hbm_read_verify_segments<0>", "LDG.E.128")):
        block = out.split(kernel, 1)[1].split("\n\n", 1)[0]
        assert mnemonic in block, (kernel, block[:400])
