"""bench.py's N>1 plumbing on CPU: world_size 2 over gloo (barrier, max-over-ranks, whole-job aggregation), and the
reference arm under a multi-rank launch (rank 0 alone works and prints)."""

import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def free_port():
    import socket

    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        return sk.getsockname()[1]


def torchrun(nproc, script, *args, port=None, timeout=240):
    port = port or free_port()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), script, *args]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT)


def test_max_over_ranks_and_weak_scaling_value_with_gloo():
    p = torchrun(2, os.path.join(ROOT, "tests", "_dist_worker.py"))
    assert p.returncode == 0, p.stderr[-2000:]
    import re

    lines = [json.loads(m) for m in re.findall(r"\{[^{}]*\}", p.stdout)]  # ranks share one pipe: do not rely on newlines
    assert sorted(l["rank"] for l in lines) == [0, 1]
    for l in lines:
        assert l["slowest_ms"] == 15.0                       # the slower rank decides, on every rank
        want = (4 << 30) * 7 * 2 / 15e-3 / 1e9               # both ranks' bytes over the slowest rank's time
        assert abs(l["value"] - want) < 1e-6 * want


def test_single_rank_helpers():
    sys.path.insert(0, ROOT)
    import bench

    assert bench.max_over_ranks(3.5, 1) == 3.5
    assert bench.whole_job_gbs(10 ** 9, 2, 4, 2.0) == 4.0


def test_reference_arm_under_torchrun_prints_once_from_rank_zero():
    p = torchrun(2, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0",
                 "--cpu-sample-gib", "1")
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [json.loads(l) for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    line = lines[0]
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and line["unit"] == "GB/s" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"] == {"value": line["value"], "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
