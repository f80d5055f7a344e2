#!/usr/bin/env python
"""Experiment: time-to-usable for N GPUs with ONE probe process versus one probe process PER GPU
(each seeing only its own device through CUDA_VISIBLE_DEVICES)."""
import json
import os
import subprocess
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpumounter_b200 import probe

exe = probe.probe_executable()
nv = {d.uuid: d for d in probe.enumerate_nvml()}
uuids = [d.uuid for d in probe.enumerate_cuda()]
common = ["--bytes", "1G", "--iters", "1", "--dma-bytes", "16M", "--skip-p2p"]


def single(n):
    env = dict(os.environ, CUDA_VISIBLE_DEVICES=",".join(uuids[:n]))
    t0 = time.monotonic_ns()
    p = subprocess.run([exe, *sum((["--uuid", f"{u}:{nv[u].minor}"] for u in uuids[:n]), []), *common], env=env,
                       capture_output=True, text=True)
    j = json.loads(p.stdout.strip().splitlines()[-1])
    return (j["mono_usable_ns"] - t0) / 1e9, j["ms_cuda_init"], max(h["ms_context"] for h in j["hbm"]), j["status"]


def fanout(n, extra_env=None):
    t0 = time.monotonic_ns()
    procs = []
    for u in uuids[:n]:
        env = dict(os.environ, CUDA_VISIBLE_DEVICES=u, **(extra_env or {}))
        procs.append(subprocess.Popen([exe, "--uuid", f"{u}:{nv[u].minor}", *common], env=env, stdout=subprocess.PIPE, text=True))
    outs = [json.loads(p.communicate()[0].strip().splitlines()[-1]) for p in procs]
    t_end = time.monotonic_ns()
    usable = max(j["mono_usable_ns"] for j in outs)
    return ((usable - t0) / 1e9, max(j["ms_cuda_init"] for j in outs), max(j["hbm"][0]["ms_context"] for j in outs),
            [j["status"] for j in outs], (t_end - t0) / 1e9)


for n in (1, 2, 4, 8):
    if n > len(uuids):
        break
    for rep in range(3):
        print(json.dumps({"n": n, "rep": rep, "single": single(n), "fanout": fanout(n)}), flush=True)
