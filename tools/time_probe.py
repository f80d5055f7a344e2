#!/usr/bin/env python
"""Wall-clock a cold gpumounter-probe run (fresh process: cuInit, context, alloc, kernels, teardown)."""
import json
import subprocess
import sys
import time

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from gpumounter_b200 import probe

args = sys.argv[1:]
t0 = time.perf_counter()
p = subprocess.run([probe.probe_executable(), *args], capture_output=True, text=True)
wall = time.perf_counter() - t0
j = json.loads(p.stdout.strip().splitlines()[-1])
out = {"wall_s": round(wall, 4), "exit": p.returncode, "status": j["status"], "n_gpus": j["n_gpus"],
       "ms": {k: j[k] for k in j if k.startswith("ms_")}}
if j["hbm"]:
    h = j["hbm"][0]
    out["hbm0"] = {k: h[k] for k in h if k.startswith("ms_") or k.endswith("_gbs")}
if "p2p" in j:
    q = j["p2p"]
    out["p2p"] = {k: q[k] for k in q if k.startswith("ms_") or k in ("matrix_match", "alltoall_gbs", "bidir_gbs", "write_gbs", "read_gbs", "memcpy_peer_gbs", "can_access", "nvml_nvlink", "nvlink_active")}
print(json.dumps(out))
