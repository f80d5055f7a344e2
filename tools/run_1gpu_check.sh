#!/bin/bash
# One call on a 1-GPU box: the GPU suite (not -x: every failure is wanted), the bench line, the pod diagnostics.
#   gpurun --timeout 1000 -- 'bash tools/run_1gpu_check.sh'
mkdir -p gpurun_out
bash tools/pod_diag.sh > gpurun_out/pod_diag.txt 2>&1
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; tail -15 gpurun_out/pytest_gpu.log
timeout 420 python bench.py --steps 10 --warmup 3 2> gpurun_out/bench.err | tail -1 > gpurun_out/bench.json; echo "bench rc=${PIPESTATUS[0]}"
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench.json"))
    print("value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms/call", round(d["e2e"]["ms_per_call"], 2), d["e2e"]["phases_ms_last_call"])
    print("cold", round(d["e2e"]["cold_call"]["ms_per_call"], 2), d["e2e"]["cold_call"]["phases_ms_last_call"])
    print("kernels", {k: round(v["gbs"]) for k, v in d["kernels"].items()}, "roofline", d["roofline"]["frac"], "clocks", d.get("clocks"))
    m = d.get("mount", {})
    print("mount_ok", d.get("mount_ok"), {k: m.get(k) for k in ("unavailable", "traceback")})
    u = m.get("mount_to_usable", {})
    print("cold in-pod", u.get("target"), u.get("p50_s"), u.get("median_phases_ms"))
    r = m.get("mount_to_usable_resident_worker", {})
    print("resident", r.get("p50_s"), r.get("median_phases_ms"))
except Exception as e:
    print("bench summary failed:", e); print(open("gpurun_out/bench.err").read()[-3000:])
PY
timeout 120 python tools/alloc_costs.py 2>/dev/null > gpurun_out/alloc_costs.json; tail -c 1500 gpurun_out/alloc_costs.json
