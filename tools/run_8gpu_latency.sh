#!/bin/bash
# One call on an N-GPU box: unmount timing with the NVML busy check in both arms, the two-stage hook with a resident
# worker at 1/2/4/N GPUs, real-GPU churn 1 -> N -> 1, and the host-view / pair tests.
N=${1:-8}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_probe_gpu.py -m gpu -x -q -k "host_view or pairs or wrong_device_set" 2>&1 | tail -3
timeout 600 python tools/mount_latency.py --iters 30 --gpus 1,2,4,8 --churn-rounds 0 --resident-iters 10 --churn-real-rounds 1 \
    --out gpurun_out/mount_latency_resident_${N}gpu.json > gpurun_out/ml_${N}gpu.log 2>&1
tail -c 400 gpurun_out/ml_${N}gpu.log
python - <<PY
import json
d = json.load(open("gpurun_out/mount_latency_resident_${N}gpu.json"))
for k in ("mount", "unmount"):
    for g, v in d[k].items():
        print(k, g, {a: round(b["p50_ms"], 2) for a, b in v.items() if isinstance(b, dict)}, round(v["speedup_p50"], 2))
for mode, rows in d.get("mount_to_usable_resident_worker", {}).items():
    for g, v in rows.items():
        print(mode, g, "p50_s", round(v["p50_s"], 4), "p99_s", round(v["p99_s"], 4), {a: round(b, 2) for a, b in v["median_phases_ms"].items()})
print(json.dumps(d.get("churn_real")))
PY
