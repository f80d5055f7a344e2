#!/bin/bash
# One call on a 1-GPU box after a round of CPU-only changes: the GPU tests, the bench line, and the latency numbers whose
# code changed without a GPU to run it on (first touch ahead of the large cudaMalloc; one NVML attach per probe).
#   gpurun --timeout 900 -- 'bash tools/run_1gpu_check.sh'
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 10 --warmup 3 2> gpurun_out/bench.err | tail -1 > gpurun_out/bench.json
python - <<PY
import json
d = json.load(open("gpurun_out/bench.json"))
print("value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms/call", round(d["e2e"]["ms_per_call"], 1), d["e2e"].get("phases_ms_last_call"))
print("roofline", d["roofline"], "clocks", d.get("clocks"))
PY
timeout 300 python tools/mount_latency.py --iters 30 --gpus 1,2,4,8 --churn-rounds 0 --usable-iters 10 --resident-iters 20 \
    --churn-real-rounds 20 --out gpurun_out/mount_latency_check.json > gpurun_out/ml_check.log 2>&1
python - <<PY
import json
d = json.load(open("gpurun_out/mount_latency_check.json"))
u = d["mount_to_usable"]["1"]
print("cold in-pod", u["p50_s"], u["median_phases_ms"])
for mode, rows in d["mount_to_usable_resident_worker"].items():
    for g, v in rows.items():
        print(mode, g, v["p50_s"], v["median_phases_ms"])
print(d["churn_real"])
PY
gpumounter_b200/gpumounter-probe --bytes 1G --iters 1 | python -c "
import json,sys; j=json.load(sys.stdin); print({k: j[k] for k in ('status','ms_nodes','ms_cuda_init','ms_nvml','ms_hbm','ms_p2p','ms_to_usable','ms_total')})"
# the same mount-to-usable, but inside a REAL cgroup-v1 devices group that denies GPU nodes until they are mounted
# (never run on a GPU box yet: if cuInit needs a device the whitelist lacks, the probe says which open() was refused)
timeout 200 python tools/mount_latency.py --iters 3 --gpus 1 --churn-rounds 0 --usable-iters 5 --real-cgroup \
    --out gpurun_out/mount_latency_real_cgroup.json > gpurun_out/ml_real_cgroup.log 2>&1; tail -c 600 gpurun_out/ml_real_cgroup.log
timeout 120 python tools/alloc_costs.py 2>/dev/null > gpurun_out/alloc_costs.json; tail -c 1200 gpurun_out/alloc_costs.json
